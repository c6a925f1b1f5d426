"""Shared helpers of the test suite."""
import numpy as np

from pyrt import types as T


def frame_rays(width, height, fov=60.0):
    z = T.camera_zdir(height, fov)
    return z, T.camera_rays(width, height, z)


def assert_colours_close(got, ref, tol=1e-4):
    """|delta| <= 1e-4 * max(1, |ref|) on xyz, w == 1; PPM within 1 LSB per channel."""
    got = np.asarray(got, dtype=np.float32).reshape(-1, 4)
    ref = np.asarray(ref, dtype=np.float32).reshape(-1, 4)
    d = np.abs(got[:, :3].astype(np.float64) - ref[:, :3].astype(np.float64))
    lim = tol * np.maximum(1.0, np.abs(ref[:, :3].astype(np.float64)))
    bad = d > lim
    assert not bad.any(), "%d channels out of tolerance, max |d| = %g" % (bad.sum(), d.max())
    assert np.all(got[:, 3] == 1.0)
    q = np.abs(T.quantize_ppm(got).astype(np.int64) - T.quantize_ppm(ref).astype(np.int64))
    assert q.max(initial=0) <= 1, "PPM differs by %d LSB" % q.max()


def random_rays(n, seed, origin_scale=30.0, centre=(0.0, 0.0, -60.0)):
    """Rays with random origins around the synthetic scene and random directions of random length."""
    from pyrt.types import RAY_DTYPE

    rng = np.random.Generator(np.random.PCG64(seed))
    rays = np.zeros(n, dtype=RAY_DTYPE)
    rays["start"][:, :3] = (rng.normal(size=(n, 3)) * origin_scale + np.asarray(centre)).astype(np.float32)
    rays["start"][:, 3] = 1.0
    d = rng.normal(size=(n, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    d *= rng.uniform(0.1, 2000.0, size=(n, 1))
    rays["direction"][:, :3] = d.astype(np.float32)
    return rays
