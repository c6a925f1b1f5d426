"""rt_trace_static: the entry point that allocates nothing (static scratch inside librt_b200.so, include/rt_b200.h).

Collected last on purpose.  Through the C ABI, against rt_trace -- the existing entry
point that runs the same kernel with buffers it allocates -- bit for bit (hit ids and t are integer / exact work in this
repository's parity bar), on calls in a row that share the scratch array."""
import numpy as np
import pytest

from helpers import random_rays
from pyrt import scenes as S

pytestmark = pytest.mark.gpu


def _camera_like_rays(n, seed):
    """Rays that start at the world origin (inside any scene's extent): no preparation beyond rt_upload_scene is needed."""
    r = random_rays(n, seed)
    r["start"][:, :3] = 0.0
    return r


def test_static_trace_equals_allocating_trace_on_calls_in_a_row(rt, ctx):
    objs, lights = S.synthetic_scene(4000, 600)
    ctx.upload_scene(objs, lights)
    first, second = random_rays(20000, 5), random_rays(7000, 11)
    want = {id(r): ctx.trace(r, 2) for r in (first, second)}  # also prepares the scene for origins this far out
    for accel in (2, -1, 1, 0):
        for rays in (first, second, first):  # the shorter batch runs over the leftovers of the longer one and back
            gid, gt = ctx.trace_static(rays, accel)
            hid, ht = want[id(rays)]
            assert np.array_equal(gid, hid) and np.array_equal(gt, ht)
    assert (want[id(first)][0] >= 0).sum() > 1000  # the comparison is not one of misses only


def test_static_trace_right_after_upload_and_across_contexts(rt, ctx):
    """No rt_trace before it: rays from inside the scene need only the resident scene.  A second context on the same GPU
    shares the array (calls in sequence)."""
    objs, lights = S.synthetic_scene(3000, 300)
    rays = _camera_like_rays(30000, 3)
    ctx.upload_scene(objs, lights)
    a_id, a_t = ctx.trace_static(rays)
    other = rt.Context(0)
    try:
        o2, l2 = S.synthetic_scene(500, 100)
        other.upload_scene(o2, l2)
        b_id, b_t = other.trace_static(rays[:9000])
        a2_id, a2_t = ctx.trace_static(rays)
        w_id, w_t = other.trace(rays[:9000])
        assert np.array_equal(b_id, w_id) and np.array_equal(b_t, w_t)
    finally:
        other.close()
    h_id, h_t = ctx.trace(rays)
    assert np.array_equal(a_id, h_id) and np.array_equal(a_t, h_t)
    assert np.array_equal(a2_id, h_id) and np.array_equal(a2_t, h_t)


def test_static_trace_limits(rt, ctx):
    objs, lights = S.synthetic_scene(200, 50)
    ctx.upload_scene(objs, lights)
    ctx.trace_static(_camera_like_rays(0, 1))  # the per-device set-up alone
    full = _camera_like_rays(rt.TRACE_STATIC_MAX_RAYS, 2)
    gid, gt = ctx.trace_static(full)
    hid, ht = ctx.trace(full)
    assert np.array_equal(gid, hid) and np.array_equal(gt, ht)
    with pytest.raises(Exception, match="exceed the static scratch"):
        ctx.trace_static(_camera_like_rays(rt.TRACE_STATIC_MAX_RAYS + 1, 2))
    far = _camera_like_rays(16, 4)
    far["start"][:, 0] = 1.0e6  # would need the scene prepared again (an allocation): refused, not done
    with pytest.raises(Exception, match="not prepared for these rays"):
        ctx.trace_static(far)
    fresh = rt.Context(0)
    try:
        with pytest.raises(Exception, match="no scene"):
            fresh.trace_static(_camera_like_rays(4, 1))
    finally:
        fresh.close()
