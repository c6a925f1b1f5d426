"""CPU tier: the CUDA path's device code (rt_device.cuh / rt_pixel.cuh / rt_scene.cpp) compiled for the host
(tests/emul) against the oracle.  This checks, without a GPU, that the conservative filters and the BVH never
lose a hit the reference arithmetic would find, and that the per-pixel state machine -- including the
dead-ray skipping -- is output-identical to the reference kernel."""
import numpy as np
import pytest

import emul
from helpers import frame_rays, random_rays
from pyrt import scenes as S
from pyrt import types as T


def _scene(name):
    return S.kat2_scene() if name == "kat2" else S.load_reference_scene(name)


@pytest.mark.parametrize("name", list(S.REFERENCE_SCENES) + ["simpleScene3L", "kat2"])
@pytest.mark.parametrize("accel", [0, 1, 2])
def test_device_code_bit_exact_on_reference_scenes(oracle_mod, name, accel):
    objs, lights = _scene(name)
    _, rays = frame_rays(128, 72)
    for mode in (0, 1, 2):
        ref, hid, ht, st = oracle_mod.port_render(mode, 30, objs, lights, rays)
        for skip_dead in (0, 1):
            out, eid, et, c = emul.render(objs, lights, rays, 30, mode, accel, skip_dead)
            assert np.array_equal(eid, hid) and np.array_equal(et, ht)
            assert np.array_equal(out, ref)
            assert (c["ref_primary"], c["ref_shadow"], c["ref_reflect"]) == (st.rays_primary, st.rays_shadow, st.rays_reflect)
            assert c["traced_shadow"] <= c["ref_shadow"] and c["traced_reflect"] <= c["ref_reflect"]
            if not skip_dead and mode == 2:
                assert c["traced_shadow"] == c["ref_shadow"] and c["traced_reflect"] == c["ref_reflect"]


@pytest.mark.parametrize("second_type", [T.BOX, T.SPHERE, T.TRIANGLE])
def test_bvh_and_cull_equal_brute_force_synthetic(oracle_mod, second_type):
    objs, lights = S.synthetic_scene(3000, 300, second_type=second_type)
    _, rays = frame_rays(96, 54)
    ref, hid, ht, st = oracle_mod.port_render(2, 4, objs, lights, rays)
    assert st.rays_reflect > st.hit_pixels  # reflective materials are exercised
    if second_type == T.TRIANGLE:
        assert np.isin(hid[hid >= 0], np.flatnonzero(objs["type"] == T.TRIANGLE)).mean() > 0.2
    for accel in (0, 1, 2):
        out, eid, et, c = emul.render(objs, lights, rays, 4, 2, accel, 1)
        assert np.array_equal(eid, hid) and np.array_equal(et, ht) and np.array_equal(out, ref)
    for leaf in (1, 2, 8, 16):
        out, eid, et, c = emul.render(objs, lights, rays, 4, 2, 2, 1, leaf=leaf)
        assert np.array_equal(eid, hid) and np.array_equal(out, ref)


@pytest.mark.parametrize("second_type", [T.SPHERE, T.BOX, T.TRIANGLE])
def test_filters_never_drop_a_noisy_hit(oracle_mod, second_type):
    """Adversarial for the bounds: tiny, strongly anisotropic objects very far from the ray origins, where the
    reference's discriminant (spheres) / slab arithmetic is dominated by rounding noise (rt_scene.cpp header).
    Rays are aimed at the jittered image of the object origin -- for the triangle extension that is a point ON its
    hypotenuse, so about half of them graze the edge."""
    rng = np.random.Generator(np.random.PCG64(7))
    n = 400
    objs, _ = S.synthetic_scene(n, 0, seed=11, second_type=second_type)
    # shrink every object by 20x about its centre: mv' = mv * S(1/20), mvInverse' = S(20) * mvInverse
    mv = objs["mv"].reshape(n, 4, 4).copy()
    mi = objs["mvInverse"].reshape(n, 4, 4).copy()
    mv[:, :3, :] *= np.float32(1 / 16.0)     # columns 0..2 of mv
    mi[:, :, :3] *= np.float32(16.0)         # rows 0..2 of mvInverse (incl. translation)
    objs["mv"], objs["mvInverse"] = mv.reshape(n, 16), mi.reshape(n, 16)
    # rays aimed at (jittered) object centres from far away, arbitrary direction length
    centre = mv[:, 3, :3]
    k = 6000
    pick = rng.integers(0, n, k)
    o = (rng.normal(size=(k, 3)) * 150.0).astype(np.float32)
    tgt = centre[pick] + (rng.normal(size=(k, 3)) * 0.03).astype(np.float32)
    rays = np.zeros(k, dtype=T.RAY_DTYPE)
    rays["start"][:, :3] = o
    rays["start"][:, 3] = 1.0
    rays["direction"][:, :3] = (tgt - o) * rng.uniform(0.5, 30.0, size=(k, 1)).astype(np.float32)
    hid, ht = oracle_mod.port_trace(objs, rays)
    assert (hid >= 0).sum() > (k // 10 if second_type == T.SPHERE else k // 50), (hid >= 0).sum()
    lights = np.zeros(0, dtype=T.LIGHT_DTYPE)
    for accel in (1, 2):
        _, eid, et, _ = emul.render(objs, lights, rays, 0, 0, accel, 1)
        assert np.array_equal(eid, hid) and np.array_equal(et, ht)


def test_random_rays_nearest_hit(oracle_mod):
    objs, _ = S.synthetic_scene(2000, 400)
    rays = random_rays(5000, 3)
    hid, ht = oracle_mod.port_trace(objs, rays)
    lights = np.zeros(0, dtype=T.LIGHT_DTYPE)
    for accel in (0, 1, 2):
        _, eid, et, _ = emul.render(objs, lights, rays, 0, 0, accel, 1)
        assert np.array_equal(eid, hid) and np.array_equal(et, ht)


def test_tie_rules_coincident_objects(oracle_mod):
    """Exact ties: duplicated spheres (highest index wins) and duplicated boxes (lowest index wins), SAR:108,147."""
    base, lights = S.kat2_scene()
    objs = np.concatenate([base[[0]], base[[1]], base[[0]], base[[1]], base[[2]], base[[2]]])
    _, rays = frame_rays(160, 90)
    ref, hid, ht, _ = oracle_mod.port_render(2, 4, objs, lights, rays)
    assert set(np.unique(hid)) == {-1, 0, 3, 5}
    for accel in (0, 1, 2):
        out, eid, et, _ = emul.render(objs, lights, rays, 4, 2, accel, 1, leaf=1)
        assert np.array_equal(eid, hid) and np.array_equal(out, ref)


def test_degenerate_inputs(oracle_mod):
    objs, lights = S.kat2_scene()
    _, rays = frame_rays(64, 36)
    # no lights: shade() returns black, reflection still runs
    nol = lights[:0]
    ref, hid, *_ = oracle_mod.port_render(2, 4, objs, nol, rays)
    for accel in (0, 1, 2):
        out, eid, *_ = emul.render(objs, nol, rays, 4, 2, accel, 1)
        assert np.array_equal(out, ref) and np.array_equal(eid, hid)
    # no objects: every pixel keeps {0,0,0,1}
    out, eid, *_ = emul.render(objs[:0], lights, rays, 4, 2, 2, 1)
    assert np.all(out == np.array([0, 0, 0, 1], dtype=np.float32)) and np.all(eid == -1)
    # axis-parallel rays (direction component exactly 0 -> the copysign branch of the slab test, SAR:36-43),
    # origins inside a box / inside a sphere, unknown primitive type (skipped by the kernels' switch)
    r = np.zeros(7, dtype=T.RAY_DTYPE)
    r["start"][:, 3] = 1.0
    r["start"][:, :3] = [(0, 0, 0), (0, 0, 0), (1, 1, 0), (0, 0, -10), (3, 2, -7), (5, 0, -10), (0, 5, -10)]
    r["direction"][:, :3] = [(0, 0, -1), (0, 0, 1), (0, 0, -3), (0, 1, 0), (1, 0, 0), (-1, 0, 0), (0, -2, 0)]
    weird = objs.copy()
    weird = np.concatenate([weird, weird[[1]]])
    weird[3]["type"] = 7
    hid, ht = oracle_mod.port_trace(weird, r)
    assert (hid >= 0).sum() >= 5 and not (hid == 3).any()
    for accel in (0, 1, 2):
        _, eid, et, _ = emul.render(weird, lights, r, 0, 0, accel, 1)
        assert np.array_equal(eid, hid) and np.array_equal(et, ht)
    # max_bounces 0 and 1 exercise the uint wrap of `bounces` (SAR:258,268,281)
    _, rays = frame_rays(96, 54)
    for mb in (0, 1, 2):
        ref, *_ = oracle_mod.port_render(2, mb, objs, lights, rays)
        for sd in (0, 1):
            out, *_ = emul.render(objs, lights, rays, mb, 2, 2, sd)
            assert np.array_equal(out, ref)


def test_unsupported_scenes_are_rejected():
    objs, lights = S.kat2_scene()
    _, rays = frame_rays(8, 8)
    bad = objs.copy()
    bad[0]["mvInverse"][3] = 0.25  # projective row
    with pytest.raises(RuntimeError):
        emul.render(bad, lights, rays)
    bad = objs.copy()
    bad[1]["mvInverse"][:] = 0
    bad[1]["mvInverse"][15] = 1
    with pytest.raises(RuntimeError):
        emul.render(bad, lights, rays)


@pytest.mark.parametrize("name", list(S.REFERENCE_SCENES) + ["simpleScene3L", "kat2"])
def test_wavefront_state_machine_bit_exact(oracle_mod, name):
    """rt_wave.cuh (one ray in flight per path, queues, backward light walk) == the reference work-item."""
    objs, lights = _scene(name)
    _, rays = frame_rays(128, 72)
    for mode in (0, 1, 2):
        for mb in ((30, 0, 1, 2) if mode == 2 else (30,)):
            ref, hid, ht, st = oracle_mod.port_render(mode, mb, objs, lights, rays)
            out, c, gens = emul.render_wave(objs, lights, rays, mb, mode, 2)
            assert np.array_equal(out, ref)
            assert (c["ref_primary"], c["ref_shadow"], c["ref_reflect"]) == (st.rays_primary, st.rays_shadow, st.rays_reflect)
            assert c["hit_pixels"] == st.hit_pixels


def test_wavefront_state_machine_synthetic(oracle_mod):
    objs, lights = S.synthetic_scene(2500, 300, n_lights=3)
    _, rays = frame_rays(96, 54)
    for mode in (1, 2):
        ref, *_ = oracle_mod.port_render(mode, 4, objs, lights, rays)
        for accel in (1, 2):
            out, c, gens = emul.render_wave(objs, lights, rays, 4, mode, accel)
            assert np.array_equal(out, ref)
    out, c, gens = emul.render_wave(objs, lights[:0], rays, 4, 2, 2)
    ref, *_ = oracle_mod.port_render(2, 4, objs, lights[:0], rays)
    assert np.array_equal(out, ref)


def test_triangle_extension_geometry(oracle_mod):
    """The type-2 extension against first principles: an axis-aligned unit triangle at z = -5, scale 2, seen from the
    origin.  A pixel hits iff its ray meets the plane inside x >= -1, y >= -1, x + y <= 0 (object units x 2); two
    coincident triangles tie towards the lower index (box rule); types >= 3 stay skipped like in the reference."""
    mv = np.eye(4, dtype=np.float32)
    mv[0, 0] = mv[1, 1] = mv[2, 2] = 2.0
    mv[3, :3] = (0, 0, -5)
    inv = np.linalg.inv(mv.astype(np.float64)).astype(np.float32)
    tri = T.make_object(T.TRIANGLE, mv, inv, ambient=(0.2, 0.3, 0.4), diffuse=(0.5, 0.5, 0.5))
    dead = T.make_object(3, mv, inv)
    objs = np.stack([dead, tri, tri])
    lights = np.stack([T.make_light((0, 0, 0, 1), ambient=(1, 1, 1), diffuse=(1, 1, 1))])
    W, H = 96, 96
    z, rays = frame_rays(W, H)
    ref, hid, ht, _ = oracle_mod.port_render(2, 2, objs, lights, rays)
    d = rays["direction"].astype(np.float64)
    t = -5.0 / d[:, 2]
    x, y = t * d[:, 0], t * d[:, 1]
    inside = (x > -1 + 1e-6) & (y > -1 + 1e-6) & (x + y < -1e-6)
    outside = (x < -1 - 1e-6) | (y < -1 - 1e-6) | (x + y > 1e-6)
    assert inside.sum() > 50
    assert (hid[inside] == 1).all() and (hid[outside] == -1).all()
    assert np.allclose(ht[inside], t[inside], rtol=1e-6)
    for accel in (0, 1, 2):
        out, eid, et, _ = emul.render(objs, lights, rays, 2, 2, accel, 1, leaf=1)
        assert np.array_equal(eid, hid) and np.array_equal(et, ht) and np.array_equal(out, ref)


# ------------------------------------------------------------------ perspective grids (rt_pgrid.cuh), host emulation
@pytest.mark.parametrize("second_type", [T.BOX, T.SPHERE, T.TRIANGLE])
def test_perspective_grids_bit_exact_synthetic(oracle_mod, second_type):
    """accel 3: camera rays and point-light shadow rays look up one cell of a perspective grid around their apex
    (host-built here with the same per-object functions the device builder uses), the rest walks the BVH."""
    objs, lights = S.synthetic_scene(3000, 300, second_type=second_type, n_lights=3)
    W, H = 96, 54
    z, rays = frame_rays(W, H)
    for mode in (1, 2):
        ref, _, _, st = oracle_mod.port_render(mode, 4, objs, lights, rays)
        for res, tile in ((16, (8, 4)), (64, (8, 4)), (200, (3, 5))):
            out, c, _ = emul.render_wave(objs, lights, rays, 4, mode, 3, camera=(W, H, z), grid_res=res, tile=tile)
            assert np.array_equal(out, ref)
            assert (c["ref_primary"], c["ref_shadow"], c["ref_reflect"]) == (st.rays_primary, st.rays_shadow, st.rays_reflect)
            assert c["pair_tests"] > 0
    # arbitrary rays: no camera grid, the shadow rays still use the light grids
    rr = random_rays(3000, 9)
    ref, *_ = oracle_mod.port_render(2, 4, objs, lights, rr)
    out, c, _ = emul.render_wave(objs, lights, rr, 4, 2, 3, camera=None, grid_res=64)
    assert np.array_equal(out, ref)


@pytest.mark.parametrize("name", list(S.REFERENCE_SCENES) + ["simpleScene3L", "kat2"])
def test_perspective_grids_reference_scenes(oracle_mod, name):
    """Large objects close to the apex (a light next to a unit cube, the camera inside nothing but looking at objects that
    fill the frame): rectangles that cover whole faces, objects partly behind the apex plane."""
    objs, lights = _scene(name)
    W, H = 128, 72
    z, rays = frame_rays(W, H)
    for mode in (1, 2):
        ref, *_ = oracle_mod.port_render(mode, 30, objs, lights, rays)
        out, c, _ = emul.render_wave(objs, lights, rays, 30, mode, 3, camera=(W, H, z), grid_res=32)
        assert np.array_equal(out, ref)


def test_perspective_grids_light_inside_objects_and_directional(oracle_mod):
    """A light buried in objects (its grids list them in every cell), a light at infinity (position.w == 0: no grids, its
    shadow rays use the BVH), a light exactly on the camera apex."""
    objs, lights = S.synthetic_scene(1500, 200, n_lights=4)
    lights = lights.copy()
    lights[0]["position"] = (0.3, -0.2, -60.0, 1.0)   # in the middle of the object cloud
    lights[1]["position"] = (0.2, 0.9, 0.4, 0.0)      # directional
    lights[2]["position"] = (0.0, 0.0, 0.0, 1.0)      # at the camera
    W, H = 80, 45
    z, rays = frame_rays(W, H)
    for mode in (1, 2):
        ref, *_ = oracle_mod.port_render(mode, 3, objs, lights, rays)
        out, c, _ = emul.render_wave(objs, lights, rays, 3, mode, 3, camera=(W, H, z), grid_res=48)
        assert np.array_equal(out, ref)


def test_perspective_grids_never_drop_a_noisy_hit(oracle_mod):
    """The adversarial scene of test_filters_never_drop_a_noisy_hit seen through the grids: tiny anisotropic spheres whose
    silhouette the reference's cancelling discriminant blurs, a camera 150 units away and lights at the far side."""
    rng = np.random.Generator(np.random.PCG64(17))
    n = 300
    objs, lights = S.synthetic_scene(n, 0, seed=11, second_type=T.SPHERE, n_lights=2)
    mv = objs["mv"].reshape(n, 4, 4).copy()
    mi = objs["mvInverse"].reshape(n, 4, 4).copy()
    mv[:, :3, :] *= np.float32(1 / 16.0)
    mi[:, :, :3] *= np.float32(16.0)
    objs["mv"], objs["mvInverse"] = mv.reshape(n, 16), mi.reshape(n, 16)
    objs["mat"]["absorption"] = 0.5
    centre = mv[:, 3, :3]
    # a pinhole frame whose pixels are aimed at the (jittered) projections of the objects: dense sampling of the silhouettes
    W, H = 4096, 2304
    z = T.camera_zdir(H)
    px = np.clip(np.round(centre[:, 0] / -centre[:, 2] * -z + W / 2), 2, W - 3).astype(np.int64)
    py = np.clip(np.round(H / 2 - centre[:, 1] / -centre[:, 2] * -z), 2, H - 3).astype(np.int64)
    rows = np.unique(np.concatenate([py - 1, py, py + 1]))
    rays = T.camera_rays(W, H, z, rows=rows).reshape(len(rows), W)
    pick = []
    row_of = {int(r): i for i, r in enumerate(rows)}
    for x, y in zip(px, py):
        for dy in (-1, 0, 1):
            for dx in (-2, -1, 0, 1, 2):
                pick.append((row_of[int(y + dy)], int(x + dx)))
    pick = np.array(sorted(set(pick)))
    sub = rays[pick[:, 0], pick[:, 1]].copy()
    ref, hid, ht, _ = oracle_mod.port_render(2, 2, objs, lights, sub)
    assert (hid >= 0).sum() > 200
    # the grid lookup only needs the camera's (W, H, zdir): the rays are a subset of that frame
    out, c, _ = emul.render_wave(objs, lights, sub, 2, 2, 3, camera=(W, H, z), grid_res=256)
    assert np.array_equal(out, ref)


def test_uniform_grid_walk_bit_exact(oracle_mod):
    """rt_grid3.cuh: the DDA walk over a uniform grid for the rays without a common apex, several cell sizes, rays that start
    outside the grid, axis-parallel rays."""
    objs, lights = S.synthetic_scene(3000, 300, n_lights=3)
    W, H = 96, 54
    z, rays = frame_rays(W, H)
    ref, *_ = oracle_mod.port_render(2, 4, objs, lights, rays)
    for dens in (0.25, 1.0, 8.0):
        out, c, _ = emul.render_wave(objs, lights, rays, 4, 2, 3, camera=(W, H, z), grid_res=64, grid3=dens)
        assert np.array_equal(out, ref) and c["cells"] > 0
    rr = random_rays(3000, 9)
    rr["direction"][::7, 0] = 0.0   # walks that never leave their x slab
    rr["direction"][::11, 1:3] = 0.0
    rr["direction"][::11, 0] = 3.0
    ref, *_ = oracle_mod.port_render(2, 4, objs, lights, rr)
    out, c, _ = emul.render_wave(objs, lights, rr, 4, 2, 3, camera=None, grid_res=64, grid3=2.0)
    assert np.array_equal(out, ref)
