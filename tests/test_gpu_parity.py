"""GPU tier: the CUDA path through the C ABI (ctypes -> librt_b200.so) against the oracle on the same inputs.

Bar (BASELINE.json north_star / SURVEY.md 8(d)): hit ids bit-exact; t and colours |d| <= 1e-4 * max(1,|ref|)
(FP32; the only non-bit-exact operation on the CUDA side is powf), PPM <= 1 LSB per channel; missed pixels
exactly (0,0,0,1).  Full-size frames are checked through size-independent properties (BVH == filter == brute
force, sharded == unsharded, checksums) because the CPU oracle cannot finish them."""
import json
import os

import numpy as np
import pytest

from helpers import assert_colours_close, frame_rays, random_rays
from pyrt import scenes as S
from pyrt import types as T

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
ALL_SCENES = list(S.REFERENCE_SCENES) + ["simpleScene3L", "kat2"]


def _scene(name):
    return S.kat2_scene() if name == "kat2" else S.load_reference_scene(name)


@pytest.mark.parametrize("name", ALL_SCENES)
def test_reference_scenes_all_modes_all_searches(rt, ctx, oracle_mod, name):
    objs, lights = _scene(name)
    W, H = 320, 180
    z, rays = frame_rays(W, H)
    ctx.set_option(rt.OPT_COUNT_RAYS, 1)
    ctx.upload_scene(objs, lights)
    ctx.set_camera(W, H, z)
    for mode in (0, 1, 2):
        ref, hid, ht, st = oracle_mod.port_render(mode, 30, objs, lights, rays)
        for accel in (0, 1, 2, -1):
            out = ctx.render(30, mode, accel)
            assert_colours_close(out, ref)
            assert np.array_equal(out[hid < 0], ref[hid < 0])
            s = ctx.stats()
            assert (s.rays_primary, s.rays_shadow, s.rays_reflect) == (st.rays_primary, st.rays_shadow, st.rays_reflect)
            assert s.hit_pixels == st.hit_pixels and s.pixels == W * H
            gid, gt = ctx.render_aov(accel)
            assert np.array_equal(gid, hid) and np.array_equal(gt, ht)


@pytest.mark.parametrize("name", ALL_SCENES + ["synthetic"])
def test_wavefront_pipeline_equals_megakernel_and_oracle(rt, ctx, oracle_mod, name):
    """The wavefront pipeline (ray queues + ballot compaction, rt_wavefront.cu) gives the same bits as the
    megakernel on the GPU and the oracle's ids / t / ray counts, for every mode and both accelerated searches."""
    if name == "synthetic":
        objs, lights = S.synthetic_scene(5000, 500, n_lights=3)
        W, H, mb = 200, 112, 4
    else:
        objs, lights = _scene(name)
        W, H, mb = 320, 180, 30
    z, rays = frame_rays(W, H)
    ctx.set_option(rt.OPT_COUNT_RAYS, 1)
    ctx.upload_scene(objs, lights)
    ctx.set_camera(W, H, z)
    for mode in (0, 1, 2):
        ref, hid, ht, st = oracle_mod.port_render(mode, mb, objs, lights, rays)
        ctx.set_option(rt.OPT_PIPELINE, 0)
        mega = ctx.render(mb, mode, 2).copy()
        sm = ctx.stats().as_dict()
        for accel in (1, 2, 3):
            ctx.set_option(rt.OPT_PIPELINE, 1)
            out = ctx.render(mb, mode, accel)
            s = ctx.stats()
            assert s.pipeline_used == 1
            assert np.array_equal(out, mega)
            assert_colours_close(out, ref)
            assert (s.rays_primary, s.rays_shadow, s.rays_reflect) == (st.rays_primary, st.rays_shadow, st.rays_reflect)
            assert (s.traced_shadow, s.traced_reflect, s.hit_pixels) == (sm["traced_shadow"], sm["traced_reflect"], sm["hit_pixels"])
            gid, gt = ctx.render_aov(accel)
            assert np.array_equal(gid, hid) and np.array_equal(gt, ht)
    ctx.set_option(rt.OPT_PIPELINE, -1)


def test_wavefront_ragged_and_sharded(rt, oracle_mod):
    objs, lights = S.synthetic_scene(3000, 300)
    W, H = 131, 77
    z, rays = frame_rays(W, H)
    ref, *_ = oracle_mod.port_render(2, 4, objs, lights, rays)
    frame = np.full((W * H, 4), -7.0, dtype=np.float32)
    for s in range(3):
        with rt.Context(0) as c:
            c.set_option(rt.OPT_PIPELINE, 1)
            c.upload_scene(objs, lights)
            c.set_camera(W, H, z)
            c.set_shard(8, s, 3)
            c.render(4, 2, 2, out=frame)
            assert c.stats().pipeline_used == 1
    assert_colours_close(frame, ref)
    with rt.Context(0) as c:
        c.set_option(rt.OPT_PIPELINE, 1)
        c.upload_scene(objs, lights)
        c.set_rays(rays[:1000])
        assert_colours_close(c.render(4, 2, 2), ref[:1000])
        c.set_rays(rays[:0])
        assert c.render(4, 2, 2).shape == (0, 4)


def test_camera_rays_equal_host_rays(rt, ctx, oracle_mod):
    """Device ray generation is bit-identical to main()'s loop (OpenCL-Raytracer.cpp:68-72)."""
    objs, lights = S.load_reference_scene("roundedCube")
    W, H = 200, 120
    z, rays = frame_rays(W, H)
    ctx.upload_scene(objs, lights)
    ctx.set_camera(W, H, z)
    a = ctx.render(30, 2, 1).copy()
    ctx.set_option(rt.OPT_DETECT_PINHOLE, 0)  # really upload the host rays
    ctx.set_rays(rays)
    b = ctx.render(30, 2, 1)
    assert np.array_equal(a, b)


def test_pinhole_ray_vector_is_recognised(rt, ctx):
    """rt_set_rays with exactly the ray vector main() builds (OpenCL-Raytracer.cpp:68-72) -- what the unmodified drop-in
    constructor receives -- switches to device ray generation: camera grid, bands and lanes apply, nothing is uploaded.
    Any other vector (one ray nudged, a transposed grid, a sub-rectangle) is uploaded as it is."""
    objs, lights = S.synthetic_scene(3000, 300)
    W, H = 320, 181
    z, rays = frame_rays(W, H)
    ctx.set_option(rt.OPT_COUNT_RAYS, 1)
    ctx.upload_scene(objs, lights)
    ctx.set_camera(W, H, z)
    ref = ctx.render(4, 2, -1).copy()
    ctx.set_rays(rays)
    out = ctx.render(4, 2, -1)
    s = ctx.stats()
    assert np.array_equal(out, ref) and s.pixels == W * H
    assert s.pair_tests > 0 and s.node_visits > 0   # camera grid for the primary rays, BVH for the reflection rays
    nodes_cam = s.node_visits
    for bad in ("nudge", "sub"):
        r2 = rays.copy()
        if bad == "nudge":
            r2["direction"][W * 7 + 5, 0] += 0.25
        else:
            r2 = r2.reshape(H, W)[:, :W - 1].reshape(-1).copy()
        ctx.set_rays(r2)
        out2 = ctx.render(4, 2, -1)
        assert ctx.stats().node_visits > 2 * nodes_cam  # primary rays walked the BVH: the vector was uploaded
        if bad == "sub":
            assert np.array_equal(out2.reshape(H, W - 1, 4), ref.reshape(H, W, 4)[:, :W - 1])


def test_golden_vectors(rt, ctx):
    kat = json.load(open(os.path.join(HERE, "golden", "kat.json")))
    for case in kat["cases"]:
        objs, lights = _scene(case["scene"])
        W, H = case["width"], case["height"]
        z = T.camera_zdir(H)
        ctx.set_option(rt.OPT_COUNT_RAYS, 1)
        ctx.set_option(rt.OPT_SKIP_DEAD_RAYS, 0)
        ctx.upload_scene(objs, lights)
        ctx.set_camera(W, H, z)
        out = ctx.render(case["max_bounces"], case["mode"], -1)
        for px in case["pixels"]:
            got = out[px["jj"] * W + px["ii"], :3]
            ref = np.asarray(px["rgb"], dtype=np.float32)
            assert np.all(np.abs(got - ref) <= 1e-4 * np.maximum(1, np.abs(ref))), (case["name"], px, got)
        if "sum_rgb" in case:
            assert np.allclose(out[:, :3].astype(np.float64).sum(axis=0), case["sum_rgb"], rtol=1e-6)
            s = ctx.stats()
            assert [s.rays_primary, s.rays_shadow, s.rays_reflect] == case["rays"]
            assert [s.rays_primary, s.traced_shadow, s.traced_reflect] == case["rays"]  # nothing skipped
        ctx.set_option(rt.OPT_SKIP_DEAD_RAYS, 1)


@pytest.mark.parametrize("second_type", [T.BOX, T.SPHERE, T.TRIANGLE])
def test_synthetic_scene_against_oracle(rt, ctx, oracle_mod, second_type):
    objs, lights = S.synthetic_scene(6000, 500, second_type=second_type)
    W, H = 160, 90
    z, rays = frame_rays(W, H)
    ref, hid, ht, st = oracle_mod.port_render(2, 4, objs, lights, rays)
    ctx.set_option(rt.OPT_COUNT_RAYS, 1)
    ctx.upload_scene(objs, lights)
    ctx.set_camera(W, H, z)
    for accel in (0, 1, 2):
        for skip in (0, 1):
            ctx.set_option(rt.OPT_SKIP_DEAD_RAYS, skip)
            out = ctx.render(4, 2, accel)
            assert_colours_close(out, ref)
            s = ctx.stats()
            assert (s.rays_primary, s.rays_shadow, s.rays_reflect) == (st.rays_primary, st.rays_shadow, st.rays_reflect)
            if not skip:
                assert (s.traced_shadow, s.traced_reflect) == (st.rays_shadow, st.rays_reflect)
        gid, gt = ctx.render_aov(accel)
        assert np.array_equal(gid, hid) and np.array_equal(gt, ht)


def test_trace_random_rays(rt, ctx, oracle_mod):
    objs, lights = S.synthetic_scene(4000, 600)
    rays = random_rays(20000, 5)
    hid, ht = oracle_mod.port_trace(objs, rays)
    ctx.upload_scene(objs, lights)
    for accel in (0, 1, 2):
        gid, gt = ctx.trace(rays, accel)
        assert np.array_equal(gid, hid) and np.array_equal(gt, ht)


def test_full_size_scene_bvh_equals_filter(rt, ctx):
    """cfg5 geometry (100k boxes + 4096 spheres) on a 512x288 frame: BVH == bounding-sphere filter, bit for bit
    (ids, t, colours); plus a band of exact brute force."""
    objs, lights = S.synthetic_scene()
    W, H = 512, 288
    z = T.camera_zdir(H)
    ctx.upload_scene(objs, lights)
    ctx.set_camera(W, H, z)
    a = ctx.render(4, 2, 2).copy()
    ida, ta = ctx.render_aov(2)
    b = ctx.render(4, 2, 1)
    idb, tb = ctx.render_aov(1)
    assert np.array_equal(ida, idb) and np.array_equal(ta, tb) and np.array_equal(a, b)
    assert (ida >= 0).mean() > 0.5
    rays = T.camera_rays(W, H, z, rows=[100, 101, 200])
    idc, tc = ctx.trace(rays, 0)
    sel = np.r_[100 * W:102 * W, 200 * W:201 * W]
    assert np.array_equal(idc, ida[sel]) and np.array_equal(tc, ta[sel])


def test_sharded_render_equals_whole_frame(rt, oracle_mod):
    objs, lights = S.synthetic_scene(3000, 300)
    W, H = 192, 100  # 100 rows: ragged last row-block
    z = T.camera_zdir(H)
    with rt.Context(0) as c:
        c.upload_scene(objs, lights)
        c.set_camera(W, H, z)
        whole = c.render(4, 2, 2).copy()
    for n_shards, block_rows in ((2, 8), (3, 16), (8, 32)):
        frame = np.full((W * H, 4), -7.0, dtype=np.float32)
        total = 0
        for s in range(n_shards):
            with rt.Context(0) as c:
                c.upload_scene(objs, lights)
                c.set_camera(W, H, z)
                c.set_shard(block_rows, s, n_shards)
                c.render(4, 2, 2, out=frame)
                total += c.stats().pixels
        assert total == W * H
        assert np.array_equal(frame, whole)


def test_banded_readback_equals_plain_render(rt, ctx):
    """rt_render of a large whole frame is cut into bands whose device->host copy overlaps the next band's kernels
    (RT_OPT_READBACK_BANDS); the frame, the ray counts and the hit statistics must not depend on the banding."""
    objs, lights = S.synthetic_scene(3000, 300)
    W, H = 2048, 2051  # >= 4 Mpixel, ragged last band
    z = T.camera_zdir(H)
    ctx.set_option(rt.OPT_COUNT_RAYS, 1)
    ctx.upload_scene(objs, lights)
    ctx.set_camera(W, H, z)
    ctx.set_option(rt.OPT_READBACK_BANDS, 1)
    plain = ctx.render(4, 2, 2).copy()
    s0 = ctx.stats()
    ref_counts = (s0.rays_primary, s0.rays_shadow, s0.rays_reflect, s0.traced_shadow, s0.traced_reflect, s0.hit_pixels, s0.pixels)
    assert s0.pixels == W * H and s0.hit_pixels > 0
    for bands in (4, 7, 16):
        ctx.set_option(rt.OPT_READBACK_BANDS, bands)
        out = np.full((W * H, 4), -3.0, dtype=np.float32)
        ctx.render(4, 2, 2, out=out)
        s1 = ctx.stats()
        assert np.array_equal(out, plain)
        assert (s1.rays_primary, s1.rays_shadow, s1.rays_reflect, s1.traced_shadow, s1.traced_reflect, s1.hit_pixels, s1.pixels) == ref_counts


def test_two_lanes_equal_one_launch_sequence(rt, ctx, oracle_mod):
    """RT_OPT_LANES: a large camera launch runs as two halves on two streams with separate queues.  Same frame, same
    ray counts as the single launch sequence -- whole frame, ragged odd row count, and as a shard; a pixel lattice of
    the frame is also checked against the oracle."""
    objs, lights = S.synthetic_scene(3000, 300)
    W, H = 2048, 1537
    z = T.camera_zdir(H)
    ctx.set_option(rt.OPT_COUNT_RAYS, 1)
    ctx.set_option(rt.OPT_READBACK_BANDS, 1)
    ctx.upload_scene(objs, lights)
    ctx.set_camera(W, H, z)
    res = {}
    for lanes in (1, 2):
        ctx.set_option(rt.OPT_LANES, lanes)
        out = ctx.render(4, 2, 2).copy()
        s = ctx.stats()
        res[lanes] = (out, (s.rays_primary, s.rays_shadow, s.rays_reflect, s.traced_shadow, s.traced_reflect, s.hit_pixels, s.pixels))
        ctx.set_shard(16, 1, 2)
        part = np.full((W * H, 4), -2.0, dtype=np.float32)
        ctx.render(4, 2, 2, out=part)
        res[lanes] += (part,)
        ctx.set_shard(1, 0, 1)
    assert res[1][1] == res[2][1] and res[2][1][-1] == W * H
    assert np.array_equal(res[1][0], res[2][0])
    assert np.array_equal(res[1][2], res[2][2])
    rows = np.arange(3, H, 97)
    rays = T.camera_rays(W, H, z, rows=rows)[::53]
    idx = (rows[:, None] * W + np.arange(W)[None, :]).reshape(-1)[::53]
    ref, *_ = oracle_mod.port_render(2, 4, objs, lights, rays)
    assert_colours_close(res[2][0][idx], ref)


def test_banded_readback_of_a_sharded_frame(rt):
    """The multi-GPU end-to-end path bands too: a shard's row-blocks are rendered and copied in bands (split adapted from
    the previous frame's kernel : copy time).  Two shards, every band count incl. auto, two frames each (the second one
    uses the adapted split): together they must reproduce the unsharded, unbanded frame bit for bit."""
    objs, lights = S.synthetic_scene(3000, 300)
    W, H = 2048, 2069  # 130 row-blocks of 16, the last one ragged; each shard >= 2 Mpixel
    z = T.camera_zdir(H)
    with rt.Context(0) as c:
        c.set_option(rt.OPT_READBACK_BANDS, 1)
        c.upload_scene(objs, lights)
        c.set_camera(W, H, z)
        whole = c.render(4, 2, 2).copy()
    for bands in (0, 3, 8):
        out = np.full((W * H, 4), -7.0, dtype=np.float32)
        total = 0
        for s in range(2):
            with rt.Context(0) as c:
                c.set_option(rt.OPT_READBACK_BANDS, bands)
                c.set_option(rt.OPT_COUNT_RAYS, 1)
                c.upload_scene(objs, lights)
                c.set_camera(W, H, z)
                c.set_shard(16, s, 2)
                c.render(4, 2, 2, out=out)
                first = out.copy()
                c.render(4, 2, 2, out=out)  # second frame: split adapted to the measured copy time
                assert np.array_equal(out, first)
                total += c.stats().pixels
        assert total == W * H
        assert np.array_equal(out, whole)


@pytest.mark.parametrize("second_type", [T.BOX, T.TRIANGLE])
def test_device_built_tree_equals_oracle(rt, ctx, oracle_mod, second_type):
    """RT_OPT_BVH_BUILDER = 1: Morton-order radix tree built on the device (rt_lbvh.cu).  A different tree, the same hits."""
    objs, lights = S.synthetic_scene(6000, 500, second_type=second_type)
    W, H = 160, 90
    z, rays = frame_rays(W, H)
    ref, hid, ht, st = oracle_mod.port_render(2, 4, objs, lights, rays)
    ctx.set_option(rt.OPT_BVH_BUILDER, 1)
    ctx.set_option(rt.OPT_COUNT_RAYS, 1)
    ctx.upload_scene(objs, lights)
    ctx.set_camera(W, H, z)
    for pipeline in (0, 1):
        ctx.set_option(rt.OPT_PIPELINE, pipeline)
        out = ctx.render(4, 2, 2)
        assert_colours_close(out, ref)
        s = ctx.stats()
        assert s.node_visits > 0 and s.pipeline_used == pipeline
        assert (s.rays_primary, s.rays_shadow, s.rays_reflect) == (st.rays_primary, st.rays_shadow, st.rays_reflect)
    gid, gt = ctx.render_aov(2)
    assert np.array_equal(gid, hid) and np.array_equal(gt, ht)
    rr = random_rays(20000, 5)
    rid, rtt = oracle_mod.port_trace(objs, rr)
    gid, gt = ctx.trace(rr, 2)
    assert np.array_equal(gid, rid) and np.array_equal(gt, rtt)
    # two objects: the smallest tree the device builder makes (one inner node, two leaves)
    ctx.upload_scene(objs[:2], lights)
    gid, gt = ctx.trace(rr, 2)
    rid, rtt = oracle_mod.port_trace(objs[:2], rr)
    assert np.array_equal(gid, rid) and np.array_equal(gt, rtt)


def test_device_built_tree_full_size(rt, ctx):
    """cfg5 geometry: the device-built tree and the host SAH tree give the same frame, bit for bit."""
    objs, lights = S.synthetic_scene()
    W, H = 512, 288
    z = T.camera_zdir(H)
    ctx.upload_scene(objs, lights)
    ctx.set_camera(W, H, z)
    a = ctx.render(4, 2, 2).copy()
    ida, ta = ctx.render_aov(2)
    ctx.set_option(rt.OPT_BVH_BUILDER, 1)
    ctx.upload_scene(objs, lights)
    b = ctx.render(4, 2, 2)
    idb, tb = ctx.render_aov(2)
    assert np.array_equal(ida, idb) and np.array_equal(ta, tb) and np.array_equal(a, b)


def test_shared_host_frame_collects_every_shard(rt):
    """Multi-GPU end-to-end path: one page-locked shared-memory frame, every shard's rt_render copies its own row-blocks
    into it (here: four shards rendered one after the other on the same GPU)."""
    objs, lights = S.synthetic_scene(3000, 300)
    W, H = 256, 150
    z = T.camera_zdir(H)
    with rt.Context(0) as c:
        c.upload_scene(objs, lights)
        c.set_camera(W, H, z)
        whole = c.render(4, 2, 2).copy()
    owner = rt.SharedFrame(W * H)
    try:
        owner.array[:] = -5.0
        for s in range(4):
            peer = rt.SharedFrame(W * H, name=owner.name)  # what another rank does with the broadcast name
            with rt.Context(0) as c:
                c.upload_scene(objs, lights)
                c.set_camera(W, H, z)
                c.set_shard(16, s, 4)
                c.render(4, 2, 2, out=peer.array)
            peer.close()
        assert np.array_equal(owner.array, whole)
    finally:
        owner.close()


def test_quantize_is_byte_identical_to_ppm_rule(rt, ctx):
    rng = np.random.Generator(np.random.PCG64(1))
    rgba = rng.uniform(-0.2, 1.3, size=(50000, 4)).astype(np.float32)
    rgba[:8, :3] = [[0, 1, 0.5], [1 / 255, 2 / 255, 254 / 255], [0.999999, 1.000001, 0.00392156], [76 / 255] * 3,
                    [0.3, 0.375826597, 0.536895216], [255.5 / 255, 3, -1], [1e-9, 0.99607843, 0.99607844], [0.5, 0.25, 0.75]]
    q, neg = ctx.quantize_rgb8(rgba)
    ref = T.quantize_ppm(rgba)
    assert neg == int((ref < 0).sum()) and neg > 0          # the exporter would print these as negative numbers
    assert np.array_equal(q, np.maximum(ref, 0).astype(np.uint8))
    for n in (1, 2, 3, 5, 4099):                            # ragged tails of the four-pixels-per-thread kernel
        q, _ = ctx.quantize_rgb8(rgba[:n])
        assert np.array_equal(q, np.maximum(ref[:n], 0).astype(np.uint8))
    # straight from the device framebuffer of the last render: nothing is uploaded
    objs, lights = S.load_reference_scene("roundedCube")
    ctx.upload_scene(objs, lights)
    ctx.set_camera(333, 77, T.camera_zdir(77))
    frame = ctx.render(30, 2, -1)
    q, neg = ctx.quantize_rgb8()
    assert neg == 0 and np.array_equal(q, T.quantize_ppm(frame).astype(np.uint8))


def test_edge_cases(rt, ctx, oracle_mod):
    objs, lights = S.kat2_scene()
    z, rays = frame_rays(70, 37)  # neither dimension a multiple of the tile
    ctx.upload_scene(objs, lights[:0])  # zero lights
    ctx.set_camera(70, 37, z)
    ref, *_ = oracle_mod.port_render(2, 4, objs, lights[:0], rays)
    assert_colours_close(ctx.render(4, 2, 2), ref)
    ctx.upload_scene(objs[:0], lights)  # zero objects
    out = ctx.render(4, 2, -1)
    assert np.all(out == np.array([0, 0, 0, 1], dtype=np.float32))
    unknown = objs.copy()
    unknown["type"] = 7  # objects of unknown type only: the reference's switch skips them all (SAR:87,112)
    ctx.upload_scene(unknown, lights)
    for accel in (0, 1, 2):
        for pipeline in (0, 1):
            ctx.set_option(rt.OPT_PIPELINE, pipeline)
            assert np.all(ctx.render(4, 2, accel) == np.array([0, 0, 0, 1], dtype=np.float32))
    ctx.set_option(rt.OPT_PIPELINE, -1)
    ctx.upload_scene(objs, lights)
    for mb in (0, 1, 2):  # uint wrap of `bounces`
        ref, *_ = oracle_mod.port_render(2, mb, objs, lights, rays)
        assert_colours_close(ctx.render(mb, 2, 1), ref)
    ctx.set_rays(rays[:0])  # empty ray set
    assert ctx.render(4, 2, 1).shape == (0, 4)
    ctx.set_rays(rays[:33])  # ragged: not a multiple of the warp
    ref, *_ = oracle_mod.port_render(2, 4, objs, lights, rays[:33])
    assert_colours_close(ctx.render(4, 2, 2), ref)


def test_errors_are_reported_not_hidden(rt, ctx):
    objs, lights = S.kat2_scene()
    with pytest.raises(rt.RtError):
        ctx.render(4, 2, 1)  # nothing uploaded
    bad = objs.copy()
    bad[0]["mvInverse"][7] = 0.5
    with pytest.raises(rt.RtError) as e:
        ctx.upload_scene(bad, lights)
    assert e.value.code == -4
    ctx.upload_scene(objs, lights)
    with pytest.raises(rt.RtError):
        ctx.render(4, 2, 1)  # no rays / camera
    ctx.set_camera(16, 16, -10.0)
    with pytest.raises(rt.RtError):
        ctx.render(4, 9, 1)  # bad mode


# ------------------------------------------------------------------ round 2: the benchmarked configuration, edge cases on the GPU
def _lattice(W, H, n_side, seed):
    """A seeded, frame-covering pixel subset: an n_side x n_side jittered lattice of distinct pixels."""
    rng = np.random.Generator(np.random.PCG64(seed))
    jj = np.minimum((np.arange(n_side) * H // n_side + rng.integers(0, max(H // n_side, 1), n_side)), H - 1)
    ii = np.minimum((np.arange(n_side) * W // n_side + rng.integers(0, max(W // n_side, 1), n_side)), W - 1)
    return jj.astype(np.int64), ii.astype(np.int64)


def _headline_against_oracle(rt, oracle_mod, n_side):
    objs, lights = S.synthetic_scene()  # cfg5-box: 100 000 boxes + 4 096 spheres, 2 lights
    W, H, mb = 7680, 4320, 4
    z = T.camera_zdir(H)
    jj, ii = _lattice(W, H, n_side, 0xC0FFEE)
    rays = T.camera_rays(W, H, z, rows=jj).reshape(len(jj), W)[:, ii].reshape(-1).copy()
    idx = (jj[:, None] * W + ii[None, :]).reshape(-1)
    ref, hid, ht, st = oracle_mod.port_render(2, mb, objs, lights, rays, nthreads=os.cpu_count() or 1)
    with rt.Context(0) as c:  # exactly as bench.py drives it: camera, depth 4, everything else on auto
        c.upload_scene(objs, lights)
        c.set_camera(W, H, z)
        frame = c.render(mb, 2, -1)
        s = c.stats()
        assert s.pipeline_used == 1 and s.accel_used == rt.ACCEL_GRID
        gid, gt = c.render_aov(-1)
        assert np.array_equal(gid[idx], hid) and np.array_equal(gt[idx], ht)  # hit ids and t: bit-exact
        assert_colours_close(frame[idx], ref)                                 # colours 1e-4, PPM <= 1 LSB
        assert np.array_equal(frame[idx][hid < 0], ref[hid < 0])              # missed pixels exactly (0,0,0,1)
        # the same pixels as caller-supplied rays (the drop-in ctor's path): same bits, and the ray counts by class
        c.set_option(rt.OPT_COUNT_RAYS, 1)
        c.set_rays(rays)
        sub = c.render(mb, 2, -1)
        s = c.stats()
        assert np.array_equal(sub, frame[idx])
        assert (s.rays_primary, s.rays_shadow, s.rays_reflect, s.hit_pixels) == (st.rays_primary, st.rays_shadow, st.rays_reflect, st.hit_pixels)
    return len(idx)


def test_headline_config_against_oracle(rt, oracle_mod):
    """The benchmarked configuration itself -- cfg5-box, set_camera(7680, 4320), depth 4, auto accel / pipeline / bands /
    lanes -- against the oracle on a seeded 64 x 64 pixel subset."""
    assert _headline_against_oracle(rt, oracle_mod, 64) == 4096


@pytest.mark.skipif(not os.environ.get("RT_SLOW"), reason="65 536-pixel subset: minutes of CPU brute force (RT_SLOW=1)")
def test_headline_config_against_oracle_65536(rt, oracle_mod):
    assert _headline_against_oracle(rt, oracle_mod, 256) == 65536


def test_headline_frame_all_searches_agree_at_full_resolution(rt, ctx):
    """7680 x 4320, cfg5-box: perspective grids + BVH == BVH == bounding-sphere filter on the WHOLE frame (ids, t, colours),
    and exact brute force on 64 rows."""
    objs, lights = S.synthetic_scene()
    W, H = 7680, 4320
    z = T.camera_zdir(H)
    ctx.upload_scene(objs, lights)
    ctx.set_camera(W, H, z)
    a = ctx.render(4, 2, 3).copy()
    ida, ta = ctx.render_aov(3)
    b = ctx.render(4, 2, 2)
    assert np.array_equal(a, b)
    del b
    idb, tb = ctx.render_aov(2)
    assert np.array_equal(ida, idb) and np.array_equal(ta, tb)
    b = ctx.render(4, 2, 1)
    assert np.array_equal(a, b)
    del b
    idb, tb = ctx.render_aov(1)
    assert np.array_equal(ida, idb) and np.array_equal(ta, tb)
    rows = np.arange(64) * (H // 64) + 11
    rays = T.camera_rays(W, H, z, rows=rows)
    idc, tc = ctx.trace(rays, 0)
    sel = (rows[:, None] * W + np.arange(W)[None, :]).reshape(-1)
    assert np.array_equal(idc, ida[sel]) and np.array_equal(tc, ta[sel])


@pytest.mark.parametrize("pipeline", [0, 1])
def test_tie_rules_coincident_objects_gpu(rt, ctx, oracle_mod, pipeline):
    """Exact ties on the GPU's own merge code (filter_flush's owner merge, the BVH kernel's deferred candidates, the grid
    kernel's list walk): duplicated spheres -- highest index wins (SAR:108) -- and duplicated boxes -- lowest index wins
    (SAR:147) -- plus a cloud of coincident pairs large enough for full 32-wide exact batches."""
    base, lights = S.kat2_scene()
    small = np.concatenate([base[[0]], base[[1]], base[[0]], base[[1]], base[[2]], base[[2]]])
    cloud, lights2 = S.synthetic_scene(600, 200, n_lights=2)
    cloud = np.concatenate([cloud, cloud[::-1], cloud[::3]])  # every object two or three times, in shuffled index order
    for objs, lts, (W, H) in ((small, lights, (160, 90)), (cloud, lights2, (192, 108))):
        z, rays = frame_rays(W, H)
        ref, hid, ht, st = oracle_mod.port_render(2, 4, objs, lts, rays)
        ctx.set_option(rt.OPT_PIPELINE, pipeline)
        ctx.set_option(rt.OPT_COUNT_RAYS, 1)
        ctx.upload_scene(objs, lts)
        ctx.set_camera(W, H, z)
        for accel in (0, 1, 2, 3):
            out = ctx.render(4, 2, accel)
            assert ctx.stats().pipeline_used == pipeline
            assert_colours_close(out, ref)
            gid, gt = ctx.render_aov(accel)
            assert np.array_equal(gid, hid) and np.array_equal(gt, ht)
    assert set(np.unique(hid)) != {-1}


@pytest.mark.parametrize("pipeline", [0, 1])
def test_degenerate_rays_gpu(rt, ctx, oracle_mod, pipeline):
    """Axis-parallel rays (a direction component exactly 0: the copysign branch of the slab test, SAR:36-43), origins
    inside a box / a sphere, an object of unknown type -- through both pipelines and every search."""
    objs, lights = S.kat2_scene()
    r = np.zeros(7, dtype=T.RAY_DTYPE)
    r["start"][:, 3] = 1.0
    r["start"][:, :3] = [(0, 0, 0), (0, 0, 0), (1, 1, 0), (0, 0, -10), (3, 2, -7), (5, 0, -10), (0, 5, -10)]
    r["direction"][:, :3] = [(0, 0, -1), (0, 0, 1), (0, 0, -3), (0, 1, 0), (1, 0, 0), (-1, 0, 0), (0, -2, 0)]
    r = np.concatenate([r] * 40)  # more than one warp, ragged
    weird = np.concatenate([objs, objs[[1]]])
    weird[3]["type"] = 7
    hid, ht = oracle_mod.port_trace(weird, r)
    ref, *_ = oracle_mod.port_render(2, 3, weird, lights, r)
    assert (hid >= 0).sum() >= 5 * 40 and not (hid == 3).any()
    ctx.set_option(rt.OPT_PIPELINE, pipeline)
    ctx.upload_scene(weird, lights)
    ctx.set_rays(r)
    for accel in (0, 1, 2, 3):
        gid, gt = ctx.render_aov(accel)
        assert np.array_equal(gid, hid) and np.array_equal(gt, ht)
        assert_colours_close(ctx.render(3, 2, accel), ref)
        tid, tt = ctx.trace(r, accel)
        assert np.array_equal(tid, hid) and np.array_equal(tt, ht)


def test_perspective_grids_lights_everywhere(rt, ctx, oracle_mod):
    """Grids around a light buried in the object cloud, a light on the camera apex, a directional light (no grids: BVH),
    more lights than light codes (lights 15.. keep the BVH), several grid resolutions; SHADE and SAR modes."""
    objs, lights = S.synthetic_scene(4000, 400, n_lights=4)
    lights = lights.copy()
    lights[0]["position"] = (0.3, -0.2, -60.0, 1.0)
    lights[1]["position"] = (0.2, 0.9, 0.4, 0.0)
    lights[2]["position"] = (0.0, 0.0, 0.0, 1.0)
    many = np.concatenate([lights] * 5)[:18]
    W, H = 160, 90
    z, rays = frame_rays(W, H)
    ctx.set_option(rt.OPT_PIPELINE, 1)
    ctx.set_option(rt.OPT_COUNT_RAYS, 1)
    for lts, mb in ((lights, 3), (many, 1)):
        for mode in (1, 2):
            ref, hid, ht, st = oracle_mod.port_render(mode, mb, objs, lts, rays)
            for res in (32, 512, 1024):
                ctx.set_option(rt.OPT_GRID_RES, res)
                ctx.upload_scene(objs, lts)
                ctx.set_camera(W, H, z)
                out = ctx.render(mb, mode, 3)
                s = ctx.stats()
                assert s.accel_used == rt.ACCEL_GRID and s.pair_tests > 0
                assert_colours_close(out, ref)
                assert (s.rays_primary, s.rays_shadow, s.rays_reflect) == (st.rays_primary, st.rays_shadow, st.rays_reflect)
                if lts is many:
                    break


def test_zero_copy_frame_equals_copied_frame(rt, ctx):
    """rt_render into page-locked, device-addressable memory: the kernels store the pixels straight into the caller's frame
    (RT_OPT_ZERO_COPY).  Same bits as the copy path -- whole frame, ragged shard, explicit rays, both pipelines."""
    objs, lights = S.synthetic_scene(3000, 300)
    W, H = 1024, 515
    z = T.camera_zdir(H)
    ctx.upload_scene(objs, lights)
    ctx.set_camera(W, H, z)
    pinned = rt.PinnedBuffer((W * H, 4))
    try:
        for pipeline in (1, 0):
            ctx.set_option(rt.OPT_PIPELINE, pipeline)
            ctx.set_option(rt.OPT_ZERO_COPY, 0)
            ref = ctx.render(4, 2, -1).copy()
            ctx.set_option(rt.OPT_ZERO_COPY, 1)
            pinned.array[:] = -9.0
            ctx.render(4, 2, -1, out=pinned.array)
            assert np.array_equal(pinned.array, ref)
            pinned.array[:] = -9.0
            for s in range(3):
                ctx.set_shard(16, s, 3)
                ctx.render(4, 2, -1, out=pinned.array)
            ctx.set_shard(1, 0, 1)
            assert np.array_equal(pinned.array, ref)
        ctx.set_option(rt.OPT_PIPELINE, -1)
        ctx.set_option(rt.OPT_ZERO_COPY, -1)
        pageable = np.full((W * H, 4), -9.0, dtype=np.float32)  # not device-addressable: the copy path, silently
        ctx.render(4, 2, -1, out=pageable)
        assert np.array_equal(pageable, ref)
    finally:
        pinned.close()


def test_uniform_grid_walk_equals_bvh(rt, ctx, oracle_mod):
    """RT_OPT_GRID3_DENSITY: the rays without a common apex walk a uniform 3-D grid (DDA, rt_grid3.cuh) instead of the BVH
    (off by default).  Same hits, same frame -- camera rays, caller rays, a shard."""
    objs, lights = S.synthetic_scene(6000, 500, n_lights=3)
    lights = lights.copy()
    lights[1]["position"] = (0.3, 0.8, 0.5, 0.0)  # directional: its shadow rays are class B too
    W, H = 200, 112
    z, rays = frame_rays(W, H)
    ref, hid, ht, st = oracle_mod.port_render(2, 4, objs, lights, rays)
    ctx.set_option(rt.OPT_PIPELINE, 1)
    ctx.set_option(rt.OPT_COUNT_RAYS, 1)
    for density in (50, 200, 1600):
        ctx.set_option(rt.OPT_GRID3_DENSITY, density)
        ctx.upload_scene(objs, lights)
        ctx.set_camera(W, H, z)
        out = ctx.render(4, 2, 3)
        s = ctx.stats()
        assert_colours_close(out, ref)
        assert (s.rays_primary, s.rays_shadow, s.rays_reflect) == (st.rays_primary, st.rays_shadow, st.rays_reflect)
        ctx.set_option(rt.OPT_DETECT_PINHOLE, 0)  # caller rays: the primary rays are class B too
        ctx.set_rays(rays)
        assert np.array_equal(ctx.render(4, 2, 3), out)
        gid, gt = ctx.render_aov(3)
        assert np.array_equal(gid, hid) and np.array_equal(gt, ht)
    rr = random_rays(20000, 5)
    rid, rtt = oracle_mod.port_trace(objs, rr)
    ctx.set_rays(rr)
    gid, gt = ctx.render_aov(3)
    assert np.array_equal(gid, rid) and np.array_equal(gt, rtt)


def test_scene_that_changes_every_frame(rt, oracle_mod):
    """RT_OPT_BVH_BUILDER = 1 prepares the scene on the device -- repack, per-object bounds (rt_bounds.cu), tree
    (rt_lbvh.cu), grids (rt_pgrid.cu).  Objects move between frames; every frame must equal, bit for bit, the frame of a fresh
    context that prepared the same scene on the host (binned SAH), and the first one the oracle."""
    objs, lights = S.synthetic_scene(6000, 500, n_lights=2)
    W, H = 192, 108
    z, rays = frame_rays(W, H)
    rng = np.random.Generator(np.random.PCG64(3))
    with rt.Context(0) as dyn:
        dyn.set_option(rt.OPT_BVH_BUILDER, 1)
        dyn.set_camera(W, H, z)
        for frame in range(4):
            if frame:
                # rigid motion of a third of the objects: mv' = T(d) mv, mvInverse' = mvInverse T(-d)
                mv = objs["mv"].reshape(-1, 4, 4).copy()
                mi = objs["mvInverse"].reshape(-1, 4, 4).copy()
                sel = rng.random(len(objs)) < 0.33
                d = (rng.normal(size=(len(objs), 3)) * 0.7).astype(np.float32)
                d[~sel] = 0
                mv[:, 3, :3] += d
                mi[:, 3, :3] -= np.einsum("nkj,nk->nj", mi[:, :3, :3], d)
                objs = objs.copy()
                objs["mv"], objs["mvInverse"] = mv.reshape(-1, 16), mi.reshape(-1, 16)
            dyn.upload_scene(objs, lights)
            for pipeline in (1, 0):
                dyn.set_option(rt.OPT_PIPELINE, pipeline)
                got = dyn.render(4, 2, -1).copy()
                with rt.Context(0) as fresh:
                    fresh.set_option(rt.OPT_PIPELINE, pipeline)
                    fresh.upload_scene(objs, lights)
                    fresh.set_camera(W, H, z)
                    assert np.array_equal(got, fresh.render(4, 2, -1))
            if frame == 0:
                ref, *_ = oracle_mod.port_render(2, 4, objs, lights, rays)
                assert_colours_close(got, ref)
        # the device path rejects what the host path rejects, naming the object
        bad = objs.copy()
        bad[77]["mvInverse"][7] = 0.5
        with pytest.raises(rt.RtError) as e:
            dyn.upload_scene(bad, lights)
        assert e.value.code == -4 and "object 77" in str(e.value)
        bad = objs.copy()
        bad[5]["mvInverse"][:] = 0
        bad[5]["mvInverse"][15] = 1
        with pytest.raises(rt.RtError) as e:
            dyn.upload_scene(bad, lights)
        assert e.value.code == -4 and "object 5" in str(e.value)
        unknown = objs.copy()
        unknown["type"][::3] = 9  # types the kernels' switch skips: leaves that are never hit
        dyn.upload_scene(unknown, lights)
        with rt.Context(0) as fresh:
            fresh.upload_scene(unknown, lights)
            fresh.set_camera(W, H, z)
            assert np.array_equal(dyn.render(4, 2, -1), fresh.render(4, 2, -1))


def test_tail_launch_equals_generation_chain(rt, ctx, oracle_mod):
    """RT_OPT_TAIL_GEN: from generation g on one launch finishes every live path (wave_tail) instead of a trace / advance pair
    per generation.  Same frame, same ray counts, for every g, three lights (stale-specular walk), SHADE and SAR, both the
    BVH and the grids."""
    objs, lights = S.synthetic_scene(5000, 500, n_lights=3)
    W, H = 200, 112
    z, rays = frame_rays(W, H)
    ctx.set_option(rt.OPT_PIPELINE, 1)
    ctx.set_option(rt.OPT_COUNT_RAYS, 1)
    ctx.upload_scene(objs, lights)
    ctx.set_camera(W, H, z)
    for mode, mb in ((2, 4), (1, 0)):
        ref, hid, ht, st = oracle_mod.port_render(mode, mb, objs, lights, rays)
        for accel in (2, 3):
            ctx.set_option(rt.OPT_TAIL_GEN, 0)
            chain = ctx.render(mb, mode, accel).copy()
            s0 = ctx.stats()
            assert_colours_close(chain, ref)
            for g in (1, 2, 3, 4, 5, 8, 40):
                ctx.set_option(rt.OPT_TAIL_GEN, g)
                out = ctx.render(mb, mode, accel)
                s = ctx.stats()
                assert np.array_equal(out, chain), (mode, accel, g)
                counts = (s.rays_primary, s.rays_shadow, s.rays_reflect, s.traced_shadow, s.traced_reflect, s.hit_pixels)
                assert counts == (s0.rays_primary, s0.rays_shadow, s0.rays_reflect, s0.traced_shadow, s0.traced_reflect, s0.hit_pixels)
                assert s.kernel_launches <= s0.kernel_launches
    ctx.set_option(rt.OPT_TAIL_GEN, -1)


def test_two_contexts_one_process_filter_kernel(rt):
    """One process, one context per device (what CudaRaytracer::UseDevices does): the wavefront filter kernel needs the 64 KB
    dynamic shared-memory opt-in on EVERY device it runs on, not only on the first one (advisor finding, round 1).  Needs two
    GPUs; bench.py runs the same check at N > 1 where the driver sees it (`one_process_multi_gpu`)."""
    import ctypes

    n = ctypes.c_int(0)
    assert rt.load().rt_device_count(ctypes.byref(n)) == 0
    if n.value < 2:
        pytest.skip("needs two GPUs")
    objs, lights = S.synthetic_scene(3000, 300)
    frames = []
    for dev in (1, 0, 1):
        with rt.Context(dev) as c:
            c.set_option(rt.OPT_PIPELINE, 1)
            c.upload_scene(objs, lights)
            c.set_camera(320, 180, T.camera_zdir(180))
            frames.append(c.render(4, 2, 1).copy())
    assert np.array_equal(frames[0], frames[1]) and np.array_equal(frames[0], frames[2])


def test_clustered_scene_keeps_the_bvh(rt, ctx, oracle_mod):
    """The uniform grid is only taken when its lists stay short.  A scene with one tight cluster far from a few stragglers
    (the teapot in the stadium) would put hundreds of objects into a handful of cells: the builder must refuse it and the
    reflection rays must walk the BVH (node visits > 0) -- and the frame must be right either way."""
    objs, lights = S.synthetic_scene(3000, 300)
    mv = objs["mv"].reshape(-1, 4, 4).copy()
    mi = objs["mvInverse"].reshape(-1, 4, 4).copy()
    centre = mv[:, 3, :3].copy()
    target = np.array([0.0, 0.0, -60.0], dtype=np.float32)
    d = ((target - centre) * np.float32(0.97)).astype(np.float32)   # pull 97 % of the way into one spot ...
    d[:8] = 0                                                        # ... except eight stragglers that keep the scene wide
    mv[:, 3, :3] += d
    mi[:, 3, :3] -= np.einsum("nkj,nk->nj", mi[:, :3, :3], d)
    objs["mv"], objs["mvInverse"] = mv.reshape(-1, 16), mi.reshape(-1, 16)
    W, H = 160, 90
    z, rays = frame_rays(W, H)
    ref, hid, ht, st = oracle_mod.port_render(2, 3, objs, lights, rays)
    ctx.set_option(rt.OPT_PIPELINE, 1)
    ctx.set_option(rt.OPT_COUNT_RAYS, 1)
    ctx.upload_scene(objs, lights)
    ctx.set_camera(W, H, z)
    out = ctx.render(3, 2, 3)
    s = ctx.stats()
    assert_colours_close(out, ref)
    assert s.traced_reflect > 0 and s.node_visits > 0   # BVH, not the grid walk (which counts cells in node_visits too, but
    gid, gt = ctx.render_aov(3)                          # far fewer: checked against the uniform scene below)
    assert np.array_equal(gid, hid) and np.array_equal(gt, ht)
    visits_clustered = s.node_visits / max(s.traced_reflect, 1)
    uni, _ = S.synthetic_scene(3000, 300)
    ctx.upload_scene(uni, lights)
    ctx.render(3, 2, 3)
    s2 = ctx.stats()
    assert s2.node_visits / max(s2.traced_reflect, 1) < visits_clustered  # the uniform scene takes the grid: cells, not node steps


@pytest.mark.parametrize("size", [(1001, 777), (2561, 1439), (64, 3)])
def test_readback_plans_on_odd_frames(rt, ctx, size):
    """rt_render's readback plan (launch sequences of shrinking size on three streams, two copy streams) on frames whose sizes
    divide by nothing: the default plan into a pageable and into a pinned buffer, and explicit sequence counts, all give the
    frame that rt_render_device + rt_readback give."""
    W, H = size
    objs, lights = S.synthetic_scene(20000, 1000)
    ctx.upload_scene(objs, lights)
    ctx.set_camera(W, H, T.camera_zdir(H))
    a = ctx.render(4, 2, -1).copy()
    ctx.render_device(4, 2, -1)
    b = np.empty((W * H, 4), np.float32)
    ctx.readback(b)
    assert np.array_equal(a, b)
    pinned = rt.PinnedBuffer((W * H, 4))
    try:
        ctx.render(4, 2, -1, out=pinned.array)
        assert np.array_equal(pinned.array, b)
    finally:
        pinned.close()
    for k in (1, 3, 7, 16):
        ctx.set_option(rt.OPT_READBACK_BANDS, k)
        assert np.array_equal(ctx.render(4, 2, -1), b), k
