#!/usr/bin/env python
"""Small renders through every kernel family (wavefront grids + grid walk / bvh / filter / brute, tail launch, megakernel, both
scene builders), each checked
against the oracle (ids / t bit-exact, colours 1e-4).  Written to be run under compute-sanitizer
(`compute-sanitizer --tool memcheck|racecheck python tools/sanitize.py`); the tool is closed on this GPU pool, so the
script only serves as a quick all-paths parity check on the box."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "opencl-raytracer_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)

import oracle  # noqa: E402
from pyrt import lib, scenes, types  # noqa: E402

objs, lights = scenes.synthetic_scene(600, 60)
W, H = 96, 40
z = types.camera_zdir(H)
rays = types.camera_rays(W, H, z)
ref, hid, ht, st = oracle.port_render(2, 3, objs, lights, rays)
for pipeline in (1, 0):  # wavefront, megakernel
    for builder in (0, 1):
        with lib.Context(0) as ctx:
            ctx.set_option(lib.OPT_PIPELINE, pipeline)
            ctx.set_option(lib.OPT_BVH_BUILDER, builder)
            ctx.upload_scene(objs, lights)
            ctx.set_camera(W, H, z)
            for accel in ((0, 1, 2, 3) if builder == 0 else (2, 3)):
                out = ctx.render(3, 2, accel)
                gid, gt = ctx.render_aov(accel)
                assert np.array_equal(gid, hid) and np.array_equal(gt, ht), (pipeline, builder, accel)
                d = np.abs(out[:, :3].astype(np.float64) - ref[:, :3])
                assert (d <= 1e-4 * np.maximum(1.0, np.abs(ref[:, :3]))).all(), (pipeline, builder, accel)
print("sanitize run ok")
