#!/usr/bin/env python
"""End-to-end time (rt_render into a pinned frame) of the shard one rank of N renders, against the number of readback bands.
One GPU emulates one rank: no contention for the host's ingest bandwidth, so the multi-rank optimum has at most this many bands."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "opencl-raytracer_b200")]
from pyrt import lib, scenes, types  # noqa: E402

objs, lights = scenes.synthetic_scene()
W, H = 7680, 4320
z = types.camera_zdir(H)
buf = lib.PinnedBuffer((W * H, 4))
with lib.Context(0) as c:
    c.upload_scene(objs, lights)
    c.set_camera(W, H, z)
    for shards in (1, 2, 4, 8):
        c.set_shard(16 if shards > 1 else 1, 0, shards)
        line = []
        for bands in (1, 2, 3, 4, 6):
            c.set_option(lib.OPT_READBACK_BANDS, bands)
            ms, km = [], []
            for _ in range(7):
                t0 = time.perf_counter()
                c.render(4, 2, -1, out=buf.array)
                ms.append((time.perf_counter() - t0) * 1e3)
                km.append(c.stats().kernel_ms)
            line.append("bands %d: e2e %6.2f (kernels %6.2f)" % (bands, float(np.median(ms[2:])), float(np.median(km[2:]))))
        print("shards %d | " % shards + " | ".join(line), flush=True)
buf.close()
