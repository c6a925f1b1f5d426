#!/usr/bin/env python
"""Kernel time of the headline frame -- whole, and the 1/2, 1/4, 1/8 shards one rank of a multi-GPU job renders -- against the
generation at which the single tail launch takes over (RT_OPT_TAIL_GEN).  One GPU emulates one rank of N."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "opencl-raytracer_b200")]
from pyrt import lib, scenes, types  # noqa: E402

objs, lights = scenes.synthetic_scene()
W, H = 7680, 4320
z = types.camera_zdir(H)
with lib.Context(0) as c:
    c.upload_scene(objs, lights)
    c.set_camera(W, H, z)
    ref = None
    for shards in (1, 2, 4, 8):
        c.set_shard(16 if shards > 1 else 1, 0, shards)
        for lanes in ((1, 2) if shards > 1 else (2,)):
            c.set_option(lib.OPT_LANES, lanes)
            line = []
            for tail in (0, 3, 4, 6, 7, 9, 10):
                c.set_option(lib.OPT_TAIL_GEN, tail)
                ms = []
                for _ in range(6):
                    c.render_device(4, 2, -1)
                    ms.append(c.stats().kernel_ms)
                line.append("tail %2d: %6.2f ms (%d launches)" % (tail, float(np.median(ms[2:])), c.stats().kernel_launches))
            print("shards %d lanes %d | " % (shards, lanes) + " | ".join(line), flush=True)
