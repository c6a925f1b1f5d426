#!/usr/bin/env python
"""End-to-end time (rt_render into a pinned frame) of the shard one rank of N renders, against the staggered readback plan
(RT_STAGGER_K sequences, each RT_STAGGER_Q of the one before, round-robin over RT_STAGGER_S streams).  One GPU emulates one
rank: no contention for the host's ingest bandwidth."""
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


def measure(c):
    ms, km = [], []
    for _ in range(8):
        t0 = time.perf_counter()
        c.render(4, 2, -1, out=buf.array)
        ms.append((time.perf_counter() - t0) * 1e3)
        km.append(c.stats().kernel_ms)
    return float(np.median(ms[3:])), float(np.median(km[3:]))


with lib.Context(0) as c:
    c.upload_scene(objs, lights)
    c.set_camera(W, H, z)
    for shards in (1, 2, 4, 8):
        c.set_shard(16 if shards > 1 else 1, 0, shards)
        for k in ("RT_STAGGER_K", "RT_STAGGER_Q", "RT_STAGGER_S"):
            os.environ.pop(k, None)
        e, km = measure(c)
        print("shards %d | auto: e2e %6.2f (kernels %6.2f)" % (shards, e, km), flush=True)
        os.environ["RT_STAGGER_K"] = "0"
        e, km = measure(c)
        print("shards %d | two-lane bands (round-2 default before): e2e %6.2f (kernels %6.2f)" % (shards, e, km), flush=True)
        for s in (2, 3):
            for k in (2, 3, 4, 6, 8):
                line = []
                for q in (0.6, 0.7, 0.8, 0.9):
                    os.environ.update(RT_STAGGER_K=str(k), RT_STAGGER_Q=str(q), RT_STAGGER_S=str(s))
                    e, km = measure(c)
                    line.append("q %.1f: %6.2f" % (q, e))
                print("shards %d | S %d K %d | " % (shards, s, k) + " | ".join(line), flush=True)
buf.close()
