#!/usr/bin/env python
"""Generates tests/golden/kat.json: known-answer vectors produced by the reference's OWN kernel sources
compiled as C++ (oracle/_ref/libsar_ref.so, see oracle/Makefile).  Needs /root/reference at build time, so it
runs in the build container only; the JSON it writes is committed.  KAT-1 / KAT-2 are the cases of SURVEY.md
section 4 (the hand-checked centre pixel of simpleSphere.txt is among them)."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "opencl-raytracer_b200")]
import oracle  # noqa: E402
from pyrt import scenes as S  # noqa: E402
from pyrt import types as T  # noqa: E402


def case(name, scene, W, H, mb, mode, pixels, whole=False):
    objs, lights = S.kat2_scene() if scene == "kat2" else S.load_reference_scene(scene)
    z = T.camera_zdir(H)
    rays = T.camera_rays(W, H, z)
    if whole:
        rgba = oracle.ref_render(mode, mb, objs, lights, rays)
        get = lambda ii, jj: rgba[jj * W + ii]
    else:
        rows = sorted({jj for _, jj in pixels})
        sub = oracle.ref_render(mode, mb, objs, lights, T.camera_rays(W, H, z, rows=rows))
        get = lambda ii, jj: sub[rows.index(jj) * W + ii]
    out = {"name": name, "scene": scene, "width": W, "height": H, "max_bounces": mb, "mode": mode, "zdir": float(z),
           "pixels": [{"ii": ii, "jj": jj, "rgb": [float(np.float32(v)) for v in get(ii, jj)[:3]]} for ii, jj in pixels]}
    if whole:
        _, hid, _, st = oracle.port_render(mode, mb, objs, lights, rays)
        out["hit_pixels"] = int((rgba[:, :3] != 0).any(axis=1).sum()) if mode != 0 else int((hid >= 0).sum())
        out["sum_rgb"] = [float(v) for v in rgba[:, :3].astype(np.float64).sum(axis=0)]
        out["rays"] = [int(st.rays_primary), int(st.rays_shadow), int(st.rays_reflect)]
    return out


cases = [
    case("kat1_simpleSphere_1440p", "simpleSphere", 2560, 1440, 30, 2,
         [(1280, 720), (1280, 700), (1300, 720), (1240, 690), (1338, 720), (0, 0)]),
    case("kat2_box_spheres_640x360_d4", "kat2", 640, 360, 4, 2,
         [(320, 180), (380, 150), (394, 122), (470, 70), (440, 120), (186, 262), (170, 300), (10, 10)], whole=True),
    case("roundedCube_480x270_d30", "roundedCube", 480, 270, 30, 2, [(240, 135), (200, 120), (300, 160), (150, 100)], whole=True),
    case("roundedCube_480x270_shade", "roundedCube", 480, 270, 30, 1, [(240, 135), (200, 120), (300, 160)], whole=True),
    case("multipleSpheres_480x270_d3", "multipleSpheres", 480, 270, 3, 2, [(240, 135), (100, 135), (380, 135)], whole=True),
    case("simpleScene3L_480x270_d5", "simpleScene3L", 480, 270, 5, 2, [(240, 135), (200, 100), (300, 170)], whole=True),
    case("simpleScene_480x270_hittest", "simpleScene", 480, 270, 0, 0, [(240, 135), (200, 100)], whole=True),
]
path = os.path.join(ROOT, "tests", "golden", "kat.json")
json.dump({"generator": "tools/make_kat.py (oracle/_ref = reference kernels compiled as C++, -O2 -ffp-contract=off)",
           "cases": cases}, open(path, "w"), indent=1)
print("wrote", path)
