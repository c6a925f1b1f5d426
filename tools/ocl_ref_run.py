#!/usr/bin/env python
"""GPU-box script: run the reference's UNCHANGED OpenCL kernel source on the box's own OpenCL
platform, if one enumerates, and

  * pin the CPU oracle against it: oracle/_ref (the same source compiled as C++ for the host) must agree with what a
    real OpenCL compiler makes of the kernels (OpenCL allows contraction and a few ulp on div/sqrt/pow, so the
    comparison is the north star's 1e-4 relative / 1 PPM LSB, and the count of differing pixels is reported);
  * time OpenCLRaytracer's ctor + Render() equivalent (JIT build, kernel, launch + blocking read) per config.

Prints one JSON line per config (and a first line with the probe result).  Test/baseline infrastructure only.
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "opencl-raytracer_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)

import oracle  # noqa: E402
from pyrt import scenes, types  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=3)
    ap.add_argument("--cfg5-rows", type=int, default=8, help="rows of the 8K frame traced by the brute-force reference kernel")
    args = ap.parse_args()

    ok, info = oracle.ocl_probe()
    print(json.dumps({"opencl_platform": ok, "info": info}), flush=True)
    if not ok:
        return

    table = [
        ("cfg1", "simpleSphere", 2560, 1440, 30, 2, None),
        ("cfg2", "multipleSpheres", 1920, 1080, 3, 2, None),
        ("cfg3", "roundedCube", 3840, 2160, 30, 2, None),
        ("cfg3-shade", "roundedCube", 3840, 2160, 30, 1, None),
        ("cfg4", "simpleScene", 3840, 2160, 5, 2, None),
        ("cfg4-3l", "simpleScene3L", 3840, 2160, 5, 2, None),
        ("cfg5-box", None, 7680, 4320, 4, 2, args.cfg5_rows),
    ]
    for name, scene, W, H, mb, mode, rows in table:
        if scene is None:
            objs, lights = scenes.synthetic_scene(100_000, 4096)
        else:
            objs, lights = scenes.load_reference_scene(scene)
        z = types.camera_zdir(H)
        if rows is None:
            rays = types.camera_rays(W, H, z)
            sample = "full frame"
        else:
            r0 = H // 2 - rows // 2
            rays = types.camera_rays(W, H, z, rows=np.arange(r0, r0 + rows))
            sample = "%d rows x %d px band at row %d" % (rows, W, r0)
        try:
            out, t = oracle.ocl_render(mode, mb, objs, lights, rays, frames=args.frames)
        except RuntimeError as e:
            print(json.dumps({"config": name, "error": str(e)}), flush=True)
            continue
        # reference ray count + parity against the host-compiled reference kernels on a bounded subset
        n = len(rays)
        sub = np.arange(n) if len(objs) < 100 and n <= 4_000_000 else np.linspace(0, n - 1, 4096).astype(np.int64)
        if len(objs) < 100 and n > 4_000_000:
            sub = np.arange(0, n, 4)
        ref = oracle.ref_render(mode, mb, objs, lights, rays[sub]) if oracle.have_ref() else oracle.port_render(mode, mb, objs, lights, rays[sub])[0]
        a, b = out[sub, :3].astype(np.float64), ref[:, :3].astype(np.float64)
        d = np.abs(a - b)
        tol = 1e-4 * np.maximum(1.0, np.abs(b))
        qa, qb = types.quantize_ppm(out[sub]), types.quantize_ppm(ref)
        _, _, _, st = oracle.port_render(mode, mb, objs, lights, rays[sub])
        nrays_sub = st.rays_primary + st.rays_shadow + st.rays_reflect
        nrays = nrays_sub * (n / len(sub))
        print(json.dumps({
            "config": name, "sample": sample, "pixels": int(n), "objects": int(len(objs)), "lights": int(len(lights)),
            "max_bounces": mb, "mode": mode, "kernel_ms": t["kernel_ms"], "render_ms": t["render_ms"], "build_ms": t["build_ms"],
            "reference_rays": nrays, "mrays_per_s_kernel": nrays / (t["kernel_ms"] * 1e-3) / 1e6,
            "mrays_per_s_render": nrays / (t["render_ms"] * 1e-3) / 1e6,
            "parity_vs_host_compiled_reference": {
                "pixels_compared": int(len(sub)), "bit_identical_pixels": int((a == b).all(axis=1).sum()),
                "max_abs_diff": float(d.max()), "pixels_over_1e-4": int((d > tol).any(axis=1).sum()),
                "ppm_max_lsb": int(np.abs(qa - qb).max()), "ppm_pixels_differing": int((qa != qb).any(axis=1).sum())},
        }), flush=True)


if __name__ == "__main__":
    main()
