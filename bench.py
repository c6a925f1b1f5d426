#!/usr/bin/env python
"""bench.py -- Mrays/s of the per-pixel render path (primary + shadow + reflection rays).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload cfg5-box|...]

One "step" = one frame of the workload = one IRaytracer::Render().  At N > 1 the frame is sharded by
row-blocks over the ranks (one process per GPU, torchrun); every rank's kernels store their pixels straight
into rank 0's framebuffer over NVLink (CUDA IPC peer mapping), rank 0 reads the frame back.

  value     Mrays/s, device-resident: scene, camera and framebuffer already in HBM, kernels only,
            CUDA events on the launching stream, max over ranks
  e2e       the same frame through the reference-facing call (rt_render == Render()): launch + readback of the
            whole frame into pinned host memory inside the timed region.  Scene and camera are uploaded once
            before (as OpenCLRaytracer's constructor does, OpenCLRaytracer.cpp:70-73), so h2d bytes per step = 0.
  roofline  the timed kernel against the measured FP32 FFMA peak of this GPU (rt_measure_fp32_peak);
            MEASURED_PEAKS.json has no FP32 entry, HBM peak is taken from it for the framebuffer traffic.
  cpu_baseline  the reference's own kernel source compiled for the host (oracle/_ref, "reference") or the C
            restatement ("port") on all host cores, on a bounded pixel sample of the same frame.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "opencl-raytracer_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)

from pyrt import scenes, types  # noqa: E402

METRIC = "Mrays/s (primary+shadow+reflect)"

# flops of the arithmetic each counted unit executes (DESIGN.md "Work per unit")
FLOPS_FILTER = 14.0   # bounding-sphere filter, expanded form: 7 fma per (ray, object) pair = 3.5 FFMA2 (rt_wavefront.cu)
FLOPS_EXACT = 78.0    # reference arithmetic of one ray x object test as written
FLOPS_NODE = 52.0     # two inflated slab tests of one BVH node: 10 FADD + 6 FMUL + 18 FFMA in the SASS of wave_trace_bvh
                      # (profiles/r2_sass_wave_trace_bvh_node_step.txt); the 7 min/max and 4 compares are not counted
FLOPS_CELL = 6.0      # one step of the uniform-grid walk (three compares, an add, the termination test)
FLOPS_SHADE = 100.0   # Phong terms per (hit, light)
FLOPS_ALGO_PAIR = 54.5  # minimal affine ray x object test


def workload(name: str, args):
    """-> (objs, lights, width, height, max_bounces, mode, label)"""
    if name == "cfg5-box":
        o, l = scenes.synthetic_scene(100_000, 4096)
        return o, l, 7680, 4320, 4, 2, "synthetic 100k boxes + 4096 spheres, 7680x4320, reflection depth 4"
    if name == "cfg5-tri":  # extension: type 2 unit triangles (no reference primitive)
        o, l = scenes.synthetic_scene(100_000, 4096, second_type=types.TRIANGLE)
        return o, l, 7680, 4320, 4, 2, "synthetic 100k triangles (extension) + 4096 spheres, 7680x4320, reflection depth 4"
    if name == "cfg5-small":
        o, l = scenes.synthetic_scene(100_000, 4096)
        return o, l, 1920, 1080, 4, 2, "synthetic 100k boxes + 4096 spheres, 1920x1080, reflection depth 4"
    if name == "cfg5-tiny":
        o, l = scenes.synthetic_scene(100_000, 4096)
        return o, l, 640, 360, 4, 2, "synthetic 100k boxes + 4096 spheres, 640x360, reflection depth 4"
    table = {
        "cfg1": ("simpleSphere", 2560, 1440, 30, 2),
        "cfg2": ("multipleSpheres", 1920, 1080, 3, 2),
        "cfg3": ("roundedCube", 3840, 2160, 30, 2),
        "cfg3-shade": ("roundedCube", 3840, 2160, 30, 1),
        "cfg4": ("simpleScene", 3840, 2160, 5, 2),
        "cfg4-3l": ("simpleScene3L", 3840, 2160, 5, 2),
    }
    s, w, h, mb, mode = table[name]
    o, l = scenes.load_reference_scene(s)
    return o, l, w, h, mb, mode, "%s.txt %dx%d depth %d mode %d" % (s, w, h, mb, mode)


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons of one GPU during the timed region (NVML)."""

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index = index
        self.samples, self.reasons = [], set()
        self.max_mhz = None
        self._halt = threading.Event()
        self.ok = False
        try:
            import pynvml

            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:
            self.ok = False

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        names = {
            nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
            nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
            nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
            nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap",
        }
        while not self._halt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._halt.wait(0.05)

    def stop(self):
        self._halt.set()
        if self.is_alive():
            self.join(timeout=2)
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(self.samples)}


def ncu_traffic(name):
    """dram__bytes_read + dram__bytes_write of the FIRST launch in a committed ncu summary (profiles/<name>, written by
    tools/ncu_summary.py from one `ncu --set full` capture of this bench command); (None, None) when it is not there."""
    path = os.path.join(ROOT, "profiles", name)
    unit = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    try:
        tot, seen = 0.0, 0
        for line in open(path):
            f = line.split()
            if len(f) >= 3 and f[0] in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                tot += float(f[1]) * unit[f[2]]
                seen += 1
                if seen == 2:
                    return tot, "profiles/%s: first wave_trace_bvh launch captured there (the longest trace launch of the frame)" % name
    except Exception:
        pass
    return None, None


def sample_rays(width, height, zdir, n_side=64, with_index=False):
    """A bounded, frame-covering pixel sample: an n_side x n_side lattice of pixels of the full frame
    (with_index: also their indices jj * width + ii in the frame)."""
    jj = np.linspace(0, height - 1, n_side).astype(np.int64)
    ii = np.linspace(0, width - 1, n_side).astype(np.int64)
    rows = types.camera_rays(width, height, zdir, rows=jj)
    rays = rows.reshape(len(jj), width)[:, ii].reshape(-1).copy()
    if with_index:
        return rays, (jj[:, None] * width + ii[None, :]).reshape(-1)
    return rays


def cpu_arm(objs, lights, rays, max_bounces, mode, steps, warmup, n_ref_rays=None, strict=False):
    """The reference's CPU implementation of the path on the host cores: oracle/_ref when it was built,
    else the C restatement.  strict: the contraction-free build (-O2 -ffp-contract=off, the parity target) instead
    of the fastest one.  Returns (Mrays/s, ms_per_step, kind, cores, ray_count, build flags, pixels of the last step)."""
    import oracle

    cores = os.cpu_count() or 1
    if n_ref_rays is None:
        _, _, _, st = oracle.port_render(mode, max_bounces, objs, lights, rays, nthreads=cores)
        n_ref_rays = st.rays_primary + st.rays_shadow + st.rays_reflect
    # the reference kernels skip every type but sphere / box: a scene with the triangle extension is timed on the C restatement
    use_ref = oracle.have_ref() and not bool((np.asarray(objs["type"]) > 1).any())
    fast = (not strict) and use_ref and oracle.cpu_has_avx2_fma() and os.path.exists(os.path.join(ROOT, "oracle", "_ref", "libsar_ref_fast.so"))
    kind = "reference" if use_ref else "port"
    last = [None]

    def once():
        if use_ref:
            last[0] = oracle.ref_render(mode, max_bounces, objs, lights, rays, nthreads=cores, fast=fast)
        else:
            last[0] = oracle.port_render(mode, max_bounces, objs, lights, rays, nthreads=cores)[0]

    for _ in range(warmup):
        once()
    t0 = time.perf_counter()
    for _ in range(steps):
        once()
    dt = (time.perf_counter() - t0) / max(steps, 1)
    build = "-O3 -march=x86-64-v3" if fast else "-O2 -ffp-contract=off"
    return n_ref_rays / dt / 1e6, dt * 1e3, kind, cores, n_ref_rays, build, last[0]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="cfg5-box")
    ap.add_argument("--accel", type=int, default=-1, help="-1 auto, 0 brute (exact on every pair), 1 filter, 2 bvh")
    ap.add_argument("--block-rows", type=int, default=16)
    ap.add_argument("--cpu-sample", type=int, default=0, help="side of the pixel lattice of the CPU sample (0 = auto)")
    ap.add_argument("--leaf", type=int, default=0, help="BVH leaf size (0 = library default)")
    ap.add_argument("--pipeline", type=int, default=-1, help="-1 auto, 0 megakernel, 1 wavefront")
    ap.add_argument("--builder", type=int, default=0, help="BVH builder: 0 host SAH, 1 device LBVH")
    ap.add_argument("--bands", type=int, default=0, help="readback bands of rt_render (0 = library default)")
    ap.add_argument("--lanes", type=int, default=0, help="1: one launch sequence per frame, 2: two half-frames on two streams (0 = library default)")
    ap.add_argument("--zero-copy", type=int, default=-1, help="rt_render stores into the pinned host frame directly: -1 library default, 0 off, 1 on")
    ap.add_argument("--grid-res", type=int, default=0, help="cells per side of the light grids (0 = library default)")
    ap.add_argument("--grid3", type=int, default=-1, help="uniform 3-D grid for rays without an apex: cells per 100 objects (0 = BVH, -1 = library default)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-brute-leg", action="store_true")
    ap.add_argument("--no-configs", action="store_true", help="skip the per-frame numbers of BASELINE.json's configs 1-4")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    objs, lights, W, H, max_bounces, mode, label = workload(args.workload, args)
    zdir = types.camera_zdir(H)
    config = {"workload": label, "objects": int(len(objs)), "lights": int(len(lights)), "width": W, "height": H,
              "max_bounces": max_bounces, "mode": ["hittest", "shade", "shade_and_reflect"][mode]}

    # ------------------------------------------------------------------ reference arm (CPU)
    if args.impl == "reference":
        if rank != 0:
            return
        side = args.cpu_sample or (48 if len(objs) > 10000 else 512)
        rays = sample_rays(W, H, zdir, side)
        v, ms, kind, cores, nrays, build, _ = cpu_arm(objs, lights, rays, max_bounces, mode, args.steps, args.warmup)
        sample = "%dx%d pixel lattice of the frame (%d rays/step), %s, %d threads" % (side, side, nrays, build, cores)
        print(json.dumps({
            "impl": "reference", "metric": METRIC, "value": v, "unit": "Mrays/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": config,
            "cpu_baseline": {"value": v, "unit": "Mrays/s", "cores": cores, "kind": kind, "sample": sample},
            "e2e": {"value": v, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        }))
        return

    # ------------------------------------------------------------------ our arm (CUDA)
    # stdout carries exactly one JSON line: libraries that print there (NCCL's version banner) go to stderr instead
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)
    import torch
    import torch.distributed as dist

    from pyrt import lib

    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    lib.load()
    ctx = lib.Context(local_rank)
    sampler = ClockSampler(local_rank)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    if args.leaf:
        ctx.set_option(lib.OPT_BVH_LEAF, args.leaf)
    if args.pipeline >= 0:
        ctx.set_option(lib.OPT_PIPELINE, args.pipeline)
    if args.bands:
        ctx.set_option(lib.OPT_READBACK_BANDS, args.bands)
    if args.lanes:
        ctx.set_option(lib.OPT_LANES, args.lanes)
    if args.builder:
        ctx.set_option(lib.OPT_BVH_BUILDER, args.builder)
    if args.zero_copy >= 0:
        ctx.set_option(lib.OPT_ZERO_COPY, args.zero_copy)
    if args.grid_res:
        ctx.set_option(lib.OPT_GRID_RES, args.grid_res)
    if args.grid3 >= 0:
        ctx.set_option(lib.OPT_GRID3_DENSITY, args.grid3)
    t_up0 = time.perf_counter()
    ctx.upload_scene(objs, lights)
    ctx.set_camera(W, H, zdir)
    torch.cuda.synchronize()
    upload_ms = (time.perf_counter() - t_up0) * 1e3
    if world > 1:
        ctx.set_shard(args.block_rows, rank, world)

    # framebuffer of the gather root, mapped into every rank (CUDA IPC): kernels store straight into it
    target = 0
    if world > 1:
        handle = [ctx.ipc_export_framebuffer() if rank == 0 else None]
        dist.broadcast_object_list(handle, src=0)
        if rank != 0:
            target = ctx.ipc_open(handle[0])

    # ray counts of the frame (one counted, untimed render)
    ctx.set_option(lib.OPT_COUNT_RAYS, 1)
    ctx.render_device(max_bounces, mode, args.accel, target)
    st = ctx.stats()
    counts = torch.tensor([st.rays_primary, st.rays_shadow, st.rays_reflect, st.traced_shadow, st.traced_reflect,
                           st.pair_tests, st.exact_tests, st.node_visits, st.hit_pixels, st.pixels], dtype=torch.int64, device="cuda")
    if world > 1:
        dist.all_reduce(counts)
    (r_prim, r_shadow, r_refl, t_shadow, t_refl, pair_tests, exact_tests, node_visits, hit_pixels, pixels) = [int(x) for x in counts.tolist()]
    rays_traced = r_prim + t_shadow + t_refl
    rays_reference = r_prim + r_shadow + r_refl
    accel_used = st.accel_used
    pipeline_used = st.pipeline_used
    grid_walk = bool(st.grid_walk_used)
    launches_per_step = st.kernel_launches
    ctx.set_option(lib.OPT_COUNT_RAYS, 0)

    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")  # > 126 MB L2

    last_rank_ms = []

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            fn()
        barrier()
        sampler_local = ClockSampler(local_rank)
        sampler_local.start()
        dev_ms = 0.0
        t0 = time.perf_counter()
        for _ in range(steps):
            flush.zero_()
            barrier()
            dev_ms += fn()
        barrier()
        wall_ms = (time.perf_counter() - t0) * 1e3
        clocks = sampler_local.stop()
        t = torch.tensor([dev_ms], dtype=torch.float64, device="cuda")
        if world > 1:
            every = [torch.zeros_like(t) for _ in range(world)]
            dist.all_gather(every, t)
            last_rank_ms[:] = [float(x.item()) / steps for x in every]
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()) / steps, wall_ms / steps, clocks

    def step_device():
        ctx.render_device(max_bounces, mode, args.accel, target)
        return ctx.stats().kernel_ms

    ms_dev, _, clocks = timed(step_device, args.steps, args.warmup)
    rank_ms = list(last_rank_ms)
    value = rays_traced / (ms_dev * 1e-3) / 1e6

    # e2e: Render() semantics -- launch + readback of the whole frame to pinned host memory on rank 0
    host = lib.PinnedBuffer((W * H, 4)) if rank == 0 else None

    def step_e2e():
        t0 = time.perf_counter()
        ctx.render(max_bounces, mode, args.accel, out=host.array)
        return (time.perf_counter() - t0) * 1e3

    def step_e2e_gather():  # N > 1, variant A: peer stores into rank 0's HBM over NVLink, rank 0 reads the frame back
        t0 = time.perf_counter()
        ctx.render_device(max_bounces, mode, args.accel, target)
        barrier()  # every shard has landed in rank 0's HBM
        if rank == 0:
            ctx.readback(host.array)
        return (time.perf_counter() - t0) * 1e3

    e2e_gather = None
    if world == 1:
        ms_e2e, _, _ = timed(step_e2e, args.steps, args.warmup)
    else:
        ms_gather, _, _ = timed(step_e2e_gather, args.steps, args.warmup)
        e2e_gather = {"value": rays_traced / (ms_gather * 1e-3) / 1e6, "ms_per_step": ms_gather,
                      "path": "kernels store into rank 0's framebuffer over NVLink (CUDA IPC), rank 0 copies the frame to the host"}
        # variant B (the headline at N > 1): one host frame in shared memory, page-locked in every rank; each GPU renders
        # into its own framebuffer and copies its own row-blocks to the host over its own PCIe link
        name = [None]
        shared = None
        if rank == 0:
            shared = lib.SharedFrame(W * H)
            name[0] = shared.name
        dist.broadcast_object_list(name, src=0)
        if rank != 0:
            shared = lib.SharedFrame(W * H, name=name[0])

        def step_e2e_direct():
            t0 = time.perf_counter()
            ctx.render(max_bounces, mode, args.accel, out=shared.array)
            barrier()  # every rank's rows are in the host frame
            return (time.perf_counter() - t0) * 1e3

        ms_e2e, _, _ = timed(step_e2e_direct, args.steps, args.warmup)
        if rank == 0:  # the two paths must produce the same frame
            assert np.array_equal(shared.array, host.array), "direct per-GPU readback differs from the NVLink gather"
        barrier()
        shared.close()
    e2e_value = rays_traced / (ms_e2e * 1e-3) / 1e6
    e2e_stats = ctx.stats()

    # cold start, once: what `OpenCLRaytracer tracer(...); tracer.Render();` costs (OpenCLRaytracer.cpp:13-105) -- context,
    # scene repack + BVH build on the host, H2D of the scene, first frame, readback
    cold_ms = None
    if rank == 0 and world == 1:
        t0 = time.perf_counter()
        with lib.Context(local_rank) as c4:
            c4.upload_scene(objs, lights)
            c4.set_camera(W, H, zdir)
            c4.render(max_bounces, mode, args.accel, out=host.array)
        cold_ms = (time.perf_counter() - t0) * 1e3

    # ------------------------------------------------------------------ BASELINE.json's configs 1-4 (the reference's own scenes)
    # Parity-test cases, not the headline -- but their numbers belong in front of the driver too: kernel and end-to-end time
    # per frame, the framebuffer store rate against the HBM peak (these frames are store / PCIe bound), and the unchanged
    # reference .cl on this same GPU where the box has an OpenCL platform.
    small = None
    if rank == 0 and world == 1 and args.workload == "cfg5-box" and not args.no_configs:
        small = []
        try:
            import oracle
            ocl_ok, _ = oracle.ocl_probe()
        except Exception:
            ocl_ok = False
        for name in ("cfg1", "cfg2", "cfg3", "cfg3-shade", "cfg4", "cfg4-3l"):
            o2, l2, w2, h2, mb2, mode2, label2 = workload(name, args)
            z2 = types.camera_zdir(h2)
            buf = lib.PinnedBuffer((w2 * h2, 4))
            with lib.Context(local_rank) as c7:
                c7.upload_scene(o2, l2)
                c7.set_camera(w2, h2, z2)
                c7.set_option(lib.OPT_COUNT_RAYS, 1)
                c7.render_device(mb2, mode2, -1)
                s7 = c7.stats()
                c7.set_option(lib.OPT_COUNT_RAYS, 0)
                ks, es = [], []
                for i in range(8):
                    flush.zero_()
                    torch.cuda.synchronize()
                    c7.render_device(mb2, mode2, -1)
                    ks.append(c7.stats().kernel_ms)
                for i in range(8):
                    flush.zero_()
                    torch.cuda.synchronize()
                    t0 = time.perf_counter()
                    c7.render(mb2, mode2, -1, out=buf.array)
                    es.append((time.perf_counter() - t0) * 1e3)
                k_ms, e_ms = float(np.median(ks[3:])), float(np.median(es[3:]))
                entry = {"config": name, "workload": label2, "rays_traced": s7.rays_traced, "rays_reference": s7.rays_reference,
                         "kernel_ms": k_ms, "e2e_ms": e_ms, "mrays_per_s_kernel": s7.rays_traced / (k_ms * 1e-3) / 1e6,
                         "mrays_per_s_e2e": s7.rays_traced / (e_ms * 1e-3) / 1e6,
                         "framebuffer_store_gbs": w2 * h2 * 16 / (k_ms * 1e-3) / 1e9,
                         "accel": ["brute", "filter", "bvh", "grids"][s7.accel_used], "pipeline": ["megakernel", "wavefront"][s7.pipeline_used]}
            if ocl_ok:
                try:
                    rays2 = types.camera_rays(w2, h2, z2)
                    _, tm = oracle.ocl_render(mode2 if mode2 in (1, 2) else 2, mb2, o2, l2, rays2, frames=3)
                    entry["reference_opencl_gpu"] = {"kernel_ms": tm["kernel_ms"], "render_ms": tm["render_ms"],
                                                     "kernel_speedup": tm["kernel_ms"] / k_ms, "render_speedup": tm["render_ms"] / e_ms}
                    del rays2
                except Exception as e:
                    entry["reference_opencl_gpu"] = {"unavailable": "error: %s" % e}
            buf.close()
            small.append(entry)

    # dynamic scene: what a re-upload of the whole scene costs when everything per object runs on the device
    # (RT_OPT_BVH_BUILDER = 1: repack + bounds + Morton tree + grids), and the frame rendered from that tree
    dynamic = None
    if rank == 0 and world == 1 and len(objs) >= 2:
        with lib.Context(local_rank) as c6:
            c6.set_option(lib.OPT_BVH_BUILDER, 1)
            c6.set_camera(W, H, zdir)
            ups, firsts = [], []
            for _ in range(4):
                t0 = time.perf_counter()
                c6.upload_scene(objs, lights)
                t1 = time.perf_counter()
                c6.render_device(max_bounces, mode, args.accel)  # rebuilds the grids, then the frame
                t2 = time.perf_counter()
                ups.append((t1 - t0) * 1e3)
                firsts.append((t2 - t1) * 1e3)
            c6.render_device(max_bounces, mode, args.accel)
            dynamic = {"upload_ms": float(np.median(ups[1:])), "first_frame_incl_grid_build_ms": float(np.median(firsts[1:])),
                       "steady_frame_ms": c6.stats().kernel_ms, "builder": "device: rt_bounds.cu + rt_lbvh.cu + rt_pgrid.cu",
                       "host_builder_upload_ms": upload_ms}

    # ------------------------------------------------------------------ roofline of the timed kernel
    peak_tf, peak_mhz = ctx.measure_fp32_peak()
    peak_scalar, peak_packed = ctx.fp32_peak_detail()
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    kernel_names = ({0: "wave_trace_brute", 1: "wave_trace_filter", 2: "wave_trace_bvh", 3: "wave_trace_bvh"} if pipeline_used == 1 else
                    {0: "render_mega<brute>", 1: "render_mega<filter>", 2: "render_mega<bvh>"})
    shades = t_shadow  # one Phong evaluation per traced shadow ray
    if accel_used == 3:  # perspective grids (filter tests per list entry) + uniform-grid walk or BVH for the rays without a common apex
        flops = node_visits * (FLOPS_CELL if grid_walk else FLOPS_NODE) + pair_tests * FLOPS_FILTER + exact_tests * FLOPS_EXACT + shades * FLOPS_SHADE
    elif accel_used == 2:
        flops = node_visits * FLOPS_NODE + exact_tests * (FLOPS_EXACT + FLOPS_FILTER) + shades * FLOPS_SHADE
    elif accel_used == 1:
        flops = pair_tests * FLOPS_FILTER + exact_tests * FLOPS_EXACT + shades * FLOPS_SHADE
    else:
        flops = pair_tests * FLOPS_EXACT + shades * FLOPS_SHADE
    achieved_tf = flops / world / (ms_dev * 1e-3) / 1e12
    # dram bytes of the dominant kernel's longest launch, from the committed ncu capture of this command (N = 1 only: a
    # shard's launches move other bytes)
    traffic, traffic_src = (None, None)
    if world == 1 and pipeline_used == 1 and args.workload == "cfg5-box":
        traffic, traffic_src = ncu_traffic({2: "r1_g_wave_trace_bvh_cfg5_8k.txt",
                                            3: "r2_wave_trace_dda_v2_cfg5_8k.txt" if grid_walk else "r2_wave_trace_bvh_cfg5_8k.txt"}.get(accel_used, "none"))
    roofline = {
        "kernel": "wave_trace_dda" if (accel_used == 3 and grid_walk and pipeline_used == 1) else kernel_names[accel_used], "bound": "fp32", "achieved": achieved_tf, "peak": peak_tf, "unit": "TFLOP/s",
        "frac": achieved_tf / peak_tf if peak_tf else None, "traffic": traffic, "traffic_source": traffic_src,
        "peak_source": "rt_measure_fp32_peak (register-only loops, this GPU, this run: scalar FFMA %.1f, packed FFMA2 %.1f TFLOP/s; "
                       "the larger is the peak and implies %.0f MHz)" % (peak_scalar, peak_packed, peak_mhz),
        "flops_per_launch": flops / world,
        # SURVEY.md 8(d)'s algorithmic unit beside the executed flops: 54.5 flop per exact ray x object test + 100 per shade --
        # what the path HAS to compute, which cannot be inflated by traversal or filter work
        "algorithmic": {"flops_per_launch": (exact_tests * FLOPS_ALGO_PAIR + shades * FLOPS_SHADE) / world,
                        "frac": (exact_tests * FLOPS_ALGO_PAIR + shades * FLOPS_SHADE) / world / (ms_dev * 1e-3) / 1e12 / peak_tf if peak_tf else None},
        "hbm": {"framebuffer_bytes_per_launch": pixels * 16 / world, "achieved_gbs": pixels * 16 / world / (ms_dev * 1e-3) / 1e9,
                "peak_gbs": peaks.get("hbm_gbs")},
        "note": ("the frame's dominant kernel is the trace of the reflection rays -- a uniform-grid walk (wave_trace_dda) where the scene "
                 "allows it, else the BVH (wave_trace_bvh); the rays through a common apex -- camera, point-light shadow -- look up a "
                 "perspective grid instead (wave_trace_grid).  These searches are bound by latency and issue slots, not by FP32 (ncu "
                 "summaries under profiles/r2_*); the FP32 fraction is reported for completeness, see roofline_brute for the FP32-bound "
                 "kernel") if accel_used >= 2 else "",
    }

    # brute-force leg: the reference's O(rays x objects) search (filter + exact) on a bounded band of the same frame
    roofline_brute = None
    if not args.no_brute_leg and rank == 0 and len(objs) > 48:
        rows = max(8, min(H, int(2.0e13 / (FLOPS_FILTER * len(objs) * W * 6)) // 8 * 8))
        r0 = (H // 2 - rows // 2) // 8 * 8
        band = types.camera_rays(W, H, zdir, rows=np.arange(r0, r0 + rows))
        with lib.Context(local_rank) as c2:
            c2.upload_scene(objs, lights)
            c2.set_rays(band)
            c2.set_option(lib.OPT_COUNT_RAYS, 1)
            c2.render_device(max_bounces, mode, 1)
            s2 = c2.stats()
            c2.set_option(lib.OPT_COUNT_RAYS, 0)
            best = None
            for _ in range(3):
                c2.render_device(max_bounces, mode, 1)
                k = c2.stats().kernel_ms
                best = k if best is None else min(best, k)
            fl = s2.pair_tests * FLOPS_FILTER + s2.exact_tests * FLOPS_EXACT + s2.traced_shadow * FLOPS_SHADE
            tf = fl / (best * 1e-3) / 1e12
            roofline_brute = {
                "kernel": "wave_trace_filter" if s2.pipeline_used == 1 else "render_mega<filter>", "bound": "fp32", "sample": "%d rows x %d px band at row %d" % (rows, W, r0),
                "pair_tests": s2.pair_tests, "ms": best, "achieved": tf, "peak": peak_tf, "unit": "TFLOP/s",
                "frac": tf / peak_tf if peak_tf else None,
                "reference_equivalent_tflops": s2.pair_tests * FLOPS_ALGO_PAIR / (best * 1e-3) / 1e12,
                "mrays_per_s": s2.rays_traced / (best * 1e-3) / 1e6,
            }

    # ------------------------------------------------------------------ CPU baseline (rank 0, N = 1 only) + parity of THIS run
    # The timed CPU leg is the reference's own kernel source in its contraction-free build -- the parity target -- so its
    # pixels are not thrown away: they are compared with the same pixels of the frame the GPU just produced through
    # rt_render (colours 1e-4 * max(1, |ref|), PPM <= 1 LSB), and the C restatement adds hit ids and t (bit-exact).  A
    # mismatch fails the run: a fast frame that differs from the reference's is not a result.
    cpu = None
    parity = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        import oracle

        side = args.cpu_sample or (40 if len(objs) > 10000 else 384)
        rays, pix = sample_rays(W, H, zdir, side, with_index=True)
        with lib.Context(local_rank) as c3:  # reference ray count of the sample (GPU count == oracle count, tests)
            c3.upload_scene(objs, lights)
            c3.set_rays(rays)
            c3.set_option(lib.OPT_COUNT_RAYS, 1)
            c3.set_option(lib.OPT_SKIP_DEAD_RAYS, 0)
            c3.render_device(max_bounces, mode, -1)
            nref = c3.stats().rays_reference
        v, ms, kind, cores, nrays, build, ref_px = cpu_arm(objs, lights, rays, max_bounces, mode, 1, 0, n_ref_rays=nref, strict=True)
        cpu = {"value": v, "unit": "Mrays/s", "cores": cores, "kind": kind,
               "sample": "%dx%d pixel lattice of the frame, %d reference rays, %.1f s, %s" % (side, side, nrays, ms / 1e3, build)}
        got = host.array[pix]
        d = np.abs(got[:, :3].astype(np.float64) - ref_px[:, :3].astype(np.float64))
        lim = 1e-4 * np.maximum(1.0, np.abs(ref_px[:, :3].astype(np.float64)))
        ppm = int(np.abs(types.quantize_ppm(got).astype(np.int64) - types.quantize_ppm(ref_px).astype(np.int64)).max(initial=0))
        _, hid, ht, _ = oracle.port_render(mode, max_bounces, objs, lights, rays, nthreads=cores)
        gid, gt = ctx.render_aov(args.accel)
        parity = {"pixels": int(len(pix)), "against": "oracle/_ref (%s) colours, oracle port hit ids / t" % build if kind == "reference" else "oracle port",
                  "colours_ok": bool((d <= lim).all()), "max_abs": float(d.max(initial=0.0)),
                  "max_rel": float((d / np.maximum(1.0, np.abs(ref_px[:, :3]))).max(initial=0.0)), "ppm_max_lsb": ppm,
                  "ids_exact": bool(np.array_equal(gid[pix], hid)), "t_exact": bool(np.array_equal(gt[pix], ht)),
                  "hit_pixels": int((hid >= 0).sum())}
        del gid, gt

    # ------------------------------------------------------------------ the reference's own .cl on this GPU (rank 0, N = 1)
    ref_ocl = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            import oracle

            ok, info = oracle.ocl_probe()
            if bool((np.asarray(objs["type"]) > 1).any()):
                ref_ocl = {"unavailable": "the reference kernels have no triangle primitive (type 2 is this project's extension)"}
            elif not ok:
                ref_ocl = {"unavailable": info}
            else:
                # brute force over every object: a band of rows sized for about a second of kernel time
                rows = H if len(objs) <= 48 else max(1, min(H, int(4.0e10 / (len(objs) * W * 7))))
                r0 = H // 2 - rows // 2
                band = types.camera_rays(W, H, zdir, rows=np.arange(r0, r0 + rows))
                with lib.Context(local_rank) as c5:  # reference ray count of the band
                    c5.upload_scene(objs, lights)
                    c5.set_rays(band)
                    c5.set_option(lib.OPT_COUNT_RAYS, 1)
                    c5.set_option(lib.OPT_SKIP_DEAD_RAYS, 0)
                    c5.render_device(max_bounces, mode, -1)
                    nref = c5.stats().rays_reference
                _, tm = oracle.ocl_render(mode if mode in (1, 2) else 2, max_bounces, objs, lights, band, frames=2)
                ref_ocl = {"platform": info, "sample": "%d rows x %d px band at row %d" % (rows, W, r0), "reference_rays": nref,
                           "kernel_ms": tm["kernel_ms"], "render_ms": tm["render_ms"], "jit_build_ms": tm["build_ms"],
                           "mrays_per_s_kernel": nref / (tm["kernel_ms"] * 1e-3) / 1e6,
                           "mrays_per_s_render": nref / (tm["render_ms"] * 1e-3) / 1e6,
                           "note": "shade_and_reflect_kernel.cl unchanged, JIT-built by the box's OpenCL platform, local size 32 "
                                   "(OpenCLRaytracer.cpp:80-105); counts every ray the reference traces"}
        except Exception as e:  # baseline infrastructure must never take the bench line down
            ref_ocl = {"unavailable": "error: %s" % e}

    # ------------------------------------------------------------------ N > 1: the one-process multi-GPU paths, where the driver sees them
    # (a) the C++ drop-in: build/rt_cli --gpus N (CudaRaytracer::UseDevices: one context + host thread per device) must write
    # the very bytes the 1-GPU run writes; (b) two contexts of ONE process on two devices run the shared-memory-hungry filter
    # kernel of the wavefront pipeline (function attributes are per device: round 1 opted in on the first device only).
    one_process = None
    if rank == 0 and world > 1:
        import subprocess
        import tempfile

        one_process = {}
        cli = os.path.join(ROOT, "build", "rt_cli")
        scene = os.path.join(ROOT, "baseline", "_ref", "roundedCube.txt")
        if os.path.exists(cli) and os.path.exists(scene):
            outs = []
            with tempfile.TemporaryDirectory() as td:
                for g in (1, world):
                    o = os.path.join(td, "r%d.ppm" % g)
                    r = subprocess.run([cli, "--scene", scene, "--width", "640", "--height", "356", "--bounces", "30", "--gpus", str(g),
                                        "--fast-ppm", "--out", o], capture_output=True, text=True, timeout=300)
                    outs.append((r.returncode, open(o, "rb").read() if os.path.exists(o) else b"", r.stdout))
            one_process["rt_cli_gpus"] = world
            one_process["rt_cli_byte_identical"] = bool(outs[0][0] == 0 and outs[1][0] == 0 and outs[0][1] == outs[1][1] and len(outs[0][1]) > 0
                                                        and ("Frame sharded over %d GPU(s)." % world) in outs[1][2])
        else:
            one_process["rt_cli_byte_identical"] = None
        so, sl = scenes.synthetic_scene(3000, 300)
        frames = []
        for dev in (0, 1):
            with lib.Context(dev) as cx:
                cx.set_option(lib.OPT_PIPELINE, 1)
                cx.upload_scene(so, sl)
                cx.set_camera(320, 180, types.camera_zdir(180))
                frames.append(cx.render(4, 2, 1).copy())
        one_process["two_contexts_filter_kernel_equal"] = bool(np.array_equal(frames[0], frames[1]))
        if one_process["rt_cli_byte_identical"] is False or not one_process["two_contexts_filter_kernel_equal"]:
            sys.stderr.write("ONE-PROCESS MULTI-GPU CHECK FAILED: %s\n" % json.dumps(one_process))

    if rank == 0:
        out = {
            "metric": METRIC, "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_dev, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": dict(config, accel=["brute", "filter", "bvh", "perspective grids + " + ("uniform-grid walk" if grid_walk else "bvh")][accel_used], pipeline=["megakernel", "wavefront"][pipeline_used],
                           sharding="row-blocks of %d rows round-robin over %d ranks, peer stores into rank 0 (CUDA IPC)" % (args.block_rows, world)
                           if world > 1 else "single GPU",
                           l2="framebuffer %d MB > L2; 256 MB memset between steps flushes L2 (scene is re-read from HBM each step)" % (W * H * 16 >> 20),
                           scene_upload_ms=upload_ms, bvh_builder=["host SAH", "device LBVH"][args.builder]),
            "rays": {"traced": rays_traced, "reference": rays_reference, "primary": r_prim, "shadow_traced": t_shadow,
                     "shadow_reference": r_shadow, "reflect_traced": t_refl, "reflect_reference": r_refl, "hit_pixels": hit_pixels,
                     "pair_tests": pair_tests, "exact_tests": exact_tests, "node_visits": node_visits,
                     "mrays_reference_equivalent": rays_reference / (ms_dev * 1e-3) / 1e6},
            "e2e": {"value": e2e_value, "unit": "Mrays/s", "ms_per_step": ms_e2e, "h2d_bytes_per_step": 0,
                    "d2h_bytes_per_step": W * H * 16,
                    "device_ms": {"kernels": e2e_stats.kernel_ms, "kernels_plus_copy": e2e_stats.total_ms},
                    "nvlink_gather_variant": e2e_gather,
                    "path": "rt_render into the caller's pinned frame" if world == 1 else
                            "every rank rt_render()s its row-blocks into one shared, page-locked host frame (own PCIe link per GPU)",
                    "ctor_plus_first_render_ms": cold_ms, "scene_h2d_bytes_once": int(len(objs)) * 276 + int(len(lights)) * 64,
                    "note": "rt_render == IRaytracer::Render(): launch + blocking readback (the frame is cut into launch sequences of shrinking size "
                            "on three streams, a finished one is copied under the kernels of the others); scene/camera resident as in the reference ctor, whose "
                            "one-off cost is ctor_plus_first_render_ms"},
            "gpu_launches": args.steps * launches_per_step,
            "rank_ms": rank_ms or None,
            "clocks": clocks,
            "roofline": roofline,
        }
        if roofline_brute:
            out["roofline_brute"] = roofline_brute
        if cpu:
            out["cpu_baseline"] = cpu
        if parity:
            out["parity"] = parity
        if dynamic:
            out["dynamic_scene"] = dynamic
        if small:
            out["configs"] = small
        if one_process:
            out["one_process_multi_gpu"] = one_process
        if ref_ocl:
            out["reference_opencl_gpu"] = ref_ocl
        sys.stdout.flush()
        os.write(json_fd, (json.dumps(out) + "\n").encode())
    if parity and not (parity["colours_ok"] and parity["ids_exact"] and parity["t_exact"] and parity["ppm_max_lsb"] <= 1):
        sys.stderr.write("PARITY FAILED: %s\n" % json.dumps(parity))
        sys.exit(3)
    if world > 1:
        if target:
            ctx.ipc_close(target)
        dist.barrier()
        dist.destroy_process_group()
    ctx.close()


if __name__ == "__main__":
    main()
