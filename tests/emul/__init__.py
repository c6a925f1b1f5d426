"""TEST INFRASTRUCTURE -- ctypes loader of the host emulation of the CUDA path's device code
(tests/emul/rt_emul.cpp).  CPU-only test tier; never imported by the product."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_lib = None

COUNTER_FIELDS = ("ref_primary", "ref_shadow", "ref_reflect", "traced_shadow", "traced_reflect", "pair_tests",
                  "exact_tests", "node_visits", "hit_pixels")


def lib():
    global _lib
    if _lib is None:
        subprocess.check_call(["make", "-s", "-C", _HERE])
        _lib = C.CDLL(os.path.join(_HERE, "librt_emul.so"))
        _lib.emul_last_error.restype = C.c_char_p
        _lib.emul_render.restype = C.c_int
        _lib.emul_render.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint64, C.c_uint32,
                                     C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p,
                                     C.c_void_p, C.c_int]
        _lib.emul_render_wave.restype = C.c_int
        _lib.emul_render_wave.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint64, C.c_uint32,
                                          C.c_int, C.c_int, C.c_int, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p,
                                          C.c_uint32, C.c_uint32, C.c_float, C.c_uint32, C.c_uint32, C.c_uint32, C.c_float, C.c_void_p]
        _lib.emul_bvh_info.restype = C.c_int
        _lib.emul_bvh_info.argtypes = [C.c_void_p, C.c_uint32, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
    return _lib


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None and a.size else None


def render(objs, lights, rays, max_bounces=30, mode=2, accel=2, skip_dead=1, leaf=4, nthreads=0):
    objs = np.ascontiguousarray(objs)
    lights = np.ascontiguousarray(lights)
    rays = np.ascontiguousarray(rays)
    n = rays.shape[0]
    out = np.empty((n, 4), dtype=np.float32)
    hid = np.empty(n, dtype=np.int32)
    ht = np.empty(n, dtype=np.float32)
    ctr = np.zeros(len(COUNTER_FIELDS), dtype=np.uint64)
    omax = float(np.sqrt((rays["start"][:, :3].astype(np.float64) ** 2).sum(axis=1).max())) if n else 0.0
    rc = lib().emul_render(_ptr(objs), len(objs), _ptr(lights), len(lights), _ptr(rays), n, max_bounces, mode, accel,
                           skip_dead, leaf, omax, _ptr(out), _ptr(hid), _ptr(ht), _ptr(ctr), nthreads)
    if rc != 0:
        raise RuntimeError("emul_render: %d %s" % (rc, lib().emul_last_error().decode()))
    return out, hid, ht, dict(zip(COUNTER_FIELDS, (int(x) for x in ctr)))


def render_wave(objs, lights, rays, max_bounces=30, mode=2, accel=2, leaf=4, camera=None, grid_res=64, tile=(8, 4), grid3=0.0):
    """The wavefront state machine (rt_wave.cuh) on the host; returns (rgba, counters, generations).
    accel 3: perspective grids (rt_pgrid.cuh) -- camera = (width, height, zdir) of the pinhole frame `rays` is, or None
    for arbitrary rays (primary rays then use the BVH); grid_res cells per side of the light cube maps; grid3 > 0: the
    rays without a common apex walk a uniform 3-D grid with that many cells per object (rt_grid3.cuh) instead of the BVH
    (the number of cells visited is added to the counters as "cells")."""
    objs = np.ascontiguousarray(objs)
    lights = np.ascontiguousarray(lights)
    rays = np.ascontiguousarray(rays)
    n = rays.shape[0]
    out = np.empty((n, 4), dtype=np.float32)
    ctr = np.zeros(len(COUNTER_FIELDS), dtype=np.uint64)
    gens = C.c_uint32(0)
    cells = C.c_uint64(0)
    omax = float(np.sqrt((rays["start"][:, :3].astype(np.float64) ** 2).sum(axis=1).max())) if n else 0.0
    rc = lib().emul_render_wave(_ptr(objs), len(objs), _ptr(lights), len(lights), _ptr(rays), n, max_bounces, mode, accel,
                                leaf, omax, _ptr(out), _ptr(ctr), C.byref(gens), camera[0] if camera else 0, camera[1] if camera else 0,
                                float(camera[2]) if camera else 0.0, grid_res, tile[0], tile[1], float(grid3), C.byref(cells))
    if rc != 0:
        raise RuntimeError("emul_render_wave: %d %s" % (rc, lib().emul_last_error().decode()))
    cd = dict(zip(COUNTER_FIELDS, (int(x) for x in ctr)))
    cd["cells"] = cells.value
    return out, cd, gens.value
