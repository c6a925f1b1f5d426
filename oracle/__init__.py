"""TEST INFRASTRUCTURE -- ctypes loaders for the CPU oracles.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this package.  The product path (opencl-raytracer_b200/) never does.

  ref()   oracle/_ref/libsar_ref.so       the reference's own .cl kernels compiled as C++
          (parity build, -O2 -ffp-contract=off); ref(fast=True) is the -O3 AVX2 timing build
  port()  oracle/liboracle_rt.so          the plain-C restatement (oracle_rt.c) that also
          reports hit ids, t and ray counts, which the reference never outputs
  ocl_*() oracle/libocl_ref.so            the reference's unchanged .cl source on the box's OpenCL platform
          (baseline/_ref/*.cl staged by `make ocl`), when one enumerates
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_libs: dict = {}


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def have_ref() -> bool:
    return os.path.exists(os.path.join(_HERE, "_ref", "libsar_ref.so"))


def cpu_has_avx2_fma() -> bool:
    try:
        with open("/proc/cpuinfo") as f:
            flags = f.read()
        return " avx2" in flags and " fma" in flags
    except OSError:
        return False


def ref(fast: bool = False):
    name = "libsar_ref_fast.so" if fast else "libsar_ref.so"
    if name not in _libs:
        lib = C.CDLL(os.path.join(_HERE, "_ref", name))
        lib.sar_ref_render.restype = C.c_int
        lib.sar_ref_render.argtypes = [C.c_int, C.c_uint32, C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32,
                                       C.c_void_p, C.c_uint64, C.c_void_p, C.c_int]
        _libs[name] = lib
    return _libs[name]


def ref_render(kernel: int, max_bounces: int, objs: np.ndarray, lights: np.ndarray, rays: np.ndarray,
               nthreads: int = 0, fast: bool = False) -> np.ndarray:
    """Run the reference kernel (0 hittest, 1 shade, 2 shade_and_reflect) over `rays`; returns (N,4) float32."""
    objs = np.ascontiguousarray(objs)
    lights = np.ascontiguousarray(lights)
    rays = np.ascontiguousarray(rays)
    out = np.empty((rays.shape[0], 4), dtype=np.float32)
    rc = ref(fast).sar_ref_render(kernel, max_bounces, _ptr(objs), objs.shape[0], _ptr(lights), lights.shape[0],
                                  _ptr(rays), rays.shape[0], _ptr(out), nthreads)
    if rc != 0:
        raise RuntimeError("sar_ref_render failed: %d" % rc)
    return out


class PortStats(C.Structure):
    _fields_ = [("rays_primary", C.c_uint64), ("rays_shadow", C.c_uint64), ("rays_reflect", C.c_uint64),
                ("hit_pixels", C.c_uint64)]


def port():
    if "port" not in _libs:
        lib = C.CDLL(os.path.join(_HERE, "liboracle_rt.so"))
        lib.oracle_render.restype = C.c_int
        lib.oracle_render.argtypes = [C.c_int, C.c_uint32, C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32,
                                      C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p,
                                      C.POINTER(PortStats), C.c_int]
        lib.oracle_trace.restype = C.c_int
        lib.oracle_trace.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_int]
        _libs["port"] = lib
    return _libs["port"]


def port_render(mode: int, max_bounces: int, objs, lights, rays, nthreads: int = 0):
    """CPU restatement: returns (rgba (N,4) f32, hit_id (N,) i32, hit_t (N,) f32, PortStats)."""
    objs = np.ascontiguousarray(objs)
    lights = np.ascontiguousarray(lights)
    rays = np.ascontiguousarray(rays)
    n = rays.shape[0]
    out = np.empty((n, 4), dtype=np.float32)
    hid = np.empty(n, dtype=np.int32)
    ht = np.empty(n, dtype=np.float32)
    st = PortStats()
    rc = port().oracle_render(mode, max_bounces, _ptr(objs), objs.shape[0], _ptr(lights), lights.shape[0],
                              _ptr(rays), n, _ptr(out), _ptr(hid), _ptr(ht), C.byref(st), nthreads)
    if rc != 0:
        raise RuntimeError("oracle_render failed: %d" % rc)
    return out, hid, ht, st


def port_trace(objs, rays, nthreads: int = 0):
    """Nearest hit (SAR:72-177 tie rules) of arbitrary rays: (hit_id i32, hit_t f32)."""
    objs = np.ascontiguousarray(objs)
    rays = np.ascontiguousarray(rays)
    n = rays.shape[0]
    hid = np.empty(n, dtype=np.int32)
    ht = np.empty(n, dtype=np.float32)
    rc = port().oracle_trace(_ptr(objs), objs.shape[0], _ptr(rays), n, _ptr(hid), _ptr(ht), nthreads)
    if rc != 0:
        raise RuntimeError("oracle_trace failed: %d" % rc)
    return hid, ht


# ---------------------------------------------------------------- the reference .cl on the box's OpenCL platform
_ROOT = os.path.dirname(_HERE)
_CL_FILES = {1: ("shade_kernel.cl", "shade"), 2: ("shade_and_reflect_kernel.cl", "shade_and_reflect")}


def ocl():
    if "ocl" not in _libs:
        lib = C.CDLL(os.path.join(_HERE, "libocl_ref.so"))
        lib.ocl_ref_probe.restype = C.c_int
        lib.ocl_ref_probe.argtypes = [C.c_char_p, C.c_int]
        lib.ocl_ref_render.restype = C.c_int
        lib.ocl_ref_render.argtypes = [C.c_char_p, C.c_char_p, C.c_uint32, C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32,
                                       C.c_void_p, C.c_uint64, C.c_int, C.c_void_p, C.POINTER(C.c_double),
                                       C.POINTER(C.c_double), C.POINTER(C.c_double), C.c_char_p, C.c_int]
        _libs["ocl"] = lib
    return _libs["ocl"]


def ocl_probe():
    """-> (available, description): is there an OpenCL GPU platform, and the staged kernel sources, on this box?"""
    if not os.path.exists(os.path.join(_HERE, "libocl_ref.so")):
        return False, "oracle/libocl_ref.so not built"
    if not os.path.exists(os.path.join(_ROOT, "baseline", "_ref", "shade_and_reflect_kernel.cl")):
        return False, "baseline/_ref/*.cl not staged (make -C oracle ocl in the build container)"
    buf = C.create_string_buffer(512)
    rc = ocl().ocl_ref_probe(buf, 512)
    return rc == 0, buf.value.decode(errors="replace")


def ocl_render(kernel: int, max_bounces: int, objs, lights, rays, frames: int = 1):
    """OpenCLRaytracer ctor + `frames` x Render() with the reference's unchanged kernel source.
    -> (rgba (N,4) f32, {"kernel_ms", "render_ms", "build_ms"})"""
    fname, kname = _CL_FILES[kernel]
    objs = np.ascontiguousarray(objs)
    lights = np.ascontiguousarray(lights)
    rays = np.ascontiguousarray(rays)
    out = np.empty((rays.shape[0], 4), dtype=np.float32)
    k, r, b = C.c_double(0), C.c_double(0), C.c_double(0)
    log = C.create_string_buffer(2048)
    rc = ocl().ocl_ref_render(os.path.join(_ROOT, "baseline", "_ref", fname).encode(), kname.encode(), max_bounces,
                              _ptr(objs), objs.shape[0], _ptr(lights), lights.shape[0], _ptr(rays), rays.shape[0], frames,
                              _ptr(out), C.byref(k), C.byref(r), C.byref(b), log, 2048)
    if rc != 0:
        raise RuntimeError("ocl_ref_render failed: %s" % log.value.decode(errors="replace"))
    return out, {"kernel_ms": k.value, "render_ms": r.value, "build_ms": b.value}
