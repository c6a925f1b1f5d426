"""ctypes binding of librt_b200.so (include/rt_b200.h) -- the Python face of the same C ABI the C++
`CudaRaytracer` uses.  There is no CPU fallback: if the CUDA library is missing or no sm_100 device is
present, construction raises.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from .types import LIGHT_DTYPE, OBJECT_DTYPE, RAY_DTYPE

_PKG = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
# RT_B200_LIB: an alternative build of the same library (kernel experiments); never a fallback
LIB_PATH = os.environ.get("RT_B200_LIB") or os.path.join(_PKG, "librt_b200.so")

OPT_PIPELINE, OPT_COUNT_RAYS, OPT_BVH_LEAF, OPT_SKIP_DEAD_RAYS, OPT_READBACK_BANDS, OPT_BVH_BUILDER, OPT_LANES, OPT_GRID_RES, OPT_ZERO_COPY, OPT_GRID3_DENSITY, OPT_DETECT_PINHOLE, OPT_TAIL_GEN = 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12
TRACE_STATIC_MAX_RAYS = 262144
ACCEL_AUTO, ACCEL_BRUTE, ACCEL_CULL, ACCEL_BVH, ACCEL_GRID = -1, 0, 1, 2, 3

# every symbol include/rt_b200.h declares (tests check the library exports all of them)
ABI_SYMBOLS = (
    "rt_last_error", "rt_abi_version", "rt_device_count", "rt_create", "rt_destroy", "rt_set_option", "rt_upload_scene", "rt_set_rays",
    "rt_set_camera", "rt_get_camera", "rt_set_shard", "rt_render", "rt_render_device", "rt_framebuffer", "rt_render_aov", "rt_trace", "rt_trace_static",
    "rt_get_stats", "rt_quantize_rgb8", "rt_host_alloc", "rt_host_free", "rt_host_register", "rt_host_unregister", "rt_measure_fp32_peak", "rt_fp32_peak_detail",
    "rt_ipc_export_framebuffer", "rt_ipc_open", "rt_ipc_close", "rt_readback", "rt_host_build_bvh", "rt_plan_bands", "rt_plan_sequences",
)


class Stats(C.Structure):
    _fields_ = [
        ("rays_primary", C.c_uint64), ("rays_shadow", C.c_uint64), ("rays_reflect", C.c_uint64),
        ("traced_shadow", C.c_uint64), ("traced_reflect", C.c_uint64),
        ("pair_tests", C.c_uint64), ("exact_tests", C.c_uint64), ("node_visits", C.c_uint64),
        ("pixels", C.c_uint64), ("hit_pixels", C.c_uint64),
        ("kernel_ms", C.c_float), ("total_ms", C.c_float),
        ("kernel_launches", C.c_uint32), ("accel_used", C.c_uint32), ("pipeline_used", C.c_uint32),
        ("grid_walk_used", C.c_uint32),
    ]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_ }

    @property
    def rays_traced(self):
        return self.rays_primary + self.traced_shadow + self.traced_reflect

    @property
    def rays_reference(self):
        return self.rays_primary + self.rays_shadow + self.rays_reflect


class BvhInfo(C.Structure):
    _fields_ = [("n_nodes", C.c_uint32), ("n_leaf_objs", C.c_uint32), ("depth", C.c_uint32), ("root", C.c_int32),
                ("omax", C.c_float), ("scene_radius", C.c_float), ("build_ms", C.c_double)]


# rt::BvhNode (csrc/rt_layout.h), 64 bytes
BVH_NODE_DTYPE = np.dtype([("ctr0", "<f4", 3), ("ext0", "<f4", 3), ("ctr1", "<f4", 3), ("ext1", "<f4", 3),
                           ("c0", "<i4"), ("c1", "<i4"), ("f0", "<f4"), ("f1", "<f4")])

_lib = None


def load():
    """Load librt_b200.so; raises OSError if it has not been built (python __graft_entry__.py build)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise OSError("%s not found: the CUDA library has not been built and there is no CPU fallback" % LIB_PATH)
        lib = C.CDLL(LIB_PATH)
        lib.rt_last_error.restype = C.c_char_p
        lib.rt_abi_version.restype = C.c_int
        vp, u32, u64, i32, i64 = C.c_void_p, C.c_uint32, C.c_uint64, C.c_int, C.c_int64
        sigs = {
            "rt_device_count": [C.POINTER(i32)],
            "rt_create": [C.POINTER(vp), i32],
            "rt_destroy": [vp],
            "rt_set_option": [vp, i32, i64],
            "rt_upload_scene": [vp, vp, u32, vp, u32],
            "rt_set_rays": [vp, vp, u64],
            "rt_set_camera": [vp, u32, u32, C.c_float],
            "rt_get_camera": [vp, C.POINTER(u32), C.POINTER(u32), C.POINTER(C.c_float)],
            "rt_set_shard": [vp, u32, u32, u32],
            "rt_render": [vp, u32, i32, i32, vp],
            "rt_render_device": [vp, u32, i32, i32, vp],
            "rt_framebuffer": [vp, C.POINTER(vp), C.POINTER(u64)],
            "rt_readback": [vp, vp],
            "rt_render_aov": [vp, i32, vp, vp],
            "rt_trace": [vp, vp, u64, i32, vp, vp],
            "rt_trace_static": [vp, vp, u64, i32, vp, vp],
            "rt_get_stats": [vp, C.POINTER(Stats)],
            "rt_quantize_rgb8": [vp, vp, u64, vp, C.POINTER(u64)],
            "rt_host_alloc": [C.POINTER(vp), u64],
            "rt_host_free": [vp],
            "rt_host_register": [vp, u64],
            "rt_host_unregister": [vp],
            "rt_measure_fp32_peak": [vp, C.POINTER(C.c_float), C.POINTER(C.c_float)],
            "rt_fp32_peak_detail": [vp, C.POINTER(C.c_float), C.POINTER(C.c_float)],
            "rt_ipc_export_framebuffer": [vp, vp],
            "rt_ipc_open": [vp, vp, C.POINTER(vp)],
            "rt_ipc_close": [vp, vp],
            "rt_host_build_bvh": [vp, u32, vp, u32, C.c_float, i32, i32, C.POINTER(BvhInfo), vp, u64, vp, u64],
            "rt_plan_bands": [u32, u32, u32, u32, i32, C.c_double, vp],
            "rt_plan_sequences": [u32, u32, u32, u32, u32, i32, C.c_double, C.c_double, vp, vp],
        }
        for name, args in sigs.items():
            fn = getattr(lib, name)
            fn.restype = C.c_int
            fn.argtypes = args
        _lib = lib
    return _lib


class RtError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("rt_b200 error %d: %s" % (code, msg))
        self.code = code


def _check(rc):
    if rc != 0:
        raise RtError(rc, load().rt_last_error().decode(errors="replace"))


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None and a.size else None


def host_build_bvh(objs: np.ndarray, lights: np.ndarray, omax_user: float = 0.0, leaf_size: int = 1, threads: int = 0):
    """The host scene builder of rt_upload_scene without a GPU (rt_host_build_bvh): (info, nodes, leaf_obj)."""
    from .types import LIGHT_DTYPE, OBJECT_DTYPE
    objs = np.ascontiguousarray(objs, dtype=OBJECT_DTYPE)
    lights = np.ascontiguousarray(lights, dtype=LIGHT_DTYPE)
    info = BvhInfo()
    lib = load()
    _check(lib.rt_host_build_bvh(_ptr(objs), len(objs), _ptr(lights), len(lights), omax_user, leaf_size, threads, C.byref(info), None, 0, None, 0))
    nodes = np.zeros(info.n_nodes, dtype=BVH_NODE_DTYPE)
    leaf = np.zeros(info.n_leaf_objs, dtype=np.uint32)
    _check(lib.rt_host_build_bvh(_ptr(objs), len(objs), _ptr(lights), len(lights), omax_user, leaf_size, threads, C.byref(info),
                                 _ptr(nodes), nodes.nbytes, _ptr(leaf), len(leaf)))
    return info, nodes, leaf


def plan_sequences(width: int, height: int, block_rows: int = 1, shard: int = 0, n_shards: int = 1, n_sequences: int = 0, q: float = 0.0,
                   kernel_copy_ratio: float = 3.0) -> np.ndarray:
    """Boundaries (rows, or owned row-blocks of a shard) of the launch sequences of a default rt_render (rt_plan_sequences)."""
    b = np.zeros(17, dtype=np.uint32)
    n = C.c_int(0)
    _check(load().rt_plan_sequences(width, height, block_rows, shard, n_shards, n_sequences, q, kernel_copy_ratio, _ptr(b), C.byref(n)))
    return b[: n.value + 1].copy()


def plan_bands(height: int, block_rows: int, shard: int, n_shards: int, n_bands: int, ratio: float) -> np.ndarray:
    """Band boundaries (in rows, or owned row-blocks of a shard) of a banded rt_render (rt_plan_bands)."""
    b = np.zeros(n_bands + 1, dtype=np.uint32)
    _check(load().rt_plan_bands(height, block_rows, shard, n_shards, n_bands, ratio, _ptr(b)))
    return b


class SharedFrame:
    """One host frame shared by every rank of a multi-GPU job: a POSIX shared-memory segment, page-locked in each
    process (rt_host_register), so that every GPU copies its own rows into it over its own PCIe link."""

    def __init__(self, n_pixels: int, name: str | None = None):
        from multiprocessing import shared_memory

        self.nbytes = int(n_pixels) * 16
        self.owner = name is None
        self.shm = shared_memory.SharedMemory(create=self.owner, size=self.nbytes) if self.owner else shared_memory.SharedMemory(name=name)
        self.name = self.shm.name
        if not self.owner:  # Python < 3.13 registers attached segments too; only the creating rank may unlink
            try:
                from multiprocessing import resource_tracker

                resource_tracker.unregister(self.shm._name, "shared_memory")
            except Exception:
                pass
        self.array = np.ndarray((int(n_pixels), 4), dtype=np.float32, buffer=self.shm.buf)
        self._addr = C.c_void_p(self.array.ctypes.data)
        _check(load().rt_host_register(self._addr, self.nbytes))

    def close(self):
        if self.shm is not None:
            load().rt_host_unregister(self._addr)
            self.array = None
            try:
                self.shm.close()
                if self.owner:
                    self.shm.unlink()
            except Exception:
                pass
            self.shm = None


class PinnedBuffer:
    """A pinned host array (rt_host_alloc) usable as a render target."""

    def __init__(self, shape, dtype=np.float32):
        self.shape = tuple(shape)
        self.dtype = np.dtype(dtype)
        nbytes = int(np.prod(self.shape)) * self.dtype.itemsize
        p = C.c_void_p()
        _check(load().rt_host_alloc(C.byref(p), nbytes))
        self._p = p
        buf = (C.c_char * max(nbytes, 1)).from_address(p.value)
        self.array = np.frombuffer(buf, dtype=self.dtype, count=int(np.prod(self.shape))).reshape(self.shape)

    def close(self):
        if self._p is not None:
            self.array = None
            load().rt_host_free(self._p)
            self._p = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Context:
    """One rt_ctx: a device, a scene, rays or a camera, and its framebuffer."""

    def __init__(self, device: int = -1):
        self._h = C.c_void_p()
        _check(load().rt_create(C.byref(self._h), device))
        self._n = 0
        self._keep = []

    def close(self):
        if self._h:
            load().rt_destroy(self._h)
            self._h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- configuration
    def set_option(self, option: int, value: int):
        _check(load().rt_set_option(self._h, option, int(value)))

    def upload_scene(self, objs: np.ndarray, lights: np.ndarray):
        objs = np.ascontiguousarray(objs, dtype=OBJECT_DTYPE)
        lights = np.ascontiguousarray(lights, dtype=LIGHT_DTYPE)
        _check(load().rt_upload_scene(self._h, _ptr(objs), len(objs), _ptr(lights), len(lights)))

    def set_rays(self, rays: np.ndarray):
        rays = np.ascontiguousarray(rays, dtype=RAY_DTYPE)
        _check(load().rt_set_rays(self._h, _ptr(rays), len(rays)))
        self._n = len(rays)

    def set_camera(self, width: int, height: int, zdir):
        _check(load().rt_set_camera(self._h, width, height, float(zdir)))
        self._n = width * height

    def set_shard(self, block_rows: int, shard: int, n_shards: int):
        _check(load().rt_set_shard(self._h, block_rows, shard, n_shards))

    # -- rendering
    def render(self, max_bounces: int = 30, mode: int = 2, accel: int = -1, out: np.ndarray | None = None) -> np.ndarray:
        if out is None:
            out = np.empty((self._n, 4), dtype=np.float32)
        assert out.dtype == np.float32 and out.size == self._n * 4 and out.flags.c_contiguous
        _check(load().rt_render(self._h, max_bounces, mode, accel, out.ctypes.data_as(C.c_void_p)))
        return out

    def render_device(self, max_bounces: int = 30, mode: int = 2, accel: int = -1, dev_ptr: int = 0):
        _check(load().rt_render_device(self._h, max_bounces, mode, accel, C.c_void_p(dev_ptr) if dev_ptr else None))

    def framebuffer(self):
        p, n = C.c_void_p(), C.c_uint64()
        _check(load().rt_framebuffer(self._h, C.byref(p), C.byref(n)))
        return p.value, n.value

    def readback(self, out: np.ndarray) -> np.ndarray:
        assert out.dtype == np.float32 and out.size == self._n * 4 and out.flags.c_contiguous
        _check(load().rt_readback(self._h, out.ctypes.data_as(C.c_void_p)))
        return out

    def render_aov(self, accel: int = -1):
        hid = np.empty(self._n, dtype=np.int32)
        ht = np.empty(self._n, dtype=np.float32)
        _check(load().rt_render_aov(self._h, accel, _ptr(hid), _ptr(ht)))
        return hid, ht

    def trace(self, rays: np.ndarray, accel: int = -1):
        rays = np.ascontiguousarray(rays, dtype=RAY_DTYPE)
        hid = np.empty(len(rays), dtype=np.int32)
        ht = np.empty(len(rays), dtype=np.float32)
        _check(load().rt_trace(self._h, _ptr(rays), len(rays), accel, _ptr(hid), _ptr(ht)))
        return hid, ht

    def trace_static(self, rays: np.ndarray, accel: int = -1):
        """rt_trace_static: the same search with no device allocation in the call (static scratch in the library,
        at most TRACE_STATIC_MAX_RAYS rays; calls that overlap on one GPU race)."""
        rays = np.ascontiguousarray(rays, dtype=RAY_DTYPE)
        hid = np.empty(len(rays), dtype=np.int32)
        ht = np.empty(len(rays), dtype=np.float32)
        _check(load().rt_trace_static(self._h, _ptr(rays), len(rays), accel, _ptr(hid), _ptr(ht)))
        return hid, ht

    def stats(self) -> Stats:
        st = Stats()
        _check(load().rt_get_stats(self._h, C.byref(st)))
        return st

    def quantize_rgb8(self, rgba: np.ndarray | None = None):
        """PPMExporter's quantisation on the device -> (uint8 [n, 3], number of negative channels clamped to 0).
        rgba None: the frame of the last render, straight from the device framebuffer."""
        neg = C.c_uint64(0)
        if rgba is None:
            out = np.empty((self._n, 3), dtype=np.uint8)
            _check(load().rt_quantize_rgb8(self._h, None, 0, _ptr(out), C.byref(neg)))
            return out, neg.value
        rgba = np.ascontiguousarray(rgba, dtype=np.float32).reshape(-1, 4)
        out = np.empty((len(rgba), 3), dtype=np.uint8)
        _check(load().rt_quantize_rgb8(self._h, _ptr(rgba), len(rgba), _ptr(out), C.byref(neg)))
        return out, neg.value

    def measure_fp32_peak(self):
        t, mhz = C.c_float(), C.c_float()
        _check(load().rt_measure_fp32_peak(self._h, C.byref(t), C.byref(mhz)))
        return t.value, mhz.value

    def fp32_peak_detail(self):
        """(scalar FFMA loop, packed FFMA2 loop) TFLOP/s of the last measure_fp32_peak()."""
        a, b = C.c_float(), C.c_float()
        _check(load().rt_fp32_peak_detail(self._h, C.byref(a), C.byref(b)))
        return a.value, b.value

    # -- multi-GPU plumbing: framebuffer shared between processes (CUDA IPC)
    def ipc_export_framebuffer(self) -> bytes:
        buf = (C.c_char * 64)()
        _check(load().rt_ipc_export_framebuffer(self._h, buf))
        return bytes(buf)

    def ipc_open(self, handle: bytes) -> int:
        buf = (C.c_char * 64).from_buffer_copy(handle)
        p = C.c_void_p()
        _check(load().rt_ipc_open(self._h, buf, C.byref(p)))
        return p.value

    def ipc_close(self, dev_ptr: int):
        _check(load().rt_ipc_close(self._h, C.c_void_p(dev_ptr)))
