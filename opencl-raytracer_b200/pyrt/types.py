"""numpy mirrors of the C-ABI structs in include/rt_b200.h and host-side helpers
that restate the reference's host glue (ray generation, PPM quantisation).

Host logic only -- no device code and no oracle here.
"""
from __future__ import annotations

import math

import numpy as np

# rt_material / rt_object / rt_light / rt_ray (include/rt_b200.h); all packed float32/uint32.
MATERIAL_DTYPE = np.dtype(
    [
        ("ambient", "<f4", (3,)),
        ("diffuse", "<f4", (3,)),
        ("specular", "<f4", (3,)),
        ("absorption", "<f4"),
        ("reflection", "<f4"),
        ("transparency", "<f4"),
        ("shininess", "<f4"),
    ]
)
OBJECT_DTYPE = np.dtype(
    [
        ("mat", MATERIAL_DTYPE),
        ("mv", "<f4", (16,)),
        ("mvInverse", "<f4", (16,)),
        ("type", "<u4"),
    ]
)
LIGHT_DTYPE = np.dtype(
    [
        ("ambient", "<f4", (3,)),
        ("diffuse", "<f4", (3,)),
        ("specular", "<f4", (3,)),
        ("position", "<f4", (4,)),
    ]
)
RAY_DTYPE = np.dtype([("start", "<f4", (4,)), ("direction", "<f4", (4,))])

assert MATERIAL_DTYPE.itemsize == 52
assert OBJECT_DTYPE.itemsize == 52 + 64 + 64 + 4
assert LIGHT_DTYPE.itemsize == 52
assert RAY_DTYPE.itemsize == 32

SPHERE, BOX = 0, 1
TRIANGLE = 2  # extension (no reference primitive): unit triangle x >= -0.5, y >= -0.5, x + y <= 0, z = 0
MODE_HITTEST, MODE_SHADE, MODE_SAR = 0, 1, 2
ACCEL_AUTO, ACCEL_BRUTE, ACCEL_CULL, ACCEL_BVH = -1, 0, 1, 2
MAX_FLOAT = np.float32(3.402823466e38)


def camera_zdir(height: int, fov_deg: float = 60.0) -> np.float32:
    """z component of every primary ray direction (OpenCL-Raytracer.cpp:18-33).

    main(): fov = glm::radians(60.f) * 0.5f (float); screenSpaceToViewSpace:
    halfHeight = height / 2.0f; z = -(halfHeight / tan(angle)).  `tan` is the
    unqualified C++ call on a float, i.e. the float overload.
    """
    fov = np.float32(fov_deg) * np.float32(0.01745329251994329576923690768489)
    fov = np.float32(fov * np.float32(0.5))
    half_h = np.float32(height) / np.float32(2.0)
    # tanf in float: go through float64 tan and round (glibc tanf is correctly rounded
    # to < 1 ulp; the value is passed to both the oracle and the CUDA path, so any last-ulp
    # difference to an MSVC build is common to both)
    t = np.float32(math.tan(float(fov)))
    return np.float32(-(half_h / t))


def camera_rays(width: int, height: int, zdir: np.float32, rows=None) -> np.ndarray:
    """Primary rays as main() builds them (OpenCL-Raytracer.cpp:68-72), row-major jj*W+ii.

    start = (0,0,0,1); direction = (ii - W/2, (H - jj) - H/2, zdir, 0).  `rows`
    optionally restricts to a list/array of jj values (rays of those rows, in order).
    """
    jj = np.arange(height, dtype=np.int64) if rows is None else np.asarray(rows, dtype=np.int64)
    ii = np.arange(width, dtype=np.int64)
    half_w = np.float32(width) / np.float32(2.0)
    half_h = np.float32(height) / np.float32(2.0)
    x = ii.astype(np.float32) - half_w
    y = (height - jj).astype(np.float32) - half_h
    rays = np.zeros(len(jj) * width, dtype=RAY_DTYPE)
    rays["start"][:, 3] = 1.0
    d = rays["direction"].reshape(len(jj), width, 4)
    d[:, :, 0] = x[None, :]
    d[:, :, 1] = y[:, None]
    d[:, :, 2] = zdir
    return rays


def quantize_ppm(rgba: np.ndarray) -> np.ndarray:
    """PPMExporter.cpp:19-21: min(255, (int)floorf(c * 255.f)) per channel (no lower clamp)."""
    c = np.asarray(rgba, dtype=np.float32)[..., :3]
    q = np.floor(c * np.float32(255.0)).astype(np.int64)
    return np.minimum(255, q).astype(np.int32)


def write_ppm_p3(path: str, width: int, height: int, rgb: np.ndarray) -> None:
    """ASCII P3 layout of PPMExporter.cpp:13-22 ("r g b\\n" per pixel, row 0 first)."""
    rgb = np.asarray(rgb).reshape(height * width, 3)
    with open(path, "w") as f:
        f.write("P3\n%d %d\n255\n" % (width, height))
        np.savetxt(f, rgb, fmt="%d %d %d")


def make_object(obj_type: int, mv: np.ndarray, mv_inverse: np.ndarray, **mat) -> np.ndarray:
    """One rt_object from column-major 4x4 mv / mvInverse (as 4x4 arrays m[col][row] or flat 16)."""
    o = np.zeros((), dtype=OBJECT_DTYPE)
    o["mat"]["absorption"] = 1.0  # Material.hpp:9-10 defaults
    o["mat"]["shininess"] = 1.0
    for k, v in mat.items():
        o["mat"][k] = v
    o["mv"] = np.asarray(mv, dtype=np.float32).reshape(16)
    o["mvInverse"] = np.asarray(mv_inverse, dtype=np.float32).reshape(16)
    o["type"] = obj_type
    return o


def make_light(position, ambient=(0, 0, 0), diffuse=(0, 0, 0), specular=(0, 0, 0)) -> np.ndarray:
    l = np.zeros((), dtype=LIGHT_DTYPE)
    l["ambient"], l["diffuse"], l["specular"] = ambient, diffuse, specular
    l["position"] = position
    return l
