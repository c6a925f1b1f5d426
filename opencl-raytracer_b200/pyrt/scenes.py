"""Scene sources for tests and bench: committed fixtures of the reference's four *.txt
scenes (loaded by the reference's own SceneLoader, see tools/make_fixtures.sh), the
glm-independent KAT-2 scene of SURVEY.md section 4, and the synthetic cfg5 generator
(SURVEY.md 8(d)).  Host logic only.
"""
from __future__ import annotations

import os
import struct

import numpy as np

from . import glm32
from .types import BOX, LIGHT_DTYPE, OBJECT_DTYPE, SPHERE

_ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
FIXTURE_DIR = os.path.join(_ROOT, "tests", "golden", "scenes")

REFERENCE_SCENES = ("simpleSphere", "multipleSpheres", "simpleScene", "roundedCube")


def load_scene_bin(path: str):
    """Read a .scene.bin written by tools/dump_scene.cpp -> (objects, lights) structured arrays."""
    with open(path, "rb") as f:
        raw = f.read()
    if raw[:4] != b"RTSC":
        raise ValueError("%s: not an RTSC scene file" % path)
    ver, n_o, n_l = struct.unpack_from("<III", raw, 4)
    if ver != 1:
        raise ValueError("%s: unsupported version %d" % (path, ver))
    off = 16
    objs = np.frombuffer(raw, dtype=OBJECT_DTYPE, count=n_o, offset=off).copy()
    off += n_o * OBJECT_DTYPE.itemsize
    lights = np.frombuffer(raw, dtype=LIGHT_DTYPE, count=n_l, offset=off).copy()
    if off + n_l * LIGHT_DTYPE.itemsize != len(raw):
        raise ValueError("%s: trailing bytes" % path)
    return objs, lights


def save_scene_bin(path: str, objs: np.ndarray, lights: np.ndarray) -> None:
    with open(path, "wb") as f:
        f.write(b"RTSC")
        f.write(struct.pack("<III", 1, len(objs), len(lights)))
        f.write(np.ascontiguousarray(objs).tobytes())
        f.write(np.ascontiguousarray(lights).tobytes())


def load_reference_scene(name: str):
    return load_scene_bin(os.path.join(FIXTURE_DIR, name + ".scene.bin"))


def _diag_translate(s: float, t) -> tuple:
    """mv = diag(s,s,s,1) + translation t;  mvInverse = diag(1/s) + translation -t/s (float32)."""
    s32 = np.float32(s)
    mv = np.zeros((4, 4), dtype=np.float32)
    inv = np.zeros((4, 4), dtype=np.float32)
    for i in range(3):
        mv[i, i] = s32
        inv[i, i] = np.float32(1.0) / s32
        mv[3, i] = np.float32(t[i])
        inv[3, i] = -np.float32(t[i]) / s32
    mv[3, 3] = inv[3, 3] = 1.0
    return mv, inv


def kat2_scene():
    """KAT-2 of SURVEY.md section 4: box s=4 at (0,0,-10) + two unit spheres, one light."""
    objs = np.zeros(3, dtype=OBJECT_DTYPE)
    mv, inv = _diag_translate(4.0, (0, 0, -10))
    objs[0]["type"] = BOX
    objs[0]["mv"], objs[0]["mvInverse"] = mv.reshape(16), inv.reshape(16)
    m = objs[0]["mat"]
    m["ambient"], m["diffuse"], m["specular"] = (0.1,) * 3, (0.6,) * 3, (1.0,) * 3
    m["absorption"], m["shininess"] = 0.2, 100.0
    for k, t in ((1, (3, 2, -7)), (2, (-3, -2, -7))):
        mv, inv = _diag_translate(1.0, t)
        objs[k]["type"] = SPHERE
        objs[k]["mv"], objs[k]["mvInverse"] = mv.reshape(16), inv.reshape(16)
        m = objs[k]["mat"]
        m["ambient"], m["diffuse"], m["specular"] = (1, 0, 0), (0, 1, 0), (0, 0, 1)
        m["absorption"], m["shininess"] = 0.7, 1.0
    lights = np.zeros(1, dtype=LIGHT_DTYPE)
    lights[0]["ambient"], lights[0]["diffuse"], lights[0]["specular"] = (0.3,) * 3, (0.7,) * 3, (1.0,) * 3
    lights[0]["position"] = (10, 10, 0, 1)
    return objs, lights


def synthetic_scene(n_second: int = 100_000, n_spheres: int = 4096, seed: int = 0x5EED1234, second_type: int = BOX,
                    n_lights: int = 2):
    """cfg5 of SURVEY.md 8(d): `n_spheres` spheres + `n_second` primitives of `second_type`,
    view-space centres uniform in x[-40,40] y[-22,22] z[-120,-20];
    mv = lookAt * translate(c) * rotate(theta, axis) * scale(s), s ~ U[0.15,0.6] per axis;
    8 materials cycling absorption {1,.8,.6,.4} x shininess {1,32}; `n_lights` lights.
    Object order: primitives are interleaved (sphere every (n_second/n_spheres)-th slot) so that
    both types occur throughout the index range.
    """
    rng = np.random.Generator(np.random.PCG64(seed))
    n = n_second + n_spheres
    F = np.float32
    centre = np.stack([rng.uniform(-40, 40, n), rng.uniform(-22, 22, n), rng.uniform(-120, -20, n)], axis=1).astype(F)
    # the scene file's coordinates are world space; lookAt moves them by -10 in z
    world = centre.copy()
    world[:, 2] += F(10.0)
    theta = rng.uniform(0.0, 360.0, n).astype(F) * F(0.01745329251994329576923690768489)
    axis = rng.normal(size=(n, 3)).astype(F)
    axis[np.abs(axis).sum(axis=1) == 0] = (1, 0, 0)
    sc = rng.uniform(0.15, 0.6, size=(n, 3)).astype(F)

    mv = glm32.look_at_default(n)
    mv = glm32.matmul(mv, glm32.translate(glm32.identity(n), world))
    mv = glm32.matmul(mv, glm32.rotate(glm32.identity(n), theta, axis))
    mv = glm32.matmul(mv, glm32.scale(glm32.identity(n), sc))
    inv = glm32.inverse(mv)

    types = np.full(n, second_type, dtype=np.uint32)
    if n_spheres:
        types[np.linspace(0, n - 1, n_spheres).astype(np.int64)] = SPHERE
        # collisions of linspace indices cannot happen for n_spheres <= n
    pal_abs = np.array([1.0, 0.8, 0.6, 0.4, 1.0, 0.8, 0.6, 0.4], dtype=F)
    pal_shn = np.array([1.0, 1.0, 1.0, 1.0, 32.0, 32.0, 32.0, 32.0], dtype=F)
    pal_col = rng.uniform(0.0, 1.0, size=(8, 3, 3)).astype(F)
    mid = np.arange(n) % 8

    objs = np.zeros(n, dtype=OBJECT_DTYPE)
    objs["mv"] = mv.reshape(n, 16)
    objs["mvInverse"] = inv.reshape(n, 16)
    objs["type"] = types
    objs["mat"]["ambient"] = pal_col[mid, 0] * F(0.5)
    objs["mat"]["diffuse"] = pal_col[mid, 1]
    objs["mat"]["specular"] = pal_col[mid, 2]
    objs["mat"]["absorption"] = pal_abs[mid]
    objs["mat"]["reflection"] = F(1.0) - pal_abs[mid]
    objs["mat"]["shininess"] = pal_shn[mid]

    lights = np.zeros(n_lights, dtype=LIGHT_DTYPE)
    lpos = [(30.0, 40.0, 0.0, 1.0), (-45.0, 25.0, -40.0, 1.0), (0.0, -30.0, -10.0, 1.0), (10.0, 60.0, -90.0, 1.0)]
    for i in range(n_lights):
        lights[i]["ambient"] = (0.3, 0.3, 0.3)
        lights[i]["diffuse"] = (0.7, 0.7, 0.7) if i % 2 == 0 else (0.5, 0.5, 0.6)
        lights[i]["specular"] = (1.0, 1.0, 1.0)
        lights[i]["position"] = lpos[i % len(lpos)]
    return objs, lights
