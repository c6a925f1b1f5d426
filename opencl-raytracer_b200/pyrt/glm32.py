"""Vectorised float32 restatement of the GLM 0.9.9 routines the reference's host code
uses to build ObjectData (SceneLoader.cpp:215-294, ObjectData.cpp:7-8), for generating
synthetic scenes (cfg5) without the C++ loader.  Same formulas and operation order as
compat/glm/*.hpp.  Arrays are (n, 4, 4) float32 indexed m[:, col, row] (column-major).

Host logic only.
"""
from __future__ import annotations

import numpy as np

F = np.float32


def identity(n: int) -> np.ndarray:
    m = np.zeros((n, 4, 4), dtype=F)
    for i in range(4):
        m[:, i, i] = 1.0
    return m


def matmul(a: np.ndarray, b: np.ndarray) -> np.ndarray:
    """column i = a0*b[i][0] + a1*b[i][1] + a2*b[i][2] + a3*b[i][3], left to right."""
    r = np.empty_like(a)
    for i in range(4):
        acc = a[:, 0, :] * b[:, i, 0:1]
        acc = acc + a[:, 1, :] * b[:, i, 1:2]
        acc = acc + a[:, 2, :] * b[:, i, 2:3]
        acc = acc + a[:, 3, :] * b[:, i, 3:4]
        r[:, i, :] = acc
    return r


def translate(m: np.ndarray, v: np.ndarray) -> np.ndarray:
    r = m.copy()
    r[:, 3, :] = m[:, 0, :] * v[:, 0:1] + m[:, 1, :] * v[:, 1:2] + m[:, 2, :] * v[:, 2:3] + m[:, 3, :]
    return r


def scale(m: np.ndarray, v: np.ndarray) -> np.ndarray:
    r = m.copy()
    for i in range(3):
        r[:, i, :] = m[:, i, :] * v[:, i : i + 1]
    return r


def normalize3(v: np.ndarray) -> np.ndarray:
    d = v[:, 0] * v[:, 0] + v[:, 1] * v[:, 1] + v[:, 2] * v[:, 2]
    inv = F(1.0) / np.sqrt(d, dtype=F)
    return (v * inv[:, None]).astype(F)


def rotate(m: np.ndarray, angle: np.ndarray, axis_in: np.ndarray) -> np.ndarray:
    """glm::rotate(m, angle[rad], axis): Rodrigues with temp = (1-c)*axis."""
    c = np.cos(angle).astype(F)
    s = np.sin(angle).astype(F)
    axis = normalize3(axis_in.astype(F))
    temp = ((F(1.0) - c)[:, None] * axis).astype(F)
    n = m.shape[0]
    R = np.zeros((n, 3, 3), dtype=F)
    R[:, 0, 0] = c + temp[:, 0] * axis[:, 0]
    R[:, 0, 1] = temp[:, 0] * axis[:, 1] + s * axis[:, 2]
    R[:, 0, 2] = temp[:, 0] * axis[:, 2] - s * axis[:, 1]
    R[:, 1, 0] = temp[:, 1] * axis[:, 0] - s * axis[:, 2]
    R[:, 1, 1] = c + temp[:, 1] * axis[:, 1]
    R[:, 1, 2] = temp[:, 1] * axis[:, 2] + s * axis[:, 0]
    R[:, 2, 0] = temp[:, 2] * axis[:, 0] + s * axis[:, 1]
    R[:, 2, 1] = temp[:, 2] * axis[:, 1] - s * axis[:, 0]
    R[:, 2, 2] = c + temp[:, 2] * axis[:, 2]
    r = m.copy()
    for i in range(3):
        r[:, i, :] = m[:, 0, :] * R[:, i, 0:1] + m[:, 1, :] * R[:, i, 1:2] + m[:, 2, :] * R[:, i, 2:3]
    return r


def look_at_default(n: int) -> np.ndarray:
    """glm::lookAt((0,0,10),(0,0,0),(0,1,0)) (SceneLoader.cpp:215) == translate(0,0,-10), exact."""
    m = identity(n)
    m[:, 3, 2] = -10.0
    return m


def inverse(m: np.ndarray) -> np.ndarray:
    """compute_inverse<4,4,float>: cofactors, SignA/SignB, det from column 0, times 1/det."""
    g = lambda c, r: m[:, c, r]
    Coef00 = g(2, 2) * g(3, 3) - g(3, 2) * g(2, 3)
    Coef02 = g(1, 2) * g(3, 3) - g(3, 2) * g(1, 3)
    Coef03 = g(1, 2) * g(2, 3) - g(2, 2) * g(1, 3)
    Coef04 = g(2, 1) * g(3, 3) - g(3, 1) * g(2, 3)
    Coef06 = g(1, 1) * g(3, 3) - g(3, 1) * g(1, 3)
    Coef07 = g(1, 1) * g(2, 3) - g(2, 1) * g(1, 3)
    Coef08 = g(2, 1) * g(3, 2) - g(3, 1) * g(2, 2)
    Coef10 = g(1, 1) * g(3, 2) - g(3, 1) * g(1, 2)
    Coef11 = g(1, 1) * g(2, 2) - g(2, 1) * g(1, 2)
    Coef12 = g(2, 0) * g(3, 3) - g(3, 0) * g(2, 3)
    Coef14 = g(1, 0) * g(3, 3) - g(3, 0) * g(1, 3)
    Coef15 = g(1, 0) * g(2, 3) - g(2, 0) * g(1, 3)
    Coef16 = g(2, 0) * g(3, 2) - g(3, 0) * g(2, 2)
    Coef18 = g(1, 0) * g(3, 2) - g(3, 0) * g(1, 2)
    Coef19 = g(1, 0) * g(2, 2) - g(2, 0) * g(1, 2)
    Coef20 = g(2, 0) * g(3, 1) - g(3, 0) * g(2, 1)
    Coef22 = g(1, 0) * g(3, 1) - g(3, 0) * g(1, 1)
    Coef23 = g(1, 0) * g(2, 1) - g(2, 0) * g(1, 1)

    st = lambda *a: np.stack(a, axis=1).astype(F)
    Fac0 = st(Coef00, Coef00, Coef02, Coef03)
    Fac1 = st(Coef04, Coef04, Coef06, Coef07)
    Fac2 = st(Coef08, Coef08, Coef10, Coef11)
    Fac3 = st(Coef12, Coef12, Coef14, Coef15)
    Fac4 = st(Coef16, Coef16, Coef18, Coef19)
    Fac5 = st(Coef20, Coef20, Coef22, Coef23)
    Vec0 = st(g(1, 0), g(0, 0), g(0, 0), g(0, 0))
    Vec1 = st(g(1, 1), g(0, 1), g(0, 1), g(0, 1))
    Vec2 = st(g(1, 2), g(0, 2), g(0, 2), g(0, 2))
    Vec3 = st(g(1, 3), g(0, 3), g(0, 3), g(0, 3))

    Inv0 = Vec1 * Fac0 - Vec2 * Fac1 + Vec3 * Fac2
    Inv1 = Vec0 * Fac0 - Vec2 * Fac3 + Vec3 * Fac4
    Inv2 = Vec0 * Fac1 - Vec1 * Fac3 + Vec3 * Fac5
    Inv3 = Vec0 * Fac2 - Vec1 * Fac4 + Vec2 * Fac5
    SignA = np.array([+1, -1, +1, -1], dtype=F)
    SignB = np.array([-1, +1, -1, +1], dtype=F)
    Inv = np.stack([Inv0 * SignA, Inv1 * SignB, Inv2 * SignA, Inv3 * SignB], axis=1).astype(F)

    Row0 = Inv[:, :, 0]
    Dot0 = m[:, 0, :] * Row0
    Dot1 = (Dot0[:, 0] + Dot0[:, 1]) + (Dot0[:, 2] + Dot0[:, 3])
    one_over_det = (F(1.0) / Dot1).astype(F)
    return (Inv * one_over_det[:, None, None]).astype(F)
