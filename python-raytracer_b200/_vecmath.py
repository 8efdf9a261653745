"""float32 vector / matrix helpers for the host side of the render path.

The reference does its scene set-up with PyGLM float32 types (`helperclasses.py:81-94`,
`hierarchy.py:30-37`, `mesh.py:24,53-70`).  The host code here is numpy; these helpers keep
every intermediate in float32 and use the same operation order as GLM's generic C++ so the
buffers handed to the kernels are the values the reference would have computed.
"""
import math

import numpy as np

F = np.float32


def vec3(v):
    a = np.asarray(v, dtype=np.float64).reshape(3)
    return a.astype(np.float32)


def dot(a, b):
    return F(F(F(a[0] * b[0]) + F(a[1] * b[1])) + F(a[2] * b[2]))


def cross(a, b):
    return np.array([
        F(a[1] * b[2]) - F(b[1] * a[2]),
        F(a[2] * b[0]) - F(b[2] * a[0]),
        F(a[0] * b[1]) - F(b[0] * a[1]),
    ], dtype=np.float32)


def length(a):
    return F(np.sqrt(dot(a, a)))


def normalize(a):
    with np.errstate(all="ignore"):
        inv = F(1.0) / F(np.sqrt(dot(a, a)))
        return (a * inv).astype(np.float32)


# ---- batched (n,3) versions, same per-element operation order -------------------------

def dot_rows(a, b):
    p = a * b
    return (p[:, 0] + p[:, 1]) + p[:, 2]


def cross_rows(a, b):
    return np.stack([
        a[:, 1] * b[:, 2] - b[:, 1] * a[:, 2],
        a[:, 2] * b[:, 0] - b[:, 2] * a[:, 0],
        a[:, 0] * b[:, 1] - b[:, 0] * a[:, 1],
    ], axis=1)


def normalize_rows(a):
    with np.errstate(all="ignore"):
        inv = F(1.0) / np.sqrt(dot_rows(a, a))
        return a * inv[:, None]


# ---- 4x4 matrices, column-major m[col, row] like GLM ------------------------------------

def identity():
    return np.eye(4, dtype=np.float32)


def translate(m, v):
    r = m.copy()
    c = m[0] * F(v[0])
    c = c + m[1] * F(v[1])
    c = c + m[2] * F(v[2])
    r[3] = c + m[3]
    return r


def rotate(m, angle, axis):
    a = F(angle)
    c = F(math.cos(float(a)))
    s = F(math.sin(float(a)))
    ax = normalize(np.asarray(axis, dtype=np.float32))
    temp = (F(1.0) - c) * ax
    R = np.zeros((3, 3), dtype=np.float32)
    R[0, 0] = c + temp[0] * ax[0]
    R[0, 1] = temp[0] * ax[1] + s * ax[2]
    R[0, 2] = temp[0] * ax[2] - s * ax[1]
    R[1, 0] = temp[1] * ax[0] - s * ax[2]
    R[1, 1] = c + temp[1] * ax[1]
    R[1, 2] = temp[1] * ax[2] + s * ax[0]
    R[2, 0] = temp[2] * ax[0] + s * ax[1]
    R[2, 1] = temp[2] * ax[1] - s * ax[0]
    R[2, 2] = c + temp[2] * ax[2]
    r = np.empty((4, 4), dtype=np.float32)
    for j in range(3):
        col = m[0] * R[j, 0]
        col = col + m[1] * R[j, 1]
        col = col + m[2] * R[j, 2]
        r[j] = col
    r[3] = m[3]
    return r


def scale(m, v):
    r = m.copy()
    r[0] = m[0] * F(v[0])
    r[1] = m[1] * F(v[1])
    r[2] = m[2] * F(v[2])
    return r


def inverse(m):
    """Cofactor inverse in float32, term order of GLM's compute_inverse for 4x4."""
    def s2(a, b, c, d):
        return F(F(a * b) - F(c * d))

    c00 = s2(m[2, 2], m[3, 3], m[3, 2], m[2, 3])
    c02 = s2(m[1, 2], m[3, 3], m[3, 2], m[1, 3])
    c03 = s2(m[1, 2], m[2, 3], m[2, 2], m[1, 3])
    c04 = s2(m[2, 1], m[3, 3], m[3, 1], m[2, 3])
    c06 = s2(m[1, 1], m[3, 3], m[3, 1], m[1, 3])
    c07 = s2(m[1, 1], m[2, 3], m[2, 1], m[1, 3])
    c08 = s2(m[2, 1], m[3, 2], m[3, 1], m[2, 2])
    c10 = s2(m[1, 1], m[3, 2], m[3, 1], m[1, 2])
    c11 = s2(m[1, 1], m[2, 2], m[2, 1], m[1, 2])
    c12 = s2(m[2, 0], m[3, 3], m[3, 0], m[2, 3])
    c14 = s2(m[1, 0], m[3, 3], m[3, 0], m[1, 3])
    c15 = s2(m[1, 0], m[2, 3], m[2, 0], m[1, 3])
    c16 = s2(m[2, 0], m[3, 2], m[3, 0], m[2, 2])
    c18 = s2(m[1, 0], m[3, 2], m[3, 0], m[1, 2])
    c19 = s2(m[1, 0], m[2, 2], m[2, 0], m[1, 2])
    c20 = s2(m[2, 0], m[3, 1], m[3, 0], m[2, 1])
    c22 = s2(m[1, 0], m[3, 1], m[3, 0], m[1, 1])
    c23 = s2(m[1, 0], m[2, 1], m[2, 0], m[1, 1])
    f = lambda *a: np.array(a, dtype=np.float32)
    fac0, fac1, fac2 = f(c00, c00, c02, c03), f(c04, c04, c06, c07), f(c08, c08, c10, c11)
    fac3, fac4, fac5 = f(c12, c12, c14, c15), f(c16, c16, c18, c19), f(c20, c20, c22, c23)
    v0 = f(m[1, 0], m[0, 0], m[0, 0], m[0, 0])
    v1 = f(m[1, 1], m[0, 1], m[0, 1], m[0, 1])
    v2 = f(m[1, 2], m[0, 2], m[0, 2], m[0, 2])
    v3 = f(m[1, 3], m[0, 3], m[0, 3], m[0, 3])
    inv0 = (v1 * fac0 - v2 * fac1) + v3 * fac2
    inv1 = (v0 * fac0 - v2 * fac3) + v3 * fac4
    inv2 = (v0 * fac1 - v1 * fac3) + v3 * fac5
    inv3 = (v0 * fac2 - v1 * fac4) + v2 * fac5
    sa, sb = f(1, -1, 1, -1), f(-1, 1, -1, 1)
    inv = np.stack([inv0 * sa, inv1 * sb, inv2 * sa, inv3 * sb]).astype(np.float32)
    d0 = m[0] * inv[:, 0]
    det = F(F(d0[0] + d0[1]) + F(d0[2] + d0[3]))
    with np.errstate(all="ignore"):
        return (inv * (F(1.0) / det)).astype(np.float32)


def trs(t, r_deg, s):
    """hierarchy.py:30-37: M = T * Rx * Ry * Rz * S (angles in degrees); returns (M, Minv)."""
    m = identity()
    m = translate(m, t)
    m = rotate(m, math.radians(float(r_deg[0])), (1, 0, 0))
    m = rotate(m, math.radians(float(r_deg[1])), (0, 1, 0))
    m = rotate(m, math.radians(float(r_deg[2])), (0, 0, 1))
    m = scale(m, s)
    return m, inverse(m)
