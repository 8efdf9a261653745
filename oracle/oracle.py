"""ctypes front-end of the tier-1 CPU oracle (`oracle/rt_oracle.c`).  TEST INFRASTRUCTURE ONLY:
imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.
It consumes the same `RtSceneDesc` (host pointers) the CUDA library is given."""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "_build", "liboracle.so")
_lib = None


def build(force=False):
    src = os.path.join(HERE, "rt_oracle.c")
    hdr = os.path.join(HERE, "..", "include", "b200rt.h")
    if (not force and os.path.exists(LIB_PATH)
            and os.path.getmtime(LIB_PATH) >= max(os.path.getmtime(src), os.path.getmtime(hdr))):
        return LIB_PATH
    subprocess.run(["make", "-s", "-B", "-C", HERE, "_build/liboracle.so"], check=True)
    return LIB_PATH


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(LIB_PATH)
        vp = C.c_void_p
        L.oracle_render.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int, vp, vp, C.c_int]
        L.oracle_render_pixels.argtypes = [vp, vp, vp, vp, C.c_int64, vp, vp, C.c_int]
        L.oracle_trace_rays.argtypes = [vp, C.c_int64, vp, vp, vp, vp, vp, vp, vp, vp, vp]
        L.oracle_occluded.argtypes = [vp, C.c_int64, vp, vp, vp, vp, vp]
        L.oracle_philox.argtypes = [C.c_uint64, C.c_uint64, vp]
        L.oracle_render_pixels_perturbed.argtypes = [vp, vp, vp, C.c_int64, vp, C.c_double, C.c_uint64, C.c_int]
        L.oracle_pixel_features.argtypes = [vp, vp, vp, C.c_int64, vp, C.c_int]
        _lib = L
    return _lib


def _ctr(abi):
    return abi.RtCounters()


def render(flat, abi, col0=0, ncols=None, row0=0, nrows=None, threads=0):
    """Scene.render over a rectangle -> ((ncols,nrows,3) float64, counters dict)."""
    d = flat.desc
    ncols = d.width - col0 if ncols is None else ncols
    nrows = d.height - row0 if nrows is None else nrows
    out = np.empty((ncols, nrows, 3), dtype=np.float64)
    ctr = _ctr(abi)
    rc = lib().oracle_render(C.addressof(d), col0, ncols, row0, nrows, out.ctypes.data, C.addressof(ctr), threads)
    assert rc == 0, rc
    return out, ctr.as_dict()


def render_pixels(flat, abi, cols, rows, threads=0):
    cols = np.ascontiguousarray(cols, dtype=np.int32)
    rows = np.ascontiguousarray(rows, dtype=np.int32)
    out = np.empty((len(cols), 3), dtype=np.float64)
    ctr = _ctr(abi)
    rc = lib().oracle_render_pixels(C.addressof(flat.desc), cols.ctypes.data, rows.ctypes.data, None, len(cols),
                                    out.ctypes.data, C.addressof(ctr), threads)
    assert rc == 0, rc
    return out, ctr.as_dict()


def render_pixels_perturbed(flat, cols, rows, eps=1e-6, seed=1, threads=0):
    """Stability probe (NOT reference behaviour): the same pixels with every cast_ray input ray and closest-hit position
    moved by a random vector of relative size `eps` (a few float32 ulp)."""
    cols = np.ascontiguousarray(cols, dtype=np.int32)
    rows = np.ascontiguousarray(rows, dtype=np.int32)
    out = np.empty((len(cols), 3), dtype=np.float64)
    rc = lib().oracle_render_pixels_perturbed(C.addressof(flat.desc), cols.ctypes.data, rows.ctypes.data, len(cols),
                                              out.ctypes.data, float(eps), int(seed), threads)
    assert rc == 0, rc
    return out


def unstable_pixels(flat, cols, rows, base_rgb, eps=1e-6, trials=6, threads=0):
    """Which of the given pixels are REFERENCE-UNSTABLE: the oracle's own uint8 value moves by more than 1 when its rays
    and hit points are perturbed by `eps` (relative).  `base_rgb` (n,3) float = the unperturbed oracle values.  A correct
    float32 implementation cannot be expected to agree with the reference on such a pixel."""
    base = (np.asarray(base_rgb, dtype=np.float64) * 255).astype(np.uint8).astype(int)
    unstable = np.zeros(len(cols), dtype=bool)
    for k in range(trials):
        p = (render_pixels_perturbed(flat, cols, rows, eps, seed=1000 + k, threads=threads) * 255).astype(np.uint8).astype(int)
        unstable |= np.abs(p - base).max(axis=1) > 1
    return unstable


def pixel_features(flat, cols, rows, threads=0):
    """(n,4) int32 per pixel: leaf object of the first sample's closest hit, material index, shadowed-light bits,
    material type (0 diffuse / 1 mirror / 2 refractive)."""
    cols = np.ascontiguousarray(cols, dtype=np.int32)
    rows = np.ascontiguousarray(rows, dtype=np.int32)
    out = np.empty((len(cols), 4), dtype=np.int32)
    rc = lib().oracle_pixel_features(C.addressof(flat.desc), cols.ctypes.data, rows.ctypes.data, len(cols), out.ctypes.data, threads)
    assert rc == 0, rc
    return out


def trace_rays(flat, origins, dirs, times=None):
    o = np.ascontiguousarray(origins, dtype=np.float32).reshape(-1, 3)
    d = np.ascontiguousarray(dirs, dtype=np.float32).reshape(-1, 3)
    n = o.shape[0]
    t = np.zeros(n, dtype=np.float64) if times is None else np.ascontiguousarray(times, dtype=np.float64)
    res = dict(rgb=np.empty((n, 3), np.float32), t=np.empty(n, np.float64), n=np.empty((n, 3), np.float32),
               p=np.empty((n, 3), np.float32), geom=np.empty(n, np.int32), mat=np.empty(n, np.int32))
    rc = lib().oracle_trace_rays(C.addressof(flat.desc), n, o.ctypes.data, d.ctypes.data, t.ctypes.data,
                                 res["rgb"].ctypes.data, res["t"].ctypes.data, res["n"].ctypes.data,
                                 res["p"].ctypes.data, res["geom"].ctypes.data, res["mat"].ctypes.data)
    assert rc == 0, rc
    return res


def occluded(flat, origins, dirs, tmax, times=None):
    o = np.ascontiguousarray(origins, dtype=np.float32).reshape(-1, 3)
    d = np.ascontiguousarray(dirs, dtype=np.float32).reshape(-1, 3)
    n = o.shape[0]
    t = np.zeros(n, dtype=np.float64) if times is None else np.ascontiguousarray(times, dtype=np.float64)
    tm = np.ascontiguousarray(np.broadcast_to(np.asarray(tmax, dtype=np.float64), (n,)))
    occ = np.empty(n, np.int32)
    rc = lib().oracle_occluded(C.addressof(flat.desc), n, o.ctypes.data, d.ctypes.data, t.ctypes.data,
                               tm.ctypes.data, occ.ctypes.data)
    assert rc == 0, rc
    return occ


def philox(idx, seed):
    out = (C.c_uint32 * 4)()
    lib().oracle_philox(idx, seed, out)
    return tuple(out)


def to_png_array(image):
    return (np.rot90(image, k=1, axes=(0, 1)) * 255).astype(np.uint8)
