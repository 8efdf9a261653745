"""ctypes front-end of tests/hostsim (host-compiled copy of the per-ray device functions; test aid only)."""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "_build", "libhostsim.so")
CSRC = os.path.join(HERE, "..", "..", "python-raytracer_b200", "csrc")
_lib = None


def build_variant(tag, defines):
    """A second copy of the library with other compile-time limits (e.g. a 3-entry traversal stack to exercise the spill)."""
    out = os.path.join(HERE, "_build", f"libhostsim_{tag}.so")
    os.makedirs(os.path.dirname(out), exist_ok=True)
    subprocess.run(["g++", "-O2", "-std=c++17", "-fopenmp", "-fPIC", "-shared", "-ffp-contract=off", *defines,
                    "-I/usr/local/cuda/include", "-o", out, os.path.join(HERE, "hostsim.cpp")], check=True)
    return C.CDLL(out)


def build(force=False):
    deps = [os.path.join(HERE, "hostsim.cpp")] + [os.path.join(CSRC, f) for f in
            ("rt_core.cuh", "rt_bvh.cuh", "rt_types.cuh", "rt_math.cuh", "rt_scene_build.hpp")]
    if not force and os.path.exists(LIB) and os.path.getmtime(LIB) >= max(os.path.getmtime(d) for d in deps):
        return LIB
    os.makedirs(os.path.dirname(LIB), exist_ok=True)
    subprocess.run(["g++", "-O2", "-std=c++17", "-fopenmp", "-fPIC", "-shared", "-ffp-contract=off",
                    "-I/usr/local/cuda/include", "-o", LIB, os.path.join(HERE, "hostsim.cpp")], check=True)
    return LIB


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(LIB)
        vp = C.c_void_p
        _lib.hostsim_render_pixels.argtypes = [vp, vp, vp, C.c_int64, vp, C.c_int, vp]
        _lib.hostsim_bvh.argtypes = [vp, C.c_int, vp, vp]
    return _lib


def render_pixels(flat, abi, cols, rows, mesh_mode=0, library=None):
    cols = np.ascontiguousarray(cols, dtype=np.int32)
    rows = np.ascontiguousarray(rows, dtype=np.int32)
    out = np.empty((len(cols), 3), dtype=np.float32)
    ctr = abi.RtCounters()
    L = library or lib()
    L.hostsim_render_pixels.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int, C.c_void_p]
    rc = L.hostsim_render_pixels(C.addressof(flat.desc), cols.ctypes.data, rows.ctypes.data, len(cols),
                                     out.ctypes.data, mesh_mode, C.addressof(ctr))
    assert rc == 0, rc
    return out, ctr.as_dict()


def render(flat, abi, mesh_mode=0, library=None):
    W, H = flat.desc.width, flat.desc.height
    cc, rr = np.meshgrid(np.arange(W), np.arange(H), indexing="ij")
    out, ctr = render_pixels(flat, abi, cc.reshape(-1), rr.reshape(-1), mesh_mode, library)
    return out.reshape(W, H, 3), ctr


def trace_rays(flat, origins, dirs, times=None, mesh_mode=0):
    o = np.ascontiguousarray(origins, dtype=np.float32).reshape(-1, 3)
    d = np.ascontiguousarray(dirs, dtype=np.float32).reshape(-1, 3)
    n = o.shape[0]
    t = np.zeros(n, dtype=np.float64) if times is None else np.ascontiguousarray(times, dtype=np.float64)
    res = dict(rgb=np.empty((n, 3), np.float32), t=np.empty(n, np.float32), geom=np.empty(n, np.int32))
    npm = np.zeros((n, 7), np.float32)
    L = lib()
    L.hostsim_trace_rays.argtypes = [C.c_void_p, C.c_int64] + [C.c_void_p] * 7 + [C.c_int]
    rc = L.hostsim_trace_rays(C.addressof(flat.desc), n, o.ctypes.data, d.ctypes.data, t.ctypes.data, res["rgb"].ctypes.data,
                              res["t"].ctypes.data, res["geom"].ctypes.data, npm.ctypes.data, mesh_mode)
    assert rc == 0, rc
    res["n"], res["p"], res["mat"] = npm[:, 0:3].copy(), npm[:, 3:6].copy(), npm[:, 6].astype(np.int32)
    return res


def occluded(flat, origins, dirs, tmax, times=None, mesh_mode=0):
    o = np.ascontiguousarray(origins, dtype=np.float32).reshape(-1, 3)
    d = np.ascontiguousarray(dirs, dtype=np.float32).reshape(-1, 3)
    n = o.shape[0]
    t = np.zeros(n, dtype=np.float64) if times is None else np.ascontiguousarray(times, dtype=np.float64)
    tm = np.ascontiguousarray(np.broadcast_to(np.asarray(tmax, dtype=np.float64), (n,)))
    occ = np.empty(n, np.int32)
    L = lib()
    L.hostsim_occluded.argtypes = [C.c_void_p, C.c_int64] + [C.c_void_p] * 5 + [C.c_int]
    rc = L.hostsim_occluded(C.addressof(flat.desc), n, o.ctypes.data, d.ctypes.data, t.ctypes.data, tm.ctypes.data,
                            occ.ctypes.data, mesh_mode)
    assert rc == 0, rc
    return occ


def bvh(flat, mesh=0):
    n = C.cast(flat.desc.meshes, C.POINTER(type("M", (C.Structure,), {"_fields_": [("n_verts", C.c_int32), ("n_faces", C.c_int32)]})))[0].n_faces
    nodes = np.zeros((max(n - 1, 0), 32), dtype=np.float32)      # 4-wide nodes, 128 B each (BvhNode4)
    order = np.zeros(n, dtype=np.int32)
    n4 = lib().hostsim_bvh(C.addressof(flat.desc), mesh, nodes.ctypes.data, order.ctypes.data)
    assert n4 >= 0
    return nodes[:n4].copy(), order


def job_pixels(job):
    L = lib()
    L.hostsim_job_pixels.restype = C.c_longlong
    L.hostsim_job_pixels.argtypes = [C.c_void_p] * 4
    n = L.hostsim_job_pixels(C.addressof(job), None, None, None)
    assert n >= 0
    cols, rows, valid = (np.zeros(n, dtype=np.int32) for _ in range(3))
    L.hostsim_job_pixels(C.addressof(job), cols.ctypes.data, rows.ctypes.data, valid.ctypes.data)
    return cols, rows, valid.astype(bool)
