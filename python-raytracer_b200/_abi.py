"""ctypes mirror of `include/b200rt.h` + scene flattening (object graph -> SoA host buffers).

The C ABI library (`csrc/libb200rt.so`, built by `csrc/build.py` / `__graft_entry__.build()`)
is the only compute path.  If it is missing or cannot initialise CUDA, `load_library()`
raises: there is NO CPU fallback in this package.
"""
import ctypes as C
import os

import numpy as np

from . import geometry as geom

RT_ABI_VERSION = 3
KIND_SPHERE, KIND_PLANE, KIND_BOX, KIND_MESH, KIND_NODE = range(5)
MAT_TYPES = {"mirror": 1, "refractive": 2}            # anything else shades as diffuse, scene.py:97-111
LIGHT_TYPES = {"point": 0, "directional": 1}
CSG_TYPES = {"union": 0, "intersection": 1, "difference": 2}   # anything else: no hits, hierarchy.py:49-78
GATE_AABB, GATE_SPHERE = 0, 1

STATUS_NAMES = {0: "RT_OK", 1: "RT_ERR_INVALID", 2: "RT_ERR_CUDA", 3: "RT_ERR_NOMEM",
                4: "RT_ERR_NO_MATERIAL", 5: "RT_ERR_LIMIT"}

f3 = C.c_float * 3


class RtMaterial(C.Structure):
    _fields_ = [("diffuse", f3), ("specular", f3), ("type", C.c_int32), ("_pad", C.c_int32),
                ("hardness", C.c_double), ("tint", C.c_double), ("refr_index", C.c_double)]


class RtLight(C.Structure):
    _fields_ = [("type", C.c_int32), ("colour", f3), ("vector", f3), ("power", C.c_float)]


class RtObject(C.Structure):
    _fields_ = [("kind", C.c_int32), ("n_mats", C.c_int32), ("mat0", C.c_int32), ("mat1", C.c_int32),
                ("tex", C.c_int32), ("mesh", C.c_int32), ("first_child", C.c_int32), ("n_children", C.c_int32),
                ("csg", C.c_int32), ("xform", C.c_int32), ("radius", C.c_float), ("_pad0", C.c_float),
                ("a", f3), ("b", f3), ("speed", f3), ("ax_w", f3), ("ax_h", f3), ("_pad1", C.c_float),
                ("tex_scale", C.c_double), ("radius_d", C.c_double)]


class RtXform(C.Structure):
    _fields_ = [("M", C.c_float * 16), ("Minv", C.c_float * 16)]


class RtMesh(C.Structure):
    _fields_ = [("n_verts", C.c_int32), ("n_faces", C.c_int32),
                ("verts", C.c_void_p), ("faces", C.c_void_p), ("vnormals", C.c_void_p),
                ("flat_shaded", C.c_int32), ("gate", C.c_int32),
                ("gate_min", f3), ("gate_max", f3), ("gate_center", f3), ("gate_radius", C.c_float),
                ("_pad", C.c_float * 2)]


class RtTexture(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("rgb", C.c_void_p)]


class RtSceneDesc(C.Structure):
    _fields_ = [("abi_version", C.c_int32), ("width", C.c_int32), ("height", C.c_int32),
                ("cam_pos", f3), ("cam_u", f3), ("cam_v", f3), ("cam_w", f3),
                ("cam_d", C.c_double), ("left", C.c_double), ("right", C.c_double),
                ("bottom", C.c_double), ("top", C.c_double),
                ("focal_length", C.c_double), ("aperture", C.c_double),
                ("dof_samples", C.c_int32), ("aa_samples", C.c_int32), ("jitter", C.c_int32),
                ("n_times", C.c_int32), ("motion_times", C.c_void_p),
                ("ambient", f3), ("max_depth", C.c_int32), ("jitter_seed", C.c_uint64),
                ("n_lights", C.c_int32), ("n_materials", C.c_int32), ("n_objects", C.c_int32),
                ("n_top", C.c_int32), ("n_xforms", C.c_int32), ("n_meshes", C.c_int32),
                ("n_textures", C.c_int32), ("mesh_motion", C.c_int32),
                ("lights", C.c_void_p), ("materials", C.c_void_p), ("objects", C.c_void_p),
                ("xforms", C.c_void_p), ("meshes", C.c_void_p), ("textures", C.c_void_p)]


class RtRenderJob(C.Structure):
    _fields_ = [("col0", C.c_int32), ("ncols", C.c_int32), ("row0", C.c_int32), ("nrows", C.c_int32),
                ("tile_w", C.c_int32), ("tile_h", C.c_int32), ("rank", C.c_int32), ("world", C.c_int32),
                ("direct", C.c_int32), ("mesh_mode", C.c_int32), ("max_batch_rays", C.c_int64)]


class RtCounters(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in
                ("rays_primary", "rays_shadow", "rays_reflect", "rays_refract", "bvh_nodes", "tri_tests",
                 "prim_tests", "shaded_hits", "mesh_hits_smooth", "queued_rays", "kernel_launches", "error_flags")]

    def as_dict(self):
        return {n: int(getattr(self, n)) for n, _ in self._fields_}


class RtProfile(C.Structure):
    _fields_ = [("ms_hit", C.c_double), ("ms_shadow", C.c_double), ("ms_shade", C.c_double),
                ("ms_secondary", C.c_double), ("ms_resolve", C.c_double),
                ("n_hit", C.c_uint64), ("n_shadow", C.c_uint64), ("n_shade", C.c_uint64),
                ("n_secondary", C.c_uint64), ("n_resolve", C.c_uint64),
                ("hit", RtCounters), ("shadow", RtCounters), ("shade", RtCounters), ("secondary", RtCounters),
                ("ms_l1_hit", C.c_double), ("ms_l1_shadow", C.c_double), ("ms_l1_shade", C.c_double),
                ("ms_deep", C.c_double), ("ms_tail", C.c_double),
                ("n_l1_hit", C.c_uint64), ("n_deep", C.c_uint64), ("n_tail", C.c_uint64)]


class RtIpcHandle(C.Structure):
    _fields_ = [("bytes", C.c_ubyte * 64)]


DIRECT_SHARED_FRAME = 2


class RtPeaks(C.Structure):
    _fields_ = [("fp32_fma_tflops", C.c_double), ("issue_ginst", C.c_double), ("lsu_gwavefronts", C.c_double),
                ("fp32_fma_tflops_nominal", C.c_double), ("issue_ginst_nominal", C.c_double),
                ("lsu_gwavefronts_nominal", C.c_double), ("sm_clock_mhz", C.c_double),
                ("sm_count", C.c_int32), ("_pad", C.c_int32)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_ if n != "_pad"}


DEFAULT_JITTER_SEED = 0x5EED
DEFAULT_MAX_DEPTH = 10  # scene.py:81


def _f3(v):
    return f3(float(v[0]), float(v[1]), float(v[2]))


class FlatScene:
    """Host-side SoA buffers of one Scene + the RtSceneDesc pointing into them."""

    def __init__(self):
        self.keep = []     # numpy arrays / ctypes arrays kept alive for the desc's pointers
        self.desc = None
        self.leaf_names = []


def _texture_rgb(im):
    """What `getpixel` + `pixel[0..2]` sees (simple_geometry.py:168-169): RGB bytes, alpha ignored."""
    if im.mode not in ("RGB", "RGBA"):
        raise TypeError(f"texture mode {im.mode!r}: the reference indexes getpixel() as a tuple (RGB/RGBA only)")
    a = np.asarray(im.convert("RGB"), dtype=np.uint8)
    return np.ascontiguousarray(a)


def flatten(scene, jitter_seed=DEFAULT_JITTER_SEED, max_depth=DEFAULT_MAX_DEPTH):
    """Scene object graph -> FlatScene.  Top-level objects first in scene order; the children of
    each node in one contiguous block (breadth-first), as `include/b200rt.h` specifies."""
    fs = FlatScene()
    vc = scene.vc

    mat_index, mats = {}, []

    def mat_id(m):
        k = id(m)
        if k not in mat_index:
            mat_index[k] = len(mats)
            mats.append(m)
        return mat_index[k]

    for m in scene.materials:
        mat_id(m)

    tex_index, texs = {}, []

    def tex_id(im):
        if im is None:
            return -1
        k = id(im)
        if k not in tex_index:
            tex_index[k] = len(texs)
            texs.append(_texture_rgb(im))
        return tex_index[k]

    objs_py = list(scene.objects)
    recs = [None] * len(objs_py)
    xforms, meshes, mesh_index = [], [], {}
    i = 0
    while i < len(objs_py):
        o = objs_py[i]
        r = RtObject()
        r.n_mats = len(o.materials)
        r.mat0 = mat_id(o.materials[0]) if len(o.materials) > 0 else -1
        r.mat1 = mat_id(o.materials[1]) if len(o.materials) > 1 else -1
        r.tex, r.mesh, r.first_child, r.n_children, r.csg, r.xform = -1, -1, -1, 0, 3, -1
        r.tex_scale, r.radius_d = 1.0, 0.0
        sp = getattr(o, "speed", None)
        r.speed = _f3(sp if sp is not None else (0, 0, 0))
        if isinstance(o, geom.Sphere):
            r.kind = KIND_SPHERE
            r.a = _f3(o.center)
            r.radius = float(o.radius)
            r.radius_d = float(o.radius)
        elif isinstance(o, geom.Plane):
            r.kind = KIND_PLANE
            r.a, r.b = _f3(o.point), _f3(o.normal)
            r.ax_w, r.ax_h = _f3(o.width_axis), _f3(o.height_axis)
            r.tex = tex_id(o.texture)
            r.tex_scale = float(getattr(o, "texture_scale", 1.0))
        elif isinstance(o, geom.AABB):
            r.kind = KIND_BOX
            r.a, r.b = _f3(o.minpos), _f3(o.maxpos)
            r.tex = tex_id(o.texture)
        elif isinstance(o, geom.Mesh):
            r.kind = KIND_MESH
            # instances (deep copies made for `ref` nodes share their arrays, geometry.Mesh.__deepcopy__) -> ONE RtMesh, one LBVH
            key = (id(o.verts), id(o.faces), id(o.norms), bool(o.flat_shaded), id(o.bounding_volume))
            if key not in mesh_index:
                mesh_index[key] = len(meshes)
                meshes.append(o)
            r.mesh = mesh_index[key]
        elif isinstance(o, geom.Hierarchy):
            r.kind = KIND_NODE
            r.csg = CSG_TYPES.get(o.hierarchy_type, 3)
            r.xform = len(xforms)
            xforms.append(o)
            r.first_child = len(objs_py)
            r.n_children = len(o.children)
            objs_py.extend(o.children)
            recs.extend([None] * len(o.children))
        else:
            raise TypeError(f"unsupported geometry {type(o)}")
        recs[i] = r
        fs.leaf_names.append(o.name)
        i += 1

    obj_arr = (RtObject * max(1, len(recs)))(*recs)
    mat_arr = (RtMaterial * max(1, len(mats)))()
    for k, m in enumerate(mats):
        mat_arr[k].diffuse, mat_arr[k].specular = _f3(m.diffuse), _f3(m.specular)
        mat_arr[k].type = MAT_TYPES.get(m.mat_type, 0)
        mat_arr[k].hardness, mat_arr[k].tint, mat_arr[k].refr_index = float(m.hardness), float(m.tint), float(m.refr_index)
    light_arr = (RtLight * max(1, len(scene.lights)))()
    for k, l in enumerate(scene.lights):
        light_arr[k].type = LIGHT_TYPES[l.type]
        light_arr[k].colour, light_arr[k].vector, light_arr[k].power = _f3(l.colour), _f3(l.vector), float(l.power)
    xf_arr = (RtXform * max(1, len(xforms)))()
    for k, n in enumerate(xforms):
        xf_arr[k].M = (C.c_float * 16)(*[float(x) for x in np.asarray(n.M, dtype=np.float32).reshape(-1)])
        xf_arr[k].Minv = (C.c_float * 16)(*[float(x) for x in np.asarray(n.Minv, dtype=np.float32).reshape(-1)])
    mesh_arr = (RtMesh * max(1, len(meshes)))()
    for k, m in enumerate(meshes):
        v = np.ascontiguousarray(m.verts, dtype=np.float32)
        f = np.ascontiguousarray(m.faces, dtype=np.int32)
        fs.keep += [v, f]
        mesh_arr[k].n_verts, mesh_arr[k].n_faces = v.shape[0], f.shape[0]
        mesh_arr[k].verts, mesh_arr[k].faces = v.ctypes.data, f.ctypes.data
        if m.flat_shaded or m.norms is None:
            mesh_arr[k].vnormals = None
        else:
            n = np.ascontiguousarray(m.norms, dtype=np.float32)
            fs.keep.append(n)
            mesh_arr[k].vnormals = n.ctypes.data
        mesh_arr[k].flat_shaded = 1 if m.flat_shaded else 0
        bv = m.bounding_volume
        if isinstance(bv, geom.BoundingAABB):
            mesh_arr[k].gate = GATE_AABB
            mesh_arr[k].gate_min, mesh_arr[k].gate_max = _f3(bv.minpos), _f3(bv.maxpos)
        else:
            mesh_arr[k].gate = GATE_SPHERE
            mesh_arr[k].gate_center, mesh_arr[k].gate_radius = _f3(bv.center), float(bv.radius)
    tex_arr = (RtTexture * max(1, len(texs)))()
    for k, t in enumerate(texs):
        tex_arr[k].height, tex_arr[k].width = t.shape[0], t.shape[1]
        tex_arr[k].rgb = t.ctypes.data
    fs.keep += texs
    times = np.ascontiguousarray(vc.motion_times, dtype=np.float64)

    d = RtSceneDesc()
    d.abi_version = RT_ABI_VERSION
    d.width, d.height = int(vc.width), int(vc.height)
    d.cam_pos, d.cam_u, d.cam_v, d.cam_w = _f3(vc.position), _f3(vc.u), _f3(vc.v), _f3(vc.w)
    d.cam_d, d.left, d.right, d.bottom, d.top = float(vc.d), float(vc.left), float(vc.right), float(vc.bottom), float(vc.top)
    d.focal_length, d.aperture = float(vc.focal_length), float(vc.aperture)
    d.dof_samples, d.aa_samples, d.jitter = int(vc.dof_samples), int(scene.samples), 1 if scene.jitter else 0
    d.n_times, d.motion_times = len(times), times.ctypes.data
    d.ambient = _f3(scene.ambient)
    d.max_depth = int(max_depth)
    d.jitter_seed = int(jitter_seed)
    d.n_lights, d.n_materials, d.n_objects = len(scene.lights), len(mats), len(recs)
    d.n_top, d.n_xforms, d.n_meshes, d.n_textures = len(scene.objects), len(xforms), len(meshes), len(texs)
    d.mesh_motion = 1 if getattr(scene, "mesh_motion", False) else 0
    d.lights = C.cast(light_arr, C.c_void_p)
    d.materials = C.cast(mat_arr, C.c_void_p)
    d.objects = C.cast(obj_arr, C.c_void_p)
    d.xforms = C.cast(xf_arr, C.c_void_p)
    d.meshes = C.cast(mesh_arr, C.c_void_p)
    d.textures = C.cast(tex_arr, C.c_void_p)
    fs.keep += [obj_arr, mat_arr, light_arr, xf_arr, mesh_arr, tex_arr, times]
    fs.desc = d
    fs.n_objects = len(recs)
    return fs


# ---------------------------------------------------------------------------- library

_LIB = None
LIB_PATH = os.environ.get("B200RT_LIB") or os.path.join(os.path.dirname(os.path.abspath(__file__)), "csrc", "libb200rt.so")

ENTRY_POINTS = ["rt_scene_create", "rt_scene_build_bvh", "rt_render", "rt_render_static", "rt_render_host", "rt_render_u8_host", "rt_job_extent",
                "rt_assemble_tiles", "rt_finalize_u8", "rt_scene_destroy", "rt_scene_bvh_stats",
                "rt_scene_bvh_download", "rt_last_error", "rt_abi_version", "rt_trace_rays", "rt_occluded",
                "rt_scene_profile", "rt_scene_profile_read", "rt_measure_peaks",
                "rt_shared_alloc", "rt_shared_free", "rt_shared_open", "rt_shared_close", "rt_obj_load", "rt_obj_copy", "rt_obj_free", "rt_mesh_ingest"]


class B200RTError(RuntimeError):
    pass


def load_library(path=None):
    """dlopen the C ABI library and declare signatures.  Raises if it is missing: the
    product has no other compute path."""
    global _LIB
    if _LIB is not None and path is None:
        return _LIB
    p = path or LIB_PATH
    if not os.path.exists(p):
        raise B200RTError(f"{p} not found: build it with `python __graft_entry__.py` (nvcc, sm_100a); "
                          "there is no CPU fallback")
    lib = C.CDLL(p)
    vp, i32, i64p = C.c_void_p, C.c_int32, C.POINTER(C.c_int64)
    lib.rt_scene_create.argtypes = [C.POINTER(RtSceneDesc), C.POINTER(vp)]
    lib.rt_scene_build_bvh.argtypes = [vp, vp]
    lib.rt_render.argtypes = [vp, C.POINTER(RtRenderJob), vp, C.POINTER(RtCounters), vp]
    lib.rt_render_host.argtypes = [vp, C.POINTER(RtRenderJob), vp, C.POINTER(RtCounters), vp]
    lib.rt_render_static.argtypes = [vp, C.POINTER(RtRenderJob), vp, C.POINTER(RtCounters), vp]
    lib.rt_render_u8_host.argtypes = [vp, C.POINTER(RtRenderJob), vp, C.POINTER(RtCounters), vp]
    lib.rt_job_extent.argtypes = [C.POINTER(RtRenderJob), i64p, i64p]
    lib.rt_assemble_tiles.argtypes = [C.POINTER(RtRenderJob), vp, i64p, vp, vp]
    lib.rt_finalize_u8.argtypes = [vp, i32, i32, vp, vp]
    lib.rt_scene_destroy.argtypes = [vp]
    lib.rt_scene_bvh_stats.argtypes = [vp, i32, i64p, i64p, i64p]
    lib.rt_scene_bvh_download.argtypes = [vp, i32, vp, vp]
    lib.rt_trace_rays.argtypes = [vp, C.c_int64, vp, vp, vp, vp, vp, vp, vp, vp, vp, i32, vp]
    lib.rt_occluded.argtypes = [vp, C.c_int64, vp, vp, vp, vp, vp, i32, vp]
    lib.rt_scene_profile.argtypes = [vp, i32]
    lib.rt_scene_profile_read.argtypes = [vp, C.POINTER(RtProfile)]
    lib.rt_measure_peaks.argtypes = [C.POINTER(RtPeaks), vp]
    lib.rt_shared_alloc.argtypes = [C.c_int64, C.POINTER(vp), C.POINTER(RtIpcHandle)]
    lib.rt_shared_free.argtypes = [vp]
    lib.rt_shared_open.argtypes = [C.POINTER(RtIpcHandle), C.POINTER(vp)]
    lib.rt_shared_close.argtypes = [vp]
    lib.rt_last_error.restype = C.c_char_p
    lib.rt_last_error.argtypes = []
    lib.rt_abi_version.argtypes = []
    lib.rt_obj_load.argtypes = [C.c_char_p, C.POINTER(vp), i64p, i64p, i64p]
    lib.rt_obj_copy.argtypes = [vp, vp, vp, vp]
    lib.rt_obj_free.argtypes = [vp]
    lib.rt_mesh_ingest.argtypes = [i32, vp, i32, vp, vp, C.c_float, i32, vp, vp, vp, vp, vp, C.POINTER(C.c_float), vp]
    for n in ENTRY_POINTS:
        if n not in ("rt_last_error",):
            getattr(lib, n).restype = C.c_int
    if lib.rt_abi_version() != RT_ABI_VERSION:
        raise B200RTError("libb200rt.so ABI version mismatch")
    if path is None:
        _LIB = lib
    return lib


def check(lib, rc, what):
    if rc != 0:
        msg = lib.rt_last_error()
        raise B200RTError(f"{what} failed: {STATUS_NAMES.get(rc, rc)}: {msg.decode() if msg else ''}")


def obj_load(path):
    """Native OBJ reader (`rt_obj_load`): (verts f64 (n,3), norms f64 (m,3), faces i32 (k,3)), same conventions as
    geometry.read_obj."""
    lib = load_library()
    h, nv, nn, nf = C.c_void_p(), C.c_int64(), C.c_int64(), C.c_int64()
    check(lib, lib.rt_obj_load(os.fsencode(path), C.byref(h), C.byref(nv), C.byref(nn), C.byref(nf)), "rt_obj_load")
    try:
        verts = np.empty((nv.value, 3), np.float64)
        norms = np.empty((nn.value, 3), np.float64)
        faces = np.empty((nf.value, 3), np.int32)
        check(lib, lib.rt_obj_copy(h, verts.ctypes.data, norms.ctypes.data, faces.ctypes.data), "rt_obj_copy")
    finally:
        lib.rt_obj_free(h)
    return verts, norms, faces


def mesh_ingest(raw_verts, faces, translate, scale, want_normals, stream=None):
    """Device-side Mesh.__init__ arithmetic (`rt_mesh_ingest`): returns (verts f32, vnormals f32 or None, extent min,
    extent max, centre, max_dist)."""
    lib = load_library()
    raw = np.ascontiguousarray(raw_verts, dtype=np.float64).reshape(-1, 3)
    f = np.ascontiguousarray(faces, dtype=np.int32).reshape(-1, 3)
    t = np.ascontiguousarray(translate, dtype=np.float32).reshape(3)
    verts = np.empty((raw.shape[0], 3), np.float32)
    vn = np.empty((raw.shape[0], 3), np.float32) if want_normals else None
    mn, mx, cen = np.empty(3, np.float32), np.empty(3, np.float32), np.empty(3, np.float32)
    md = C.c_float()
    rc = lib.rt_mesh_ingest(raw.shape[0], raw.ctypes.data, f.shape[0], f.ctypes.data, t.ctypes.data, float(np.float32(scale)),
                            1 if want_normals else 0, verts.ctypes.data, None if vn is None else vn.ctypes.data,
                            mn.ctypes.data, mx.ctypes.data, cen.ctypes.data, C.byref(md), stream)
    check(lib, rc, "rt_mesh_ingest")
    return verts, vn, mn, mx, cen, float(md.value)
