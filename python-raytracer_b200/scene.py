"""`Scene`: drop-in for the reference's `provided/scene.py:16-79`.

Same constructor and the same `render(subimage, tasks)` contract (scene.py:35-39,73): returns a
freshly allocated float64 ndarray of shape (len(array_split(arange(W), tasks)[subimage]), H, 3),
axis 0 = column left->right, axis 1 = row bottom->top, values in [0, 1].  Everything inside
the reference's loops (scene.py:47-77: sample generation, cast_ray, shading) runs as CUDA
kernels behind the C ABI in `include/b200rt.h`; the host only flattens the scene once and
copies the image back.  There is no CPU implementation of the path in this package.
"""
import ctypes as C

import numpy as np

from . import _abi


class Scene:
    def __init__(self, vc, jitter: bool, samples: int, ambient, lights: list, materials: list, objects: list):
        self.vc = vc                # viewport camera
        self.jitter = jitter        # should rays be jittered
        self.samples = samples      # number of AA rays per pixel
        self.ambient = np.asarray(ambient, dtype=np.float32)
        self.lights = lights
        self.materials = materials
        self.objects = objects
        self.jitter_seed = _abi.DEFAULT_JITTER_SEED
        self.max_recursion = _abi.DEFAULT_MAX_DEPTH   # cast_ray default, scene.py:81
        self.tile = (8, 8)          # interleaved tile size used for load balance (kernel-side order only)
        # EXTENSION, off = reference behaviour (the reference stores a mesh's speed and never uses it, mesh.py:72-153):
        # when True, mesh objects translate by speed * time like the primitives do (JSON: top-level "mesh_motion": true)
        self.mesh_motion = False
        self.last_counters = None
        self._flat = None
        self._handle = None
        self._lib = None

    # ------------------------------------------------------------------ device scene lifetime
    def flatten(self):
        """Host SoA buffers + RtSceneDesc (also what the tests hand to the oracle)."""
        if self._flat is None:
            self._flat = _abi.flatten(self, self.jitter_seed, self.max_recursion)
        return self._flat

    def invalidate(self):
        """Call after mutating the scene graph; the next render re-uploads it."""
        self.close()
        self._flat = None

    def upload(self, stream=None):
        """rt_scene_create + rt_scene_build_bvh (once per scene)."""
        if self._handle is None:
            lib = _abi.load_library()
            fs = self.flatten()
            h = C.c_void_p()
            _abi.check(lib, lib.rt_scene_create(C.byref(fs.desc), C.byref(h)), "rt_scene_create")
            self._lib, self._handle = lib, h
            _abi.check(lib, lib.rt_scene_build_bvh(h, stream), "rt_scene_build_bvh")
        return self._handle

    def close(self):
        if self._handle is not None:
            self._lib.rt_scene_destroy(self._handle)
            self._handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ------------------------------------------------------------------ the drop-in call
    def strip_columns(self, subimage: int = 0, tasks: int = 1):
        widths = np.array_split(np.arange(self.vc.width), tasks)   # scene.py:36-37
        return widths[subimage]

    def make_job(self, col0, ncols, row0=0, nrows=None, rank=0, world=1, direct=True, mesh_mode=0,
                 max_batch_rays=0):
        job = _abi.RtRenderJob()
        job.col0, job.ncols = int(col0), int(ncols)
        job.row0, job.nrows = int(row0), int(self.vc.height if nrows is None else nrows)
        job.tile_w, job.tile_h = self.tile
        job.rank, job.world = int(rank), int(world)
        job.direct = int(direct) if not isinstance(direct, bool) else (1 if direct else 0)    # 2 = _abi.DIRECT_SHARED_FRAME
        job.mesh_mode = int(mesh_mode) | int(getattr(self, "debug_mode_bits", 0))
        job.max_batch_rays = int(max_batch_rays or getattr(self, "max_batch_rays", 0))
        return job

    def render(self, subimage: int = 0, tasks: int = 1) -> np.ndarray:
        width = self.strip_columns(subimage, tasks)
        out32 = np.empty((len(width), self.vc.height, 3), dtype=np.float32)
        if len(width) > 0:
            self.render_region(int(width[0]), len(width), 0, self.vc.height, out=out32)
        return out32.astype(np.float64)

    def render_region(self, col0, ncols, row0, nrows, out=None, mesh_mode=0, stream=None):
        """Render a pixel rectangle to a host float32 array (ncols, nrows, 3) through
        `rt_render_host` (device staging + D2H inside the call)."""
        h = self.upload(stream)
        if out is None:
            out = np.empty((ncols, nrows, 3), dtype=np.float32)
        assert out.dtype == np.float32 and out.flags["C_CONTIGUOUS"] and out.shape == (ncols, nrows, 3)
        job = self.make_job(col0, ncols, row0, nrows, mesh_mode=mesh_mode)
        ctr = _abi.RtCounters()
        rc = self._lib.rt_render_host(h, C.byref(job), out.ctypes.data, C.byref(ctr), stream)
        _abi.check(self._lib, rc, "rt_render_host")
        self.last_counters = ctr.as_dict()
        self._raise_flags(ctr)
        return out

    def _raise_flags(self, ctr):
        if ctr.error_flags & (1 << 4):
            raise IndexError("list index out of range (a shaded hit had no material; geometry/__init__.py:57)")
        if ctr.error_flags & (1 << 5):
            raise _abi.B200RTError("RT_ERR_LIMIT: the scene exceeds a compiled limit (traversal stack or CSG nesting depth)")

    def render_u8(self, subimage: int = 0, tasks: int = 1, stream=None) -> np.ndarray:
        """`render()` followed by main.py:31-33 (rot90, *255 truncated to uint8), converted ON THE DEVICE
        (`rt_render_u8_host`): returns the (H, W_sub, 3) uint8 array `PIL.Image.fromarray` takes."""
        width = self.strip_columns(subimage, tasks)
        out = np.empty((self.vc.height, len(width), 3), dtype=np.uint8)
        if len(width) == 0:
            return out
        h = self.upload(stream)
        job = self.make_job(int(width[0]), len(width), 0, self.vc.height)
        ctr = _abi.RtCounters()
        rc = self._lib.rt_render_u8_host(h, C.byref(job), out.ctypes.data, C.byref(ctr), stream)
        _abi.check(self._lib, rc, "rt_render_u8_host")
        self.last_counters = ctr.as_dict()
        self._raise_flags(ctr)
        return out

    def render_tensor(self, subimage: int = 0, tasks: int = 1):
        """`render()` that stays on the GPU: float32 CUDA tensor (W_sub, H, 3) in the reference orientation, produced by
        `torch.ops.b200rt.render` on the current torch stream (no host copy, no synchronisation)."""
        from . import torch_ext
        ops = torch_ext.load()
        width = self.strip_columns(subimage, tasks)
        if len(width) == 0:
            import torch
            return torch.empty((0, self.vc.height, 3), dtype=torch.float32, device="cuda")
        job = self.make_job(int(width[0]), len(width), 0, self.vc.height)
        return ops.render(torch_ext.scene_handle(self), torch_ext.job_tensor(job))

    def render_device(self, d_out_ptr, job, counters=None, stream=None):
        """Raw device-buffer call (`rt_render`) for callers that own device memory (bench, multi-GPU)."""
        h = self.upload(stream)
        rc = self._lib.rt_render(h, C.byref(job), d_out_ptr, None if counters is None else C.byref(counters), stream)
        _abi.check(self._lib, rc, "rt_render")

    def render_device_static(self, d_out_ptr, job, counters=None, stream=None):
        """`rt_render_static`: the same render with its scratch in the library's static device array - the call allocates
        nothing (callers under CUDA graph capture / serving loops); limits RT_STATIC_MAX_RAYS samples per pixel, RT_STATIC_MAX_LIGHTS."""
        h = self.upload(stream)
        rc = self._lib.rt_render_static(h, C.byref(job), d_out_ptr, None if counters is None else C.byref(counters), stream)
        _abi.check(self._lib, rc, "rt_render_static")

    # ------------------------------------------------------------------ debugging / unit-test aids
    def trace_rays(self, origins, dirs, times=None, mesh_mode=0):
        """cast_ray (scene.py:81) for explicit rays: returns dict(rgb, t, n, p, geom, mat)."""
        h = self.upload()
        o = np.ascontiguousarray(origins, dtype=np.float32).reshape(-1, 3)
        d = np.ascontiguousarray(dirs, dtype=np.float32).reshape(-1, 3)
        n = o.shape[0]
        t = np.zeros(n, dtype=np.float32) if times is None else np.ascontiguousarray(times, dtype=np.float32)
        res = dict(rgb=np.empty((n, 3), np.float32), t=np.empty(n, np.float32), n=np.empty((n, 3), np.float32),
                   p=np.empty((n, 3), np.float32), geom=np.empty(n, np.int32), mat=np.empty(n, np.int32))
        rc = self._lib.rt_trace_rays(h, n, o.ctypes.data, d.ctypes.data, t.ctypes.data, res["rgb"].ctypes.data,
                                     res["t"].ctypes.data, res["n"].ctypes.data, res["p"].ctypes.data,
                                     res["geom"].ctypes.data, res["mat"].ctypes.data, mesh_mode, None)
        _abi.check(self._lib, rc, "rt_trace_rays")
        return res

    def occluded(self, origins, dirs, tmax, times=None, mesh_mode=0):
        """The any-hit loop of scene.py:161-164 for explicit shadow rays."""
        h = self.upload()
        o = np.ascontiguousarray(origins, dtype=np.float32).reshape(-1, 3)
        d = np.ascontiguousarray(dirs, dtype=np.float32).reshape(-1, 3)
        n = o.shape[0]
        t = np.zeros(n, dtype=np.float32) if times is None else np.ascontiguousarray(times, dtype=np.float32)
        tm = np.ascontiguousarray(np.broadcast_to(np.asarray(tmax, dtype=np.float32), (n,)))
        occ = np.empty(n, np.int32)
        rc = self._lib.rt_occluded(h, n, o.ctypes.data, d.ctypes.data, t.ctypes.data, tm.ctypes.data,
                                   occ.ctypes.data, mesh_mode, None)
        _abi.check(self._lib, rc, "rt_occluded")
        return occ


def to_png_array(image):
    """main.py:31-33 / glue.py:23-25: rot90 then (image*255) truncated to uint8."""
    return (np.rot90(image, k=1, axes=(0, 1)) * 255).astype(np.uint8)
