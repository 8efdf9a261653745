"""Geometry records: host-side mirror of the reference's `provided/geometry/` package.

Same class names and constructor arguments as the reference (`Sphere`, `Plane`, `AABB`
simple_geometry.py:12,86,179; `Mesh` mesh.py:16; `Hierarchy` hierarchy.py:11;
`BoundingSphere`/`BoundingAABB` bounding_volumes.py:12,43).  They carry only the state the
kernels need (plus the one-off set-up the reference performs in its constructors: plane
texture axes, box corners, mesh transform / vertex normals / bounding-volume choice, node
matrices).  The per-ray methods of the reference (`intersect`, `shadow_intersect`,
`is_inside`, `get_material`, `get_diffuse`) are NOT implemented on the host: they are the
hot path and live in `csrc/` as CUDA.
"""
import math

import numpy as np

from . import _vecmath as vm

epsilon = 10 ** (-4)  # geometry/__init__.py:12


class Geometry:
    shadow_epsilon = 10 ** (-4)  # geometry/__init__.py:39

    def __init__(self, name: str, gtype: str, materials: list, speed):
        self.name = name
        self.gtype = gtype
        self.materials = materials
        self.speed = None if speed is None else vm.vec3(speed)

    def set_scene(self, scene):
        self.scene = scene

    def __repr__(self):
        return f"Geometry({self.name}, type: {self.gtype})"


class Sphere(Geometry):
    shadow_epsilon = 10 ** (-3)  # simple_geometry.py:13

    def __init__(self, name, gtype, materials, center, radius: float, speed):
        super().__init__(name, gtype, materials, speed)
        self.center = vm.vec3(center)
        self.radius = radius

    def __repr__(self):
        return f"Sphere({self.name}, center: {self.center}, radius: {self.radius})"


def _eq(a, b):
    return bool(a[0] == b[0] and a[1] == b[1] and a[2] == b[2])


class Plane(Geometry):
    def __init__(self, name, gtype, materials, point, normal, speed):
        super().__init__(name, gtype, materials, speed)
        self.point = vm.vec3(point)
        self.normal = vm.vec3(normal)
        self.texture = None
        self.texture_scale = 1.0
        n = self.normal
        # simple_geometry.py:93-103
        if _eq(n, (0, 1, 0)) or _eq(n, (0, -1, 0)) or _eq(n, (0, 0, 1)):
            self.width_axis = vm.vec3((1, 0, 0))
        elif _eq(n, (0, 0, -1)):
            self.width_axis = vm.vec3((-1, 0, 0))
        elif _eq(n, (1, 0, 0)):
            self.width_axis = vm.vec3((0, 0, -1))
        elif _eq(n, (-1, 0, 0)):
            self.width_axis = vm.vec3((0, 0, 1))
        else:
            self.width_axis = vm.normalize(vm.cross(n, vm.vec3((0, 0, 1))))
        self.height_axis = vm.normalize(vm.cross(self.width_axis, n))

    def __repr__(self):
        return f"Plane({self.name}, point: {self.point}, normal: {self.normal})"


class AABB(Geometry):
    def __init__(self, name, gtype, materials, center, dimension, speed):
        super().__init__(name, gtype, materials, speed)
        center, dimension = vm.vec3(center), vm.vec3(dimension)
        halfside = (dimension / np.float32(2)).astype(np.float32)
        self.minpos = center - halfside
        self.maxpos = center + halfside
        self.texture = None

    def __repr__(self):
        return f"AABB({self.name}, minpos: {self.minpos}, maxpos: {self.maxpos})"


class BoundingSphere:
    """bounding_volumes.py:12-16"""

    def __init__(self, center, radius: float, geometry=None):
        self.center = vm.vec3(center)
        self.radius = radius
        self.geometry = geometry


class BoundingAABB:
    """bounding_volumes.py:43-47"""

    def __init__(self, minpos, maxpos, geometry=None):
        self.minpos = vm.vec3(minpos)
        self.maxpos = vm.vec3(maxpos)


def read_obj(filepath):
    """`v` / `vn` / `f` rows of a Wavefront OBJ (what mesh.py:20 takes from igl.read_obj);
    polygons are fan-triangulated, indices made 0-based."""
    verts, norms, faces = [], [], []
    with open(filepath) as fh:
        for line in fh:
            p = line.split()
            if not p:
                continue
            if p[0] == "v":
                verts.append((float(p[1]), float(p[2]), float(p[3])))
            elif p[0] == "vn":
                norms.append((float(p[1]), float(p[2]), float(p[3])))
            elif p[0] == "f":
                idx = []
                for tok in p[1:]:
                    i = int(tok.split("/")[0])
                    idx.append(i - 1 if i > 0 else len(verts) + i)
                for k in range(1, len(idx) - 1):
                    faces.append((idx[0], idx[k], idx[k + 1]))
    return (np.array(verts, dtype=np.float64).reshape(-1, 3),
            np.array(norms, dtype=np.float64).reshape(-1, 3),
            np.array(faces, dtype=np.int32).reshape(-1, 3))


class Mesh(Geometry):
    """mesh.py:16-51.  `filepath` may also be a (verts, faces) tuple of arrays for
    procedurally generated meshes (the 1M-triangle benchmark scene).

    `ingest` picks where the load-time arithmetic runs (an explicit choice, never a fallback): "host" = the numpy
    restatement below, mirroring the reference's loops; "device" = the native OBJ reader plus `rt_mesh_ingest`
    (CUDA, bit-identical results, for meshes of 10^5..10^7 faces where the Python-side loops dominate the set-up)."""

    DEFAULT_INGEST = "host"

    def __init__(self, name, gtype, materials, translate, scale: float, filepath,
                 flat_shaded: bool = False, speed=None, ingest=None):
        super().__init__(name, gtype, materials, speed)
        ingest = ingest or Mesh.DEFAULT_INGEST
        if ingest not in ("host", "device"):
            raise ValueError(f"ingest must be 'host' or 'device', not {ingest!r}")
        if isinstance(filepath, (tuple, list)):
            verts, faces = filepath
            verts = np.asarray(verts, dtype=np.float64).reshape(-1, 3)
            self.faces = np.ascontiguousarray(faces, dtype=np.int32).reshape(-1, 3)
        elif ingest == "device":
            from . import _abi
            verts, _norms, self.faces = _abi.obj_load(filepath)
        else:
            verts, _norms, self.faces = read_obj(filepath)
        self.flat_shaded = flat_shaded
        self.norms = None
        t = vm.vec3(translate)
        if ingest == "device":
            from . import _abi
            self.verts, self.norms, mn32, mx32, center, max_dist = _abi.mesh_ingest(verts, self.faces, t, scale, not flat_shaded)
            mn, mx = mn32.astype(np.float64), mx32.astype(np.float64)
        else:
            # (vec3(v) + translate) * scale, float32 (mesh.py:24)
            self.verts = np.ascontiguousarray((verts.astype(np.float32) + t) * np.float32(scale), dtype=np.float32)
            if not flat_shaded:
                self._compute_normals()
            v = self.verts
            mx = v.max(axis=0).astype(np.float64)
            mn = v.min(axis=0).astype(np.float64)
            center = ((mx + mn) / 2).astype(np.float32)                      # mesh.py:38-41
            max_dist = float(np.sqrt(vm.dot_rows(v - center, v - center)).max())  # mesh.py:43
        ext = mx - mn
        aabb_volume = ext[0] * ext[1] * ext[2]
        sphere_volume = 4 / 3 * math.pi * max_dist ** 3
        if aabb_volume < sphere_volume:                                  # mesh.py:48-51
            self.bounding_volume = BoundingAABB(mn, mx, None)
        else:
            self.bounding_volume = BoundingSphere(center, max_dist, None)

    SHARE_ON_COPY = True

    def __deepcopy__(self, memo):
        """`ref` nodes are `copy.deepcopy` of a whole subtree in the reference (scene_parser.py:199), which duplicates every
        vertex of every mesh below it.  Here a copy is an INSTANCE: its own name / materials / speed (they change per copy:
        set_fallback_material, hierarchy.py:21-28), but the SAME vertex, face and normal arrays and bounding volume (read-only
        after construction), so `_abi.flatten` hands the device one RtMesh - one LBVH - for all copies.  Results are identical
        to duplicating the data (tests/test_host_logic.py)."""
        import copy
        if not Mesh.SHARE_ON_COPY:
            cls = self.__class__
            dup = cls.__new__(cls)
            memo[id(self)] = dup
            for k, v in self.__dict__.items():
                setattr(dup, k, copy.deepcopy(v, memo))
            return dup
        cls = self.__class__
        inst = cls.__new__(cls)
        memo[id(self)] = inst
        for k, v in self.__dict__.items():
            if k in ("verts", "faces", "norms", "bounding_volume", "scene"):
                setattr(inst, k, v)
            else:
                setattr(inst, k, copy.deepcopy(v, memo))
        return inst

    def _compute_normals(self):
        """Area-weighted vertex normals, mesh.py:53-70, accumulated per vertex in face order."""
        v, f = self.verts, self.faces
        e1 = v[f[:, 1]] - v[f[:, 0]]
        e2 = v[f[:, 2]] - v[f[:, 0]]
        c = vm.cross_rows(e1, e2)
        normal = vm.normalize_rows(c)
        area = (np.sqrt(vm.dot_rows(c, c)).astype(np.float64) / 2).astype(np.float32)
        weighted = (normal * area[:, None]).astype(np.float32)
        acc = np.zeros_like(v)
        np.add.at(acc, f.reshape(-1), np.repeat(weighted, 3, axis=0))
        self.norms = np.ascontiguousarray(vm.normalize_rows(acc), dtype=np.float32)

    def __repr__(self):
        return f"Mesh({self.name})"


class Hierarchy(Geometry):
    """hierarchy.py:11-37 (transform node + CSG operator)."""

    def __init__(self, name, gtype, materials, hierarchy_type: str, t, r, s, speed):
        super().__init__(name, gtype, materials, speed)
        self.hierarchy_type = hierarchy_type
        self.children = []
        self.make_matrices(t, r, s)

    def set_fallback_material(self, materials: list):
        """hierarchy.py:21-28: APPENDS the root's materials to every leaf's list."""
        if len(materials) == 0:
            return
        for child in self.children:
            if isinstance(child, Hierarchy):
                child.set_fallback_material(materials)
            else:
                child.materials += materials

    def make_matrices(self, t, r, s):
        self.t, self.r, self.s = vm.vec3(t), vm.vec3(r), vm.vec3(s)
        self.M, self.Minv = vm.trs(self.t, self.r, self.s)

    def set_scene(self, scene):
        self.scene = scene
        for child in self.children:
            child.set_scene(scene)

    def __repr__(self):
        return f"Hierarchy({self.name}, t: {self.t}, r: {self.r}, s: {self.s}, children: {len(self.children)})"
