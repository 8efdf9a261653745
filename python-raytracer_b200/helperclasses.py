"""Scene data model: host-side mirror of the reference's `provided/helperclasses.py`.

Same class names, constructor arguments and attribute names as the reference
(`Ray :13`, `Material :28`, `Light :50`, `ViewportCamera :69`), holding numpy float32
vectors instead of PyGLM ones.  These are plain records: they are inputs to the flattening
step (`_abi.flatten`) that hands SoA buffers to the CUDA kernels; no ray maths runs here.
"""
import math

import numpy as np

from . import _vecmath as vm


class Ray:
    """helperclasses.py:13-25"""

    def __init__(self, o, d):
        self.origin = vm.vec3(o)
        self.direction = vm.vec3(d)

    def getDistance(self, point):
        return float(vm.length(vm.vec3(point) - self.origin))

    def getPoint(self, t: float):
        return self.origin + self.direction * np.float32(t)

    def __repr__(self):
        return f"Ray(origin: {self.origin}, dir: {self.direction})"


class Material:
    """helperclasses.py:28-47"""

    def __init__(self, name: str, specular, diffuse, hardness: float, ID: int,
                 mat_type: str = "diffuse", mat_tint: float = 0.0):
        self.name = name
        self.specular = vm.vec3(specular)
        self.diffuse = vm.vec3(diffuse)
        self.hardness = hardness
        self.ID = ID
        self.mat_type = mat_type
        self.refr_index = 1.0
        self.tint = mat_tint

    @staticmethod
    def default():
        return Material("default", (0, 0, 0), (0, 0, 0), -1, -1)

    def __repr__(self):
        return (f"Material({self.name}, type: {self.mat_type}, specular: {self.specular}, "
                f"diffuse: {self.diffuse}, hardness: {self.hardness}, ID: {self.ID})")


class Light:
    """helperclasses.py:50-59"""

    def __init__(self, ltype: str, name: str, colour, vector, power: float):
        self.type = ltype
        self.name = name
        self.colour = vm.vec3(colour)
        self.vector = vm.vec3(vector)
        self.power = power

    def __repr__(self):
        return f"Light({self.name}, type: {self.type}, colour: {self.colour}, vector: {self.vector}, power: {self.power})"


class ViewportCamera:
    """helperclasses.py:69-108 (same fluent setters)."""

    def __init__(self):
        self.focal_length = 1.0
        self.aperture = 0.0
        self.dof_samples = 1
        self.motion_times = [0.0]

    def set_viewport(self, width: int, height: int) -> "ViewportCamera":
        self.width = width
        self.height = height
        self.aspect = width / height
        return self

    def set_camera(self, position, lookat, up, fov: float) -> "ViewportCamera":
        position, lookat, up = vm.vec3(position), vm.vec3(lookat), vm.vec3(up)
        cam_dir = position - lookat
        self.position = position
        self.d = 1.0
        self.top = self.d * math.tan(math.radians(fov / 2))
        self.right = self.aspect * self.top
        self.bottom = -self.top
        self.left = -self.right
        self.w = vm.normalize(cam_dir)
        self.u = vm.normalize(vm.cross(up, self.w))
        self.v = vm.cross(self.w, self.u)
        return self

    def set_lens(self, focal_length: float, aperture: float, dof_samples: int) -> "ViewportCamera":
        self.focal_length = focal_length
        self.aperture = aperture
        self.dof_samples = dof_samples
        return self

    def set_motion(self, time: float, motion_samples: int, motion_final: int) -> "ViewportCamera":
        dt = time / motion_samples
        self.motion_times = [dt * i for i in range(motion_samples)]
        self.motion_times += [time] * motion_final
        return self
