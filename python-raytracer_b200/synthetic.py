"""Deterministic synthetic benchmark scene (BASELINE.json configs[4], SURVEY.md 8d "Config 5"):
a displaced torus tessellated as a grid[0] x grid[1] quad grid (1024 x 512 -> 1 048 576 triangles,
bounding radius ~2 around (0, 1.5, 0), smooth shaded, diffuse), 64 spheres of radius U(0.1, 0.3) on a
ring around it with materials cycling diffuse / mirror (tint 0.2) / refractive (1.458, tint 0.2), a
checker ground plane y = -1, two point lights and one directional light, camera (0,4,8) -> (0,1,0),
fov 45.  Built through the same classes `scene_parser.load_scene` produces."""
import math

import numpy as np

from . import geometry as geom
from . import helperclasses as hc
from .scene import Scene


def displaced_torus(nu, nv, seed=1234):
    """(verts float64 (nu*nv,3), faces int32 (2*nu*nv,3)); closed surface, no degenerate triangles."""
    rng = np.random.default_rng(seed)
    ph = rng.uniform(0, 2 * math.pi, size=4)
    u = (np.arange(nu) / nu) * 2 * math.pi
    v = (np.arange(nv) / nv) * 2 * math.pi
    U, V = np.meshgrid(u, v, indexing="ij")
    R, r0 = 1.35, 0.5
    r = r0 * (1.0 + 0.12 * np.sin(7 * U + ph[0]) * np.sin(5 * V + ph[1]) + 0.05 * np.sin(23 * U + 11 * V + ph[2])
              + 0.02 * np.sin(61 * V - 37 * U + ph[3]))
    x = (R + r * np.cos(V)) * np.cos(U)
    z = (R + r * np.cos(V)) * np.sin(U)
    y = r * np.sin(V)
    tilt = math.radians(35.0)                      # tip the ring towards the camera
    y2 = y * math.cos(tilt) - z * math.sin(tilt)
    z2 = y * math.sin(tilt) + z * math.cos(tilt)
    verts = np.stack([x, y2 + 1.5, z2], axis=-1).reshape(-1, 3)
    i = np.arange(nu)[:, None]
    j = np.arange(nv)[None, :]
    a = (i * nv + j)
    b = (((i + 1) % nu) * nv + j)
    c = (((i + 1) % nu) * nv + (j + 1) % nv)
    d = (i * nv + (j + 1) % nv)
    faces = np.concatenate([np.stack([a, c, b], axis=-1).reshape(-1, 3), np.stack([a, d, c], axis=-1).reshape(-1, 3)])
    return verts, faces.astype(np.int32)


def make_scene(width=3840, height=2160, aa_samples=64, grid=(1024, 512), n_spheres=64, jitter=True, seed=1234,
               dof_samples=1, flat_shaded=False, ingest=None):
    rng = np.random.default_rng(seed)
    vc = hc.ViewportCamera().set_viewport(width, height) \
        .set_camera((0.0, 4.0, 8.0), (0.0, 1.0, 0.0), (0.0, 1.0, 0.0), 45.0) \
        .set_lens(1.0, 0.0, dof_samples).set_motion(0, 1, 0)
    M = hc.Material
    mats = [M("mesh", (0.4, 0.4, 0.4), (0.8, 0.45, 0.2), 32, 0),
            M("tile_a", (0, 0, 0), (0.9, 0.9, 0.9), 0, 1),
            M("tile_b", (0, 0, 0), (0.15, 0.15, 0.2), 0, 2),
            M("ball", (0.5, 0.5, 0.5), (0.2, 0.5, 0.9), 50, 3),
            M("mirror", (0.8, 0.8, 0.8), (0.9, 0.9, 0.9), 64, 4, "mirror", 0.2),
            M("glass", (0.8, 0.8, 0.8), (0.7, 0.9, 0.8), 64, 5, "refractive", 0.2)]
    mats[5].refr_index = 1.458
    objects = [geom.Plane("ground", "plane", [mats[1], mats[2]], (0.0, -1.0, 0.0), (0.0, 1.0, 0.0), None)]
    verts, faces = displaced_torus(grid[0], grid[1], seed)
    objects.append(geom.Mesh("torus", "mesh", [mats[0]], (0.0, 0.0, 0.0), 1.0, (verts, faces), flat_shaded, None, ingest=ingest))
    for k in range(n_spheres):
        ang = 2 * math.pi * k / max(1, n_spheres)
        rad = float(rng.uniform(0.1, 0.3))
        ring = 2.9 + 0.5 * math.sin(3 * ang)
        centre = (ring * math.cos(ang), -1.0 + rad + float(rng.uniform(0.0, 0.6)), ring * math.sin(ang))
        objects.append(geom.Sphere(f"s{k}", "sphere", [mats[3 + k % 3]], centre, rad, None))
    lights = [hc.Light("point", "key", (1.0, 0.95, 0.9), (6.0, 9.0, 6.0), 0.7),
              hc.Light("point", "fill", (0.6, 0.7, 1.0), (-7.0, 5.0, 3.0), 0.4),
              hc.Light("directional", "sun", (1.0, 1.0, 1.0), (0.3, -1.0, -0.4), 1.0)]
    lights[2].power = 0.3
    sc = Scene(vc, jitter, aa_samples, (0.1, 0.1, 0.1), lights, mats, objects)
    sc.jitter_seed = seed
    for o in objects:
        o.set_scene(sc)
    return sc
