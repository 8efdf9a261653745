import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests", "hostsim")):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pytest_collection_modifyitems(config, items):
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


def _merge(data, overrides):
    for k, v in (overrides or {}).items():
        if isinstance(v, dict) and isinstance(data.get(k), dict):
            data[k] = {**data[k], **v}
        else:
            data[k] = v
    return data


def manifest():
    with open(os.path.join(GOLDEN, "MANIFEST.json")) as fh:
        return json.load(fh)


def load_case(case, extra_overrides=None, ingest=None):
    """Scene of a golden case (tests/golden/MANIFEST.json), through the package's own parser."""
    import python_raytracer_b200 as rt
    rt.scene_parser.VERBOSE = False
    e = manifest()[case]
    with open(os.path.join(GOLDEN, "scenes", e["scene"])) as fh:
        data = _merge(json.load(fh), e["overrides"])
    data = _merge(data, extra_overrides)
    sc = rt.scene_parser.load_scene_dict(data, os.path.join(GOLDEN, "scenes"), ingest=ingest)
    sc.jitter_seed = e["seed"]
    return sc


def golden_png(case):
    from PIL import Image
    return np.array(Image.open(os.path.join(GOLDEN, "renders", case + ".png")).convert("RGB"))


def golden_f32(case):
    z = np.load(os.path.join(GOLDEN, "renders", case + "_f32.npz"))
    return z["cols"], z["rows"], z["rgb"]


def to_png_array(image):
    return (np.rot90(np.asarray(image, dtype=np.float64), k=1, axes=(0, 1)) * 255).astype(np.uint8)


def parity_stats(a_u8, b_u8):
    d = np.abs(a_u8.astype(int) - b_u8.astype(int)).max(axis=2)
    return float((d <= 1).mean()), int(d.max()), d


# BASELINE.json north_star: per-channel |delta| <= 1/255 on >= 99.9 % of pixels
PARITY_FRACTION = 0.999

ALL_CASES = ["TwoSpheresPlane", "MirrorRefraction", "MotionBlur", "TorusMesh_flat", "TorusMesh_smooth",
             "DepthOfField", "BoxTexture", "MeshSphereGate", "MeshInNode", "NovelScene1_small", "NovelScene2_small",
             "Fuzz1", "Fuzz8", "Fuzz12", "Fuzz18", "Fuzz19"]     # Fuzz<seed>: tests/fuzz_scenes.py rendered by the reference


# Fresh random scenes (tests/fuzz_scenes.py) compared with the oracle at test time.
FUZZ_GPU_SEEDS = [100, 101, 102, 103, 104, 105, 106, 107, 108, 110, 111, 112, 113, 115, 116, 117]
FUZZ_HOST_SEEDS = [118, 119, 120, 122, 123, 124]
# Seeds of 0..149 on which a plain >= 99.9 % threshold fails or nearly fails (round 1 left them out): the reference's CSG
# semantics are chaotic on some surfaces (e.g. the exit face of a box seen through a carved hole: Hierarchy.shadow_intersect of a
# difference goes through AABB.intersect, which returns no hit at all when the un-offset shadow origin lies 1e-7 inside the box
# and two hits when it lies 1e-7 outside, hierarchy.py:94-107 + simple_geometry.py:223), so there the shadow state is a coin
# toss in the reference itself.  They are TESTED, not hidden: `residue_report` below classifies every out-of-tolerance pixel with
# the oracle's stability probe and the tests require the statistic on the reference-stable pixels.
FUZZ_CHAOTIC_SEEDS = [10, 15, 25, 26, 28, 109, 121, 130, 132, 134, 140, 148, 149]


def u8_cr(image):
    """(cols, rows, 3) float image -> the uint8 values main.py:33 writes, still indexed [col, row]."""
    return (np.asarray(image, dtype=np.float64) * 255).astype(np.uint8).astype(int)


def residue_report(flat, ref, got, eps=1e-6, trials=6):
    """WHERE the pixels are that differ from the reference by more than 1/255 (SURVEY 8d "parity statistic", north_star
    "residue confined to fp32-vs-fp64 edge and refraction pixels").  `ref` = oracle image, `got` = image under test, both
    (W, H, 3).  Every residue pixel is
      * `unstable`: the ORACLE's own uint8 value moves by more than 1 when its rays and hit points are perturbed by `eps`
        relative (a few float32 ulp; oracle.unstable_pixels) - a reference-chaotic pixel;
      * `edge`: the first sample's (leaf object, material, shadowed-light bits) differs from a 3x3 neighbour's: silhouette,
        checker / material edge, shadow edge;
      * `specular`: it or a neighbour sees a mirror / refractive material (reflection and refraction regions);
    and `unexplained` counts the ones that are none of these.  `frac_stable` is the parity statistic over the
    reference-stable pixels: 1 - (#residue pixels that are not unstable) / (#pixels)."""
    import oracle
    d = np.abs(u8_cr(ref) - u8_cr(got)).max(axis=2)
    W, H = d.shape
    bad = np.argwhere(d > 1)
    rep = {"pixels": int(W * H), "residue": int(len(bad)), "frac": float((d <= 1).mean()), "max_delta": int(d.max()),
           "unstable": 0, "edge": 0, "specular": 0, "unexplained": 0, "frac_stable": 1.0}
    if len(bad) == 0:
        return rep
    cols, rows = bad[:, 0], bad[:, 1]
    un = oracle.unstable_pixels(flat, cols, rows, np.asarray(ref)[cols, rows], eps=eps, trials=trials)
    f0 = oracle.pixel_features(flat, cols, rows)
    edge, spec = np.zeros(len(bad), bool), f0[:, 3] != 0
    for dc in (-1, 0, 1):
        for dr in (-1, 0, 1):
            f = oracle.pixel_features(flat, np.clip(cols + dc, 0, W - 1), np.clip(rows + dr, 0, H - 1))
            edge |= (f[:, :3] != f0[:, :3]).any(axis=1)
            spec |= f[:, 3] != 0
    rep.update(unstable=int(un.sum()), edge=int(edge.sum()), specular=int(spec.sum()),
               unexplained=int((~un & ~edge & ~spec).sum()), frac_stable=float(1.0 - (~un).sum() / (W * H)))
    return rep


def unstable_fraction(flat, ref, n=2000, seed=0):
    """Share of reference-unstable pixels in a random sample of the frame (keeps the probe honest: it must not call
    everything unstable)."""
    import oracle
    W, H = np.asarray(ref).shape[:2]
    rng = np.random.default_rng(seed)
    cols, rows = rng.integers(0, W, n), rng.integers(0, H, n)
    return float(oracle.unstable_pixels(flat, cols, rows, np.asarray(ref)[cols, rows]).mean())


def load_bundled(name, overrides=None):
    """A bundled scene (tests/golden/scenes copy of the reference's scenes/<name>.json) at its SHIPPED settings."""
    import python_raytracer_b200 as rt
    rt.scene_parser.VERBOSE = False
    with open(os.path.join(GOLDEN, "scenes", name + ".json")) as fh:
        data = _merge(json.load(fh), overrides)
    sc = rt.scene_parser.load_scene_dict(data, os.path.join(GOLDEN, "scenes"))
    sc.jitter_seed = 0x5EED
    return sc


def load_fuzz(seed, width=128, height=96):
    import python_raytracer_b200 as rt
    import fuzz_scenes
    rt.scene_parser.VERBOSE = False
    sc = rt.scene_parser.load_scene_dict(fuzz_scenes.make(seed, width, height), os.path.join(GOLDEN, "scenes"))
    sc.jitter_seed = 0x5EED
    return sc


def morton63_keys(verts_f32, faces):
    """numpy restatement of csrc/rt_bvh.cuh morton63 over triangle centroids (float32 arithmetic in the same order)."""
    v = np.asarray(verts_f32, dtype=np.float32)
    lo, hi = v.min(0), v.max(0)
    ext = hi - lo
    inv = np.where(ext > 0, np.float32(1) / np.where(ext > 0, ext, 1).astype(np.float32), 0).astype(np.float32)
    a, b, c = v[faces[:, 0]], v[faces[:, 1]], v[faces[:, 2]]
    cen = ((a + b) + c) * np.float32(1.0 / 3.0)
    x = np.clip((cen - lo) * inv, 0, 1).astype(np.float32)
    q = np.minimum(x * np.float32(2097152.0), np.float32(2097151.0)).astype(np.uint64)

    def expand(u):
        u = u & np.uint64(0x1fffff)
        for sh, mask in ((32, 0x1f00000000ffff), (16, 0x1f0000ff0000ff), (8, 0x100f00f00f00f00f), (4, 0x10c30c30c30c30c3), (2, 0x1249249249249249)):
            u = (u | u << np.uint64(sh)) & np.uint64(mask)
        return u
    return (expand(q[:, 0]) << np.uint64(2)) | (expand(q[:, 1]) << np.uint64(1)) | expand(q[:, 2])


def many_meshes_scene(n=11, resolution=(96, 72)):
    """TorusMesh with `n` top-level torus meshes (flat and smooth alternating) on a ring: more than the RT_CONST_MESH = 8 headers the
    lean kernels keep in the parameter bank, so the scene runs on the general kernels (rt_scene_build.hpp convert_scene)."""
    import math
    objs = [{"materials": [1, 0], "name": "plane", "normal": [0.0, 1.0, 0.0], "position": [0.0, -1.0, 0.0], "type": "plane"}]
    for k in range(n):
        a = 2 * math.pi * k / n
        objs.append({"filepath": "./meshes/torus.obj", "flat_shaded": k % 2 == 0, "materials": [2], "name": f"torus{k}",
                     "position": [1.6 * math.cos(a), 0.3 * math.sin(3 * a), 1.6 * math.sin(a)], "scale": 0.35, "type": "mesh"})
    return load_case("TorusMesh_flat", {"objects": objs, "resolution": list(resolution)})
