"""Generate tests/golden/ from the UNMODIFIED reference (tier-0 oracle).  Dev container only.

    python oracle/make_golden.py [--only NAME ...]

Inputs are read from /root/reference (scenes, torus.obj, textures) and re-serialised into
tests/golden/{scenes,meshes,textures}/ so that the GPU box (which has no /root/reference) can
load the same scenes; outputs are the reference's own renders of those inputs, executed
through oracle/shims by oracle/ref_runner.py:

    tests/golden/renders/<case>.png       (H,W,3) uint8, exactly main.py:31-33's PNG
    tests/golden/renders/<case>_f32.npz   every 4th column/row of the float image (values are
                                          float32-exact: scene.py:73 divides a float32 vec3)
    tests/golden/MANIFEST.json            per case: scene file, overrides, seed, and whether the
                                          render is bit-identical to the author's renders/*.png

Jittered cases draw from Philox4x32-10 (oracle/philox.py) instead of the unseeded global
numpy RNG (scene.py:64), so they are reproducible; the author's jittered PNGs are only
statistically comparable and are not used.
"""
import argparse
import json
import os
import sys
import time

import numpy as np
from PIL import Image

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import ref_runner  # noqa: E402

REF = ref_runner.REFERENCE
OUT = os.path.join(HERE, "..", "tests", "golden")

SCENES = {  # golden scene name -> reference file
    "TwoSpheresPlane": "scenes/TwoSpheresPlane.json",
    "MirrorRefraction": "scenes/MirrorRefraction.json",
    "MotionBlur": "scenes/MotionBlur.json",
    "TorusMesh": "test_scenes/TorusMesh.json",   # scenes/TorusMesh.json points at a bunny.obj that is not shipped
    "DepthOfField": "scenes/DepthOfField.json",
    "NovelScene1": "scenes/NovelScene1.json",
    "NovelScene2": "scenes/NovelScene2.json",
}


def _torus_smooth():
    with open(os.path.join(REF, SCENES["TorusMesh"])) as fh:
        objs = json.load(fh)["objects"]
    for o in objs:
        if o["type"] == "mesh":
            o["flat_shaded"] = False
    return {"objects": objs}


# AABB.get_diffuse with a texture (simple_geometry.py:312-355) is used by none of the shipped scenes: this case puts the
# shipped wall textures on boxes seen from outside on five of their faces (camera, and a mirror sphere behind them)
BOX_TEXTURE = {
    "resolution": [200, 150], "AA": {"jitter": False, "samples": 1},
    "materials": [{"name": "blue", "ID": 0, "diffuse": [0.0, 0.0, 1.0], "specular": [0.5, 0.5, 0.5], "hardness": 16},
                  {"name": "mirror", "ID": 1, "diffuse": [0.1, 0.1, 0.1], "specular": [0.3, 0.3, 0.3], "hardness": 50,
                   "type": "mirror", "tint": 0.85},
                  {"name": "black", "ID": 2, "diffuse": [0.0, 0.0, 0.0], "specular": [0.0, 0.0, 0.0], "hardness": 0},
                  {"name": "white", "ID": 3, "diffuse": [1.0, 1.0, 1.0], "specular": [0.0, 0.0, 0.0], "hardness": 0}],
    "objects": [{"name": "plane", "type": "plane", "normal": [0.0, 1.0, 0.0], "position": [0.0, 0.0, 0.0], "materials": [2, 3]},
                {"name": "crate_l", "type": "box", "size": [1.6, 1.6, 1.6], "position": [-1.7, 0.8, 0.2], "materials": [0],
                 "texture": "textures/wall1.png"},
                {"name": "crate_r", "type": "box", "min": [1.0, 0.0, -0.6], "max": [2.4, 2.2, 0.9], "materials": [0],
                 "texture": "textures/wall3.png"},
                {"name": "lid", "type": "box", "size": [1.0, 0.3, 1.0], "position": [0.0, 2.6, -0.5], "materials": [0],
                 "texture": "textures/wall2.png"},
                {"name": "ball", "type": "sphere", "radius": 1.5, "position": [0.0, 1.5, -3.2], "materials": [1]}],
}

# BoundingSphere (bounding_volumes.py:12-37) is chosen by Mesh.__init__ only when the sphere is smaller than the box
# (mesh.py:45-51); the shipped torus gets the box.  An irregular octahedron (written by export_inputs, not a reference
# file) gets the sphere; smooth shaded, translated and scaled, with a mirror ball so rays also arrive from behind.
OCTA_OBJ = """# tests/golden fixture (authored for this repo): irregular octahedron
v 1.1 0 0
v -1 0 0
v 0 1.2 0
v 0 -0.9 0
v 0 0 1
v 0 0 -1.05
f 1 3 5
f 3 2 5
f 2 4 5
f 4 1 5
f 3 1 6
f 2 3 6
f 4 2 6
f 1 4 6
"""
SPHERE_GATE = {
    "resolution": [200, 150], "AA": {"jitter": False, "samples": 1},
    "materials": [{"name": "blue", "ID": 0, "diffuse": [0.1, 0.2, 0.9], "specular": [0.6, 0.6, 0.6], "hardness": 24},
                  {"name": "mirror", "ID": 1, "diffuse": [0.1, 0.1, 0.1], "specular": [0.3, 0.3, 0.3], "hardness": 50,
                   "type": "mirror", "tint": 0.85},
                  {"name": "black", "ID": 2, "diffuse": [0.0, 0.0, 0.0], "specular": [0.0, 0.0, 0.0], "hardness": 0},
                  {"name": "white", "ID": 3, "diffuse": [1.0, 1.0, 1.0], "specular": [0.0, 0.0, 0.0], "hardness": 0}],
    "objects": [{"name": "plane", "type": "plane", "normal": [0.0, 1.0, 0.0], "position": [0.0, 0.0, 0.0], "materials": [2, 3]},
                {"name": "octa", "type": "mesh", "filepath": "./meshes/octa.obj", "scale": 1.3, "position": [-0.4, 1.05, 0.2],
                 "materials": [0], "flat_shaded": False},
                {"name": "ball", "type": "sphere", "radius": 1.2, "position": [1.9, 1.2, -2.0], "materials": [1]}],
}

# Meshes as CHILDREN of transform / CSG nodes (scene_parser.py:262-275 allows it, no shipped scene does it): a rotated,
# non-uniformly scaled union holding the torus, and a difference torus - sphere (Mesh.is_inside is the base class's False).
MESH_IN_NODE = {
    "resolution": [200, 150], "AA": {"jitter": False, "samples": 1},
    "materials": [{"name": "blue", "ID": 0, "diffuse": [0.1, 0.2, 0.9], "specular": [0.6, 0.6, 0.6], "hardness": 24},
                  {"name": "red", "ID": 1, "diffuse": [0.9, 0.15, 0.1], "specular": [0.5, 0.5, 0.5], "hardness": 40},
                  {"name": "black", "ID": 2, "diffuse": [0.0, 0.0, 0.0], "specular": [0.0, 0.0, 0.0], "hardness": 0},
                  {"name": "white", "ID": 3, "diffuse": [1.0, 1.0, 1.0], "specular": [0.0, 0.0, 0.0], "hardness": 0},
                  {"name": "mirror", "ID": 4, "diffuse": [0.1, 0.1, 0.1], "specular": [0.3, 0.3, 0.3], "hardness": 50,
                   "type": "mirror", "tint": 0.85}],
    "objects": [{"name": "plane", "type": "plane", "normal": [0.0, 1.0, 0.0], "position": [0.0, 0.0, 0.0], "materials": [2, 3]},
                {"name": "tilted", "type": "node", "hierarchy_type": "union", "position": [-1.5, 1.3, 0.0],
                 "rotation": [35.0, 20.0, 10.0], "scale": [1.0, 1.4, 0.8], "materials": [1],
                 "children": [{"name": "torus", "type": "mesh", "filepath": "./meshes/torus.obj", "scale": 1.0,
                               "position": [0.0, 0.0, 0.0], "materials": [0], "flat_shaded": True},
                              {"name": "hub", "type": "sphere", "radius": 0.35, "position": [0.0, 0.0, 0.0], "materials": []}]},
                {"name": "bitten", "type": "node", "hierarchy_type": "difference", "position": [1.4, 0.9, 0.3],
                 "rotation": [60.0, 0.0, 0.0], "materials": [1],
                 "children": [{"name": "torus", "type": "mesh", "filepath": "./meshes/torus.obj", "scale": 1.1,
                               "position": [0.0, 0.0, 0.0], "materials": [], "flat_shaded": False},
                              {"name": "bite", "type": "sphere", "radius": 0.9, "position": [1.0, 0.0, 0.3], "materials": [0]}]},
                {"name": "ball", "type": "sphere", "radius": 0.9, "position": [0.2, 0.9, -3.0], "materials": [4]}],
}

# Random scenes (tests/fuzz_scenes.py): combinations of primitives, nested CSG / transform nodes, `ref` copies, meshes inside
# nodes, moving objects, textures, all three material types and AA / DOF / motion sampling that no hand-written scene has.
sys.path.insert(0, os.path.join(HERE, "..", "tests"))
import fuzz_scenes  # noqa: E402
FUZZ_SEEDS = (1, 8, 12, 18, 19)

CASES = [  # (case, scene, overrides, author's render to cross-check or None)
    ("TwoSpheresPlane", "TwoSpheresPlane", None, "renders/TwoSpheresPlane.png"),
    ("MirrorRefraction", "MirrorRefraction", None, "renders/MirrorRefraction.png"),
    ("MotionBlur", "MotionBlur", None, "renders/MotionBlur.png"),
    ("TorusMesh_flat", "TorusMesh", None, "renders/TorusMesh_flat.png"),
    ("TorusMesh_smooth", "TorusMesh", "torus_smooth", "renders/TorusMesh.png"),
    ("DepthOfField", "DepthOfField", None, None),
    ("BoxTexture", "TwoSpheresPlane", BOX_TEXTURE, None),
    ("MeshSphereGate", "TwoSpheresPlane", SPHERE_GATE, None),
    ("MeshInNode", "TwoSpheresPlane", MESH_IN_NODE, None),
    ("NovelScene1_small", "NovelScene1", {"resolution": [256, 128], "AA": {"jitter": True, "samples": 2}}, None),
    ("NovelScene2_small", "NovelScene2", {"resolution": [128, 64], "AA": {"jitter": True, "samples": 2},
                                          "DOF": {"samples": 3}, "motion": {"samples": 3, "final": 1}}, None),
] + [(f"Fuzz{s}", "TwoSpheresPlane", fuzz_scenes.make(s, 128, 96), None) for s in FUZZ_SEEDS]


def export_inputs():
    os.makedirs(os.path.join(OUT, "scenes"), exist_ok=True)
    os.makedirs(os.path.join(OUT, "meshes"), exist_ok=True)
    os.makedirs(os.path.join(OUT, "textures"), exist_ok=True)
    textures = set()

    def walk(objs):
        for o in objs:
            if "texture" in o:
                textures.add(o["texture"])
            walk(o.get("children", []))

    for name, rel in SCENES.items():
        with open(os.path.join(REF, rel)) as fh:
            data = json.load(fh)
        walk(data["objects"])
        with open(os.path.join(OUT, "scenes", name + ".json"), "w") as fh:
            json.dump(data, fh, separators=(",", ":"))
    for t in sorted(textures):
        im = Image.open(os.path.join(REF, t))
        assert t.endswith(".png")
        im.save(os.path.join(OUT, t), optimize=True)   # same pixels and mode, re-encoded
    with open(os.path.join(OUT, "meshes", "octa.obj"), "w") as out:
        out.write(OCTA_OBJ)
    # torus: keep only v / f rows
    with open(os.path.join(REF, "meshes/torus.obj")) as fh, open(os.path.join(OUT, "meshes", "torus.obj"), "w") as out:
        out.write("# tests/golden fixture: v/f rows of the reference's meshes/torus.obj\n")
        for line in fh:
            p = line.split()
            if p and p[0] in ("v", "f"):
                out.write(" ".join(p) + "\n")


def _for_reference(overrides):
    """The reference resolves mesh paths against its own (read-only) tree: fixtures authored here are given by absolute path."""
    ov = json.loads(json.dumps(overrides)) if overrides else overrides
    def walk(objs):
        for o in objs:
            if o.get("filepath") == "./meshes/octa.obj":
                o["filepath"] = os.path.abspath(os.path.join(OUT, "meshes", "octa.obj"))
            walk(o.get("children", []))
    walk((ov or {}).get("objects", []))
    return ov


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", nargs="*", default=None)
    ap.add_argument("--seed", type=lambda s: int(s, 0), default=0x5EED)
    a = ap.parse_args()
    export_inputs()
    os.makedirs(os.path.join(OUT, "renders"), exist_ok=True)
    man_path = os.path.join(OUT, "MANIFEST.json")
    manifest = json.load(open(man_path)) if os.path.exists(man_path) else {}
    for case, scene, ov, author in CASES:
        if a.only and case not in a.only:
            continue
        overrides = _torus_smooth() if ov == "torus_smooth" else ov
        t0 = time.time()
        img, cpu = ref_runner.render_reference(SCENES[scene], overrides=_for_reference(overrides), seed=a.seed)
        png = ref_runner.to_png_array(img)
        assert np.array_equal(img.astype(np.float32).astype(np.float64), img)
        Image.fromarray(png).save(os.path.join(OUT, "renders", case + ".png"), optimize=True)
        cols = np.arange(0, img.shape[0], 4)
        rows = np.arange(0, img.shape[1], 4)
        np.savez_compressed(os.path.join(OUT, "renders", case + "_f32.npz"), cols=cols, rows=rows,
                            rgb=img[np.ix_(cols, rows)].astype(np.float32))
        entry = {"scene": scene + ".json", "overrides": overrides, "seed": a.seed,
                 "width": int(img.shape[0]), "height": int(img.shape[1]),
                 "reference_cpu_seconds": round(cpu, 1), "generator": "oracle/make_golden.py"}
        if author:
            ref_png = np.array(Image.open(os.path.join(REF, author)).convert("RGB"))
            diff = np.abs(ref_png.astype(int) - png.astype(int)).max(axis=2)
            entry["author_render"] = author
            entry["author_render_max_abs_diff_u8"] = int(diff.max())
            entry["author_render_pixels_differing"] = int((diff > 0).sum())
        manifest[case] = entry
        print(f"{case}: {img.shape} {time.time()-t0:.1f}s wall {cpu:.1f}s cpu "
              f"author-diff={entry.get('author_render_max_abs_diff_u8')}", flush=True)
        with open(man_path, "w") as fh:
            json.dump(manifest, fh, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
