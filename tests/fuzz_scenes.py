"""Seeded random scenes in the reference's JSON schema (scene_parser.py): every primitive kind, transform / CSG nodes
(union, intersection, difference, nested, `ref` copies), meshes at top level and inside nodes, moving objects, textures,
checker planes, the three material types, point and directional lights, AA / DOF / motion sampling.

Test infrastructure.  Used three ways, always from the same seeds:
  * oracle/make_golden.py renders Fuzz<seed> cases with the UNMODIFIED reference (dev container) -> tests/golden fixtures
    that pin the C oracle on code paths the hand-written scenes do not combine;
  * tests/test_host_logic.py: the host-compiled device functions against the oracle on fresh seeds;
  * tests/test_gpu_parity.py: the CUDA path against the oracle on fresh seeds.
Asset paths are the reference's own relative ones (`./meshes/torus.obj`, `textures/*.png`): they resolve under
/root/reference for the reference and under tests/golden for the package's parser.
"""
import random

TEXTURES = ["textures/wall1.png", "textures/wall2.png", "textures/wall3.png", "textures/wall4.png", "textures/ground.png"]


def _v(r, lo, hi):
    return [round(r.uniform(lo, hi), 3) for _ in range(3)]


def _colour(r):
    return [round(r.uniform(0.05, 1.0), 3) for _ in range(3)]


def _materials(r):
    mats = []
    for k in range(4):
        mats.append({"name": f"d{k}", "ID": k, "diffuse": _colour(r), "specular": [round(r.uniform(0.0, 0.9), 3)] * 3,
                     "hardness": r.choice([0, 1, 8, 32, 64])})
    mats.append({"name": "mirror", "ID": 4, "diffuse": _colour(r), "specular": [0.6, 0.6, 0.6], "hardness": 32,
                 "type": "mirror", "tint": round(r.uniform(0.2, 0.9), 3)})
    mats.append({"name": "glass", "ID": 5, "diffuse": _colour(r), "specular": [0.5, 0.5, 0.5], "hardness": 16,
                 "type": "refractive", "refr_index": round(r.uniform(1.1, 1.8), 3), "tint": round(r.uniform(0.1, 0.7), 3)})
    mats.append({"name": "black", "ID": 6, "diffuse": [0.05, 0.05, 0.05], "specular": [0.0, 0.0, 0.0], "hardness": 0})
    mats.append({"name": "white", "ID": 7, "diffuse": [0.95, 0.95, 0.95], "specular": [0.0, 0.0, 0.0], "hardness": 0})
    return mats


def _leaf(r, name, where, size, mats, allow_mesh=True, moving=False):
    """One primitive around `where` (list of 3) with extent ~`size`."""
    kind = r.choice(["sphere", "sphere", "box", "box", "mesh"] if allow_mesh else ["sphere", "box"])
    o = {"name": name, "type": kind, "position": [round(w + r.uniform(-0.3, 0.3) * size, 3) for w in where], "materials": mats}
    if kind == "sphere":
        o["radius"] = round(r.uniform(0.35, 0.8) * size, 3)
    elif kind == "box":
        if r.random() < 0.5:
            o["size"] = _v(r, 0.5 * size, 1.4 * size)
        else:
            p = o.pop("position")
            o["min"] = [round(p[k] - r.uniform(0.3, 0.7) * size, 3) for k in range(3)]
            o["max"] = [round(p[k] + r.uniform(0.3, 0.7) * size, 3) for k in range(3)]
        if r.random() < 0.35:
            o["texture"] = r.choice(TEXTURES[:4])
    else:
        o["filepath"] = "./meshes/torus.obj"
        o["scale"] = round(r.uniform(0.5, 0.9) * size, 3)
        o["flat_shaded"] = r.random() < 0.5
    if moving and kind != "mesh":
        o["speed"] = _v(r, -0.4, 0.4)
    return o


def _node(r, name, where, size, mats, depth):
    n_children = r.randint(2, 3)
    node = {"name": name, "type": "node", "hierarchy_type": r.choice(["union", "intersection", "difference", "difference"]),
            "position": [round(w, 3) for w in where], "rotation": _v(r, -60.0, 60.0),
            "scale": [round(r.uniform(0.7, 1.4), 3) for _ in range(3)], "materials": mats, "children": []}
    for c in range(n_children):
        cm = r.choice([[], [r.randint(0, 3)], [r.choice([4, 5])]])
        off = [0.0, 0.0, 0.0] if c == 0 else _v(r, -0.45 * size, 0.45 * size)
        if depth > 0 and r.random() < 0.3:
            node["children"].append(_node(r, f"{name}_n{c}", off, 0.8 * size, cm, depth - 1))
        else:
            node["children"].append(_leaf(r, f"{name}_c{c}", off, 0.9 * size, cm, allow_mesh=r.random() < 0.4))
    return node


def make(seed, width=96, height=64):
    """Scene JSON (dict) of fuzz case `seed`."""
    r = random.Random(0xF0220000 + seed)
    cam_r, cam_a = r.uniform(6.0, 9.0), r.uniform(-0.9, 0.9)
    import math
    data = {
        "resolution": [width, height],
        "camera": {"position": [round(cam_r * math.sin(cam_a), 3), round(r.uniform(1.5, 4.5), 3), round(cam_r * math.cos(cam_a), 3)],
                   "lookAt": [round(r.uniform(-0.5, 0.5), 3), round(r.uniform(0.5, 1.5), 3), 0.0], "up": [0.0, 1.0, 0.0],
                   "fov": round(r.uniform(38.0, 60.0), 2)},
        "ambient": [round(r.uniform(0.0, 0.3), 3)] * 3,
        "AA": {"jitter": r.random() < 0.5, "samples": r.choice([1, 1, 2, 3])},
        "materials": _materials(r),
    }
    if r.random() < 0.5:
        data["DOF"] = {"focal_length": round(r.uniform(5.0, 9.0), 3), "aperture": round(r.uniform(0.02, 0.15), 3),
                       "samples": r.choice([1, 2, 3])}
    moving = r.random() < 0.5
    if moving:
        data["motion"] = {"time": round(r.uniform(0.5, 1.5), 3), "samples": r.choice([1, 2]), "final": r.choice([0, 1])}
    lights = []
    for k in range(r.randint(1, 3)):
        if r.random() < 0.6:
            lights.append({"name": f"p{k}", "type": "point", "colour": _colour(r), "power": round(r.uniform(0.4, 1.0), 3),
                           "position": [round(r.uniform(-6, 6), 3), round(r.uniform(3, 9), 3), round(r.uniform(-4, 7), 3)]})
        else:
            lights.append({"name": f"s{k}", "type": "directional", "colour": _colour(r), "power": 1.0,
                           "direction": [round(r.uniform(-0.8, 0.8), 3), round(r.uniform(-1.0, -0.4), 3), round(r.uniform(-0.8, 0.8), 3)]})
    data["lights"] = lights
    objs = []
    ground = {"name": "ground", "type": "plane", "normal": [0.0, 1.0, 0.0], "position": [0.0, round(r.uniform(-0.6, 0.0), 3), 0.0],
              "materials": r.choice([[6, 7], [r.randint(0, 3)], [4, 7]])}
    if r.random() < 0.4:
        ground["texture"] = TEXTURES[4]
        ground["texture_scale"] = round(r.uniform(1.0, 6.0), 2)
    objs.append(ground)
    if r.random() < 0.4:          # a second, tilted plane behind the scene
        objs.append({"name": "wall", "type": "plane", "normal": [round(r.uniform(-0.3, 0.3), 3), round(r.uniform(0.0, 0.3), 3), 1.0],
                     "position": [0.0, 0.0, round(r.uniform(-7.0, -4.0), 3)], "materials": [r.randint(0, 3), 7]})
    spots = [[x, 0.0, z] for x in (-3.0, -1.0, 1.0, 3.0) for z in (-2.0, 0.0, 2.0)]
    r.shuffle(spots)
    n_obj = r.randint(4, 7)
    first_node = None
    for k in range(n_obj):
        x, _, z = spots[k]
        size = r.uniform(0.7, 1.2)
        where = [x + r.uniform(-0.4, 0.4), size + r.uniform(0.0, 0.8), z + r.uniform(-0.4, 0.4)]
        mats = [r.choice([0, 1, 2, 3, 3, 4, 5])]
        roll = r.random()
        if roll < 0.35:
            node = _node(r, f"node{k}", where, size, mats, depth=1)
            if moving and r.random() < 0.4:
                node["speed"] = _v(r, -0.3, 0.3)
            objs.append(node)
            first_node = first_node or node["name"]
        elif roll < 0.45 and first_node is not None:
            objs.append({"name": f"ref{k}", "type": "node", "ref": first_node, "position": [round(w, 3) for w in where],
                         "rotation": _v(r, -90.0, 90.0), "scale": [round(r.uniform(0.6, 1.2), 3)] * 3, "materials": mats})
        else:
            objs.append(_leaf(r, f"obj{k}", where, size, mats, allow_mesh=True, moving=moving and r.random() < 0.5))
    data["objects"] = objs
    return data
