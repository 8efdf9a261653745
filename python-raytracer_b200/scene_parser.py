"""JSON scene -> Scene object graph: mirror of the reference's `provided/scene_parser.py`.

`load_scene(infile)` keeps the reference's schema, defaults and quirks (citations are
scene_parser.py lines): grouped defaults for AA / DOF / motion (:68-96), directional-light
power forced to 1.0 (:118), a KeyError in any light drops ALL lights (:123-125), material
defaults (:133-138), node `ref` = deep copy with new matrices and materials (:190-203),
the ROOT node's speed pushed down to every primitive leaf (:268-271,283), fallback materials
appended after parsing (:153-155), `set_scene` wiring (:160-161).
"""
import copy
import json
import os

import numpy as np
from PIL import Image

from . import geometry as geom
from . import helperclasses as hc
from . import scene as scene_mod

VERBOSE = True  # the reference prints its progress; tests switch this off


def _say(*a):
    if VERBOSE:
        print(*a)


def populateVec(array):
    if array is None:
        return None
    if len(array) in (3, 4):
        return np.asarray(array, dtype=np.float64).astype(np.float32)
    return None


def get_or(obj, keys, default, msg=""):
    if not isinstance(keys, list) or len(keys) == 0:
        try:
            return obj[keys]
        except KeyError:
            if msg != "":
                _say(msg)
            return default
    try:
        return [obj[key] for key in keys]
    except KeyError:
        if msg != "":
            _say(msg)
        return default


_search_roots = []


def _resolve(path):
    """The reference opens asset paths relative to the cwd (it must be run from its repo root);
    additionally look beside / above the scene file so scenes work from any cwd."""
    if os.path.isabs(path) or os.path.exists(path):
        return path
    for root in _search_roots:
        cand = os.path.join(root, path)
        if os.path.exists(cand):
            return cand
    return path


def load_scene(infile, ingest=None):
    """scene_parser.py:19-163.  `ingest` ("host" | "device", default geometry.Mesh.DEFAULT_INGEST) says where meshes are
    read and prepared: see geometry.Mesh."""
    _say("Parsing file:", infile)
    with open(infile) as f:
        data = json.load(f)
    return load_scene_dict(data, os.path.dirname(os.path.abspath(infile)), ingest=ingest)


_mesh_ingest = None


def load_scene_dict(data, base_dir=None, ingest=None):
    global _search_roots, _mesh_ingest
    _mesh_ingest = ingest
    _search_roots = [] if base_dir is None else [base_dir, os.path.dirname(base_dir)]

    cam_pos = populateVec(data["camera"]["position"])
    cam_lookat = populateVec(data["camera"]["lookAt"])
    cam_up = populateVec(data["camera"]["up"])
    cam_fov = data["camera"]["fov"]

    width, height = get_or(data, "resolution", [1080, 720], "No resolution found, defaulting to 1080x720.")
    ambient = populateVec(get_or(data, "ambient", [0, 0, 0], "No ambient light defined, defaulting to [0, 0, 0]"))

    try:
        jitter = data["AA"]["jitter"]
        samples = data["AA"]["samples"]
    except KeyError:
        _say("No Anti-Aliasing options found, setting to default")
        jitter, samples = False, 1

    try:
        focal_length = data["DOF"]["focal_length"]
        aperture = data["DOF"]["aperture"]
        dof_samples = data["DOF"]["samples"]
    except KeyError:
        _say("No Depth of Field options found, setting to default")
        focal_length, aperture, dof_samples = 1, 0, 1

    try:
        motion_time = data["motion"]["time"]
        motion_samples = data["motion"]["samples"]
        motion_final = data["motion"]["final"]
    except KeyError:
        _say("No motion blur options found, setting to default")
        motion_time, motion_samples, motion_final = 0, 1, 0

    vc = hc.ViewportCamera() \
        .set_viewport(width, height) \
        .set_camera(cam_pos, cam_lookat, cam_up, cam_fov) \
        .set_lens(focal_length, aperture, dof_samples) \
        .set_motion(motion_time, motion_samples, motion_final)

    lights = []
    try:
        for light in data["lights"]:
            l_type = light["type"]
            l_name = light["name"]
            l_colour = populateVec(light["colour"])
            if l_type == "point":
                l_vector = populateVec(light["position"])
                l_power = light["power"]
            elif l_type == "directional":
                l_vector = populateVec(light["direction"])
                l_power = 1.0
            else:
                _say("Unkown light type", l_type, ", skipping initialization")
                continue
            lights.append(hc.Light(l_type, l_name, l_colour, l_vector, l_power))
    except KeyError as e:
        _say("Error loading lights: ", e)
        lights = []

    materials = []
    for material in data["materials"]:
        mat = hc.Material(material["name"],
                          populateVec(get_or(material, "specular", [0, 0, 0])),
                          populateVec(get_or(material, "diffuse", [0, 0, 0])),
                          get_or(material, "hardness", 32),
                          material["ID"],
                          get_or(material, "type", "diffuse"),
                          get_or(material, "tint", 0.0))
        mat.refr_index = get_or(material, "refr_index", 1.0)
        materials.append(mat)

    objects, root_names, roots = [], [], []
    for g in data["objects"]:
        parse_geometry(g, objects, root_names, roots, materials)

    for obj in objects:
        if isinstance(obj, geom.Hierarchy):
            obj.set_fallback_material(obj.materials)

    _say("Parsing complete")
    sc = scene_mod.Scene(vc, jitter, samples, ambient, lights, materials, objects)
    # extension key (absent from the reference's schema, which ignores unknown keys): meshes move like the primitives do
    sc.mesh_motion = bool(data.get("mesh_motion", False))
    for obj in objects:
        obj.set_scene(sc)
    return sc


def parse_geometry(g, objects, root_names, roots, materials):
    g_name = g["name"]
    g_type = g["type"]
    g_pos = populateVec(get_or(g, "position", [0, 0, 0]))
    g_mats = associate_material(materials, get_or(g, "materials", []))
    g_speed = populateVec(get_or(g, "speed", None))

    if add_basic_shape(g_name, g_type, g_pos, g_speed, g_mats, g, objects):
        return
    if g_type == "node":
        g_ref = get_or(g, "ref", "")
        g_r = populateVec(get_or(g, "rotation", [0, 0, 0]))
        g_s = populateVec(get_or(g, "scale", [1, 1, 1]))
        g_htype = get_or(g, "hierarchy_type", "union")
        if g_ref == "":
            root_names.append(g_name)
            node = geom.Hierarchy(g_name, g_type, g_mats, g_htype, g_pos, g_r, g_s, g_speed)
            traverse_children(node, g["children"], materials, root_names, roots, g_speed)
            roots.append(node)
            objects.append(node)
        else:
            rid = root_names.index(g_ref) if g_ref in root_names else -1
            if rid != -1 and rid < len(roots):
                node = copy.deepcopy(roots[rid])
                node.name = g_name
                node.materials = g_mats
                node.make_matrices(g_pos, g_r, g_s)
                objects.append(node)
            else:
                _say("Node reference", g_ref, "not found, skipping creation")
    else:
        _say("Unkown object type", g_type, ", skipping initialization")


def _open_texture(path):
    im = Image.open(_resolve(path))
    im.load()
    return im


def add_basic_shape(g_name, g_type, g_pos, g_speed, g_mats, g, objects):
    if g_type == "sphere":
        objects.append(geom.Sphere(g_name, g_type, g_mats, g_pos, g["radius"], g_speed))
    elif g_type == "plane":
        plane = geom.Plane(g_name, g_type, g_mats, g_pos, populateVec(g["normal"]), g_speed)
        if "texture" in g:
            plane.texture = _open_texture(g["texture"])
            plane.texture_scale = get_or(g, "texture_scale", 1.0)
        objects.append(plane)
    elif g_type == "box":
        if "size" in g:
            box = geom.AABB(g_name, g_type, g_mats, g_pos, populateVec(g["size"]), g_speed)
        else:
            box = geom.AABB(g_name, g_type, g_mats, g_pos, (0, 0, 0), g_speed)
            box.minpos = populateVec(g["min"])
            box.maxpos = populateVec(g["max"])
        if "texture" in g:
            box.texture = _open_texture(g["texture"])
        objects.append(box)
    elif g_type == "mesh":
        objects.append(geom.Mesh(g_name, g_type, g_mats, g_pos, g["scale"], _resolve(g["filepath"]),
                                 get_or(g, "flat_shaded", False), g_speed, ingest=_mesh_ingest))
    else:
        return False
    return True


def traverse_children(node, children, materials, root_names, roots, speed):
    for g in children:
        g_name = g["name"]
        g_type = g["type"]
        g_pos = populateVec(get_or(g, "position", [0, 0, 0]))
        g_mats = associate_material(materials, get_or(g, "materials", []))
        if speed is None:
            g_speed = None
        else:
            g_speed = speed + populateVec(get_or(g, "speed", [0, 0, 0]))

        if add_basic_shape(g_name, g_type, g_pos, g_speed, g_mats, g, node.children):
            continue
        if g_type == "node":
            g_r = populateVec(get_or(g, "rotation", [0, 0, 0]))
            g_s = populateVec(get_or(g, "scale", [1, 1, 1]))
            g_htype = get_or(g, "hierarchy_type", "union")
            inner = geom.Hierarchy(g_name, g_type, g_mats, g_htype, g_pos, g_r, g_s, g_speed)
            node.children.append(inner)
            traverse_children(inner, g["children"], materials, root_names, roots, speed)
        else:
            parse_geometry(g, node.children, root_names, roots, materials)


def associate_material(mats, ids):
    return [mat for i in ids for mat in mats if i == mat.ID]
