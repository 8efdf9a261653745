"""CPU tests of the host side: parser semantics, flattening, the C ABI library's exports, and the
per-ray device functions compiled for the host (tests/hostsim) against the oracle."""
import ctypes as C
import json
import os
import re

import numpy as np
import pytest

import hostsim
import oracle
import python_raytracer_b200 as rt
from conftest import ALL_CASES, FUZZ_HOST_SEEDS, GOLDEN, PARITY_FRACTION, ROOT, load_case, load_fuzz, parity_stats, to_png_array

abi = rt._abi


def _scene_dict(**over):
    base = {"camera": {"position": [0, 1, 5], "lookAt": [0, 1, 0], "up": [0, 1, 0], "fov": 45.0},
            "materials": [{"name": "a", "ID": 0, "diffuse": [1, 0, 0]}], "objects": [], "lights": []}
    base.update(over)
    return base


def test_parser_defaults_match_reference():
    rt.scene_parser.VERBOSE = False
    sc = rt.scene_parser.load_scene_dict(_scene_dict())
    assert (sc.vc.width, sc.vc.height) == (1080, 720)            # scene_parser.py:62
    assert sc.samples == 1 and sc.jitter is False                 # :68-74
    assert sc.vc.dof_samples == 1 and sc.vc.aperture == 0 and sc.vc.focal_length == 1   # :76-85
    assert sc.vc.motion_times == [0.0]                            # :87-96, helperclasses.py:103-108
    m = sc.materials[0]
    assert m.hardness == 32 and m.tint == 0.0 and m.refr_index == 1.0 and m.mat_type == "diffuse"  # :133-138
    assert np.array_equal(m.specular, [0, 0, 0])


def test_parser_grouped_defaults_and_light_quirks():
    rt.scene_parser.VERBOSE = False
    # a missing key inside AA resets BOTH values (scene_parser.py:68-74)
    sc = rt.scene_parser.load_scene_dict(_scene_dict(AA={"samples": 4}))
    assert sc.samples == 1 and sc.jitter is False
    # directional power is forced to 1.0 (:118); an unknown type is skipped (:120)
    lights = [{"type": "directional", "name": "d", "colour": [1, 1, 1], "direction": [0, -1, 0], "power": 7.0},
              {"type": "spot", "name": "s", "colour": [1, 1, 1]},
              {"type": "point", "name": "p", "colour": [1, 1, 1], "position": [0, 5, 0], "power": 0.5}]
    sc = rt.scene_parser.load_scene_dict(_scene_dict(lights=lights))
    assert [l.type for l in sc.lights] == ["directional", "point"] and sc.lights[0].power == 1.0
    # a KeyError in one light drops ALL lights (:123-125)
    lights[2].pop("power")
    sc = rt.scene_parser.load_scene_dict(_scene_dict(lights=lights))
    assert sc.lights == []


def test_motion_times():
    vc = rt.helperclasses.ViewportCamera().set_motion(2.0, 16, 1)
    assert len(vc.motion_times) == 17 and vc.motion_times[-1] == 2.0 and vc.motion_times[1] == 0.125


def test_node_speed_pushdown_fallback_materials_and_ref_copy():
    sc = load_case("NovelScene1_small")
    bike = sc.objects[5]
    assert isinstance(bike, rt.geometry.Hierarchy) and bike.name == "bike"
    # the ROOT speed reaches every primitive leaf, nested nodes pass it on unchanged (scene_parser.py:268-283)
    inner = bike.children[0].children[0].children[0].children[0]
    assert isinstance(inner, rt.geometry.Sphere) and np.array_equal(inner.speed, [0, 0, 0.5])
    # set_fallback_material APPENDS the root's materials (hierarchy.py:21-28)
    outer = bike.children[0].children[0].children[0].children[1]
    assert [m.name for m in outer.materials] == ["blue"]
    assert [m.name for m in inner.materials] == ["gloss_black", "blue"]
    bike2 = sc.objects[6]
    assert bike2.name == "bike2" and bike2 is not bike and bike2.children[0] is not bike.children[0]
    assert [m.name for m in bike2.children[0].children[0].children[0].children[1].materials] == ["gold"]
    assert np.array_equal(bike2.t, [-2.0, -2.5, 2.0])


def test_plane_axes_and_box_corners():
    P = rt.geometry.Plane
    assert np.array_equal(P("p", "plane", [], (0, 0, 0), (0, 1, 0), None).width_axis, [1, 0, 0])
    assert np.array_equal(P("p", "plane", [], (0, 0, 0), (1, 0, 0), None).width_axis, [0, 0, -1])
    assert np.array_equal(P("p", "plane", [], (0, 0, 0), (1, 0, 0), None).height_axis, [0, -1, 0])
    b = rt.geometry.AABB("b", "box", [], (1.5, 4.0, -15.0), (6, 6, 6), None)
    assert np.array_equal(b.minpos, [-1.5, 1.0, -18.0]) and np.array_equal(b.maxpos, [4.5, 7.0, -12.0])


def test_trs_matrix_inverse():
    M, Minv = rt._vecmath.trs((1, 2, 3), (15, 30, 45), (0.7, 1.0, 2.0))
    assert M.dtype == np.float32
    # column-major storage m[col,row]: as a maths matrix it is M.T
    assert np.allclose(M.T @ Minv.T, np.eye(4), atol=1e-6)
    assert np.allclose(M[3, :3], [1, 2, 3])


def test_mesh_host_setup_matches_reference_semantics():
    sc = load_case("TorusMesh_smooth")
    mesh = sc.objects[1]
    assert mesh.verts.shape == (64, 3) and mesh.faces.shape == (128, 3) and mesh.norms.shape == (64, 3)
    assert np.allclose(np.linalg.norm(mesh.norms, axis=1), 1.0, atol=1e-6)
    # mesh.py:45-51: torus is flat -> AABB volume < sphere volume
    assert isinstance(mesh.bounding_volume, rt.geometry.BoundingAABB)


def test_flatten_layout():
    sc = load_case("NovelScene1_small")
    fs = sc.flatten()
    d = fs.desc
    assert d.n_top == 9 and d.n_objects == 91 and d.n_xforms == 38 and d.n_textures == 5
    objs = C.cast(d.objects, C.POINTER(abi.RtObject))
    assert C.sizeof(abi.RtObject) == 128 and C.sizeof(abi.RtMaterial) == 56 and C.sizeof(abi.RtLight) == 32
    for i in range(d.n_objects):
        o = objs[i]
        if o.kind == abi.KIND_NODE:
            assert d.n_top <= o.first_child and o.first_child + o.n_children <= d.n_objects
    assert objs[0].kind == abi.KIND_PLANE and objs[0].tex == 0 and objs[0].tex_scale == 4.0
    # inverted box (scenes/NovelScene1.json:402-403) is kept as given
    assert objs[7].kind == abi.KIND_BOX and objs[7].a[2] > objs[7].b[2]


def test_header_and_ctypes_agree_on_entry_points():
    with open(os.path.join(ROOT, "include", "b200rt.h")) as fh:
        hdr = fh.read()
    declared = set(re.findall(r"^(?:int|const char\*)\s+(rt_\w+)\s*\(", hdr, flags=re.M))
    assert declared == set(abi.ENTRY_POINTS)


def test_library_loads_and_exports_every_declared_symbol():
    # no compute calls here (no GPU in this container): dlopen + symbol presence + ABI version
    assert os.path.exists(abi.LIB_PATH), "build with `python __graft_entry__.py`"
    lib = abi.load_library()
    for name in abi.ENTRY_POINTS:
        assert hasattr(lib, name), name
    assert lib.rt_abi_version() == abi.RT_ABI_VERSION


def test_native_obj_reader_equals_python_reader(tmp_path):
    """rt_obj_load (host code of the C ABI library) against geometry.read_obj: bundled torus, and a file with the row
    kinds OBJ exporters produce (comments, CRLF, v/vt/vn corners, negative indices, quads and a pentagon, exponents)."""
    from python_raytracer_b200 import geometry
    torus = os.path.join(GOLDEN, "meshes", "torus.obj")
    for a, b in zip(geometry.read_obj(torus), abi.obj_load(torus)):
        assert a.dtype == b.dtype and np.array_equal(a, b)
    rng = np.random.default_rng(5)
    rows = ["# made by a test", "o thing", ""]
    for k in range(40):
        x, y, z = rng.normal(size=3).tolist()
        rows.append(f"v {x!r} {y:.7e} {z:.3f} 1.0" if k % 3 else f"v  {x:.17g}\t{y:.9f}   {z!r}")
        rows.append(f"vn {x:.4f} {y:.4f} {z:.4f}")
        rows.append(f"vt {x:.4f} {y:.4f}")
        if k >= 5:
            rows.append([f"f {k} {k - 1} {k - 2}", f"f {k}/1/1 {k - 1}/2/2 {k - 2}//3 {k - 3}/4", "f -1 -2 -3 -4 -5",
                         f"f {k}//{k} -2//1 {k - 4}//2"][k % 4])
    rows += ["s off", "usemtl m", "f 1 2 3"]
    path = tmp_path / "t.obj"
    path.write_bytes("\r\n".join(rows).encode())
    want, got = geometry.read_obj(str(path)), abi.obj_load(str(path))
    assert want[0].shape == (40, 3) and want[2].shape[0] > 40
    for a, b in zip(want, got):
        assert a.dtype == b.dtype and np.array_equal(a, b)
    bad = tmp_path / "bad.obj"
    bad.write_text("v 0 0 0\nv 1 0 0\nv 0 1 x\nf 1 2 3\n")
    with pytest.raises(abi.B200RTError, match="bad.obj:3"):
        abi.obj_load(str(bad))
    bad.write_text("v 0 0 0\nv 1 0 0\nv 0 1 0\nf 1 2 4\n")
    with pytest.raises(abi.B200RTError, match="out of range"):
        abi.obj_load(str(bad))
    with pytest.raises(abi.B200RTError, match="cannot open"):
        abi.obj_load(str(tmp_path / "missing.obj"))


def test_missing_library_fails_loudly(tmp_path):
    with pytest.raises(abi.B200RTError):
        abi.load_library(str(tmp_path / "nope.so"))


@pytest.mark.parametrize("case", ALL_CASES)
def test_device_functions_on_host_match_oracle(case):
    """The RT_HD per-ray functions the kernels call (fp32, LBVH traversal, folded per-level clamp,
    iterative CSG walk), compiled for the host, against the oracle: the parity statistic of the gpu tests."""
    sc = load_case(case)
    fs = sc.flatten()
    ref, c_ref = oracle.render(fs, abi)
    img, c_sim = hostsim.render(fs, abi)
    frac, dmax, _ = parity_stats(to_png_array(ref), to_png_array(img))
    assert frac >= 0.9995, (frac, dmax)
    for k in ("rays_primary",):
        assert c_ref[k] == c_sim[k]
    for k in ("rays_shadow", "rays_reflect", "rays_refract", "shaded_hits"):
        assert abs(c_ref[k] - c_sim[k]) <= max(4, 2e-4 * c_ref[k]), (k, c_ref[k], c_sim[k])
    assert c_sim["error_flags"] == 0


def test_more_top_level_meshes_than_the_parameter_bank_holds():
    """11 top-level meshes: convert_scene routes the scene to the general kernels (the lean ones read at most RT_CONST_MESH = 8 mesh
    headers from the parameter bank); host twin against the oracle, and the BVHs only prune."""
    from conftest import many_meshes_scene
    sc = many_meshes_scene()
    fs = sc.flatten()
    assert fs.desc.n_meshes >= 1 and sum(1 for o in sc.objects if type(o).__name__ == "Mesh") == 11
    ref, c_ref = oracle.render(fs, abi)
    img, c_sim = hostsim.render(fs, abi)
    frac, dmax, _ = parity_stats(to_png_array(ref), to_png_array(img))
    assert frac >= 0.999, (frac, dmax)
    assert c_sim["error_flags"] == 0 and c_sim["tri_tests"] > 0
    plain, _ = hostsim.render(fs, abi, mesh_mode=3)
    assert np.array_equal(img, plain)


@pytest.mark.parametrize("seed", FUZZ_HOST_SEEDS)
def test_device_functions_on_host_match_oracle_on_random_scenes(seed):
    """Fresh random scenes (nested CSG, `ref` copies, meshes in nodes, moving objects, textures, DOF / motion / jitter):
    same statistic as above, plus: the BVHs and the child-bounds pruning only prune (bit-identical to the plain loops)."""
    sc = load_fuzz(seed)
    fs = sc.flatten()
    ref, c_ref = oracle.render(fs, abi)
    img, c_sim = hostsim.render(fs, abi)
    frac, dmax, _ = parity_stats(to_png_array(ref), to_png_array(img))
    assert frac >= 0.9995, (seed, frac, dmax)
    assert c_ref["rays_primary"] == c_sim["rays_primary"] and c_sim["error_flags"] == 0
    for k in ("rays_shadow", "rays_reflect", "rays_refract", "shaded_hits"):
        assert abs(c_ref[k] - c_sim[k]) <= max(8, 1e-3 * c_ref[k]), (k, c_ref[k], c_sim[k])
    plain, _ = hostsim.render(fs, abi, mesh_mode=3)
    # identical, except where a far-away ray (ground plane hit at the horizon, t ~ 1e4) makes a primitive's float32 test report
    # a hit that is geometrically impossible (cancellation in the sphere discriminant): the conservative boxes skip that
    # primitive, the plain loop evaluates it.  Seen on 4 of 12288 pixels of seed 123, colour difference 5e-5.
    differ = np.abs(np.asarray(img, dtype=np.float32) - plain).max(axis=2) > 0
    assert differ.mean() <= 1e-3 and np.abs(np.asarray(img, dtype=np.float32) - plain).max() < 1e-3


def test_per_ray_functions_on_host_match_oracle_on_random_rays():
    """closest hit, cast_ray colour and any-hit of rays that are not camera rays, through the host twins of rt_trace_rays / rt_occluded."""
    sc = load_fuzz(FUZZ_HOST_SEEDS[0])
    fs = sc.flatten()
    rng = np.random.default_rng(5)
    n = 20000
    o = rng.uniform([-5, 0.2, -4], [5, 5, 6], size=(n, 3)).astype(np.float32)
    tgt = rng.uniform([-3.5, 0, -2.5], [3.5, 2.5, 2.5], size=(n, 3))
    d = (tgt - o).astype(np.float32)
    times = rng.choice(np.asarray(sc.vc.motion_times, dtype=np.float64), size=n)
    ref, got = oracle.trace_rays(fs, o, d, times), hostsim.trace_rays(fs, o, d, times)
    hit_ref, hit_got = np.isfinite(ref["t"]), np.isfinite(got["t"])
    assert (hit_ref == hit_got).mean() >= 0.999
    both = hit_ref & hit_got
    assert (got["geom"][both] == ref["geom"][both]).mean() >= 0.999
    sel = np.flatnonzero(both)[got["geom"][both] == ref["geom"][both]]
    assert np.allclose(got["t"][sel], ref["t"][sel], rtol=2e-4, atol=2e-4)
    assert (np.abs(got["rgb"] - ref["rgb"]).max(axis=1) <= 1.0 / 255.0).mean() >= 0.995
    tmax = np.where(rng.random(n) < 0.5, 1.0, np.inf)
    assert (oracle.occluded(fs, o, d, tmax, times) == hostsim.occluded(fs, o, d, tmax, times)).mean() >= 0.9995


def test_bvh_traversal_equals_brute_force_on_host():
    sc = load_case("TorusMesh_smooth")
    fs = sc.flatten()
    a, ca = hostsim.render(fs, abi, mesh_mode=0)
    b, cb = hostsim.render(fs, abi, mesh_mode=1)
    assert np.array_equal(a, b)
    assert ca["tri_tests"] < cb["tri_tests"] / 10


@pytest.mark.parametrize("case", ALL_CASES)
def test_object_bvh_equals_linear_object_loop_on_host(case):
    """The object BVH over the bounded top-level objects (replaces the `for obj in self.objects` loops of
    scene.py:87,160) must give the image of the plain in-order loop bit for bit (equal-t ties included)."""
    sc = load_case(case)
    fs = sc.flatten()
    a, ca = hostsim.render(fs, abi, mesh_mode=0)
    b, cb = hostsim.render(fs, abi, mesh_mode=2)
    assert np.array_equal(a, b)
    for k in ("rays_primary", "rays_shadow", "rays_reflect", "rays_refract", "shaded_hits"):
        assert ca[k] == cb[k], k
    assert ca["error_flags"] == 0 and cb["error_flags"] == 0


def test_object_bvh_synthetic_scene_fewer_tests_same_image():
    from python_raytracer_b200 import synthetic
    sc = synthetic.make_scene(width=96, height=54, aa_samples=2, grid=(32, 16), n_spheres=64)
    fs = sc.flatten()
    a, ca = hostsim.render(fs, abi, mesh_mode=0)
    b, cb = hostsim.render(fs, abi, mesh_mode=2)
    assert np.array_equal(a, b)
    assert ca["prim_tests"] < cb["prim_tests"] / 4


def test_traversal_stack_spill_gives_the_same_image():
    """Walks deeper than the shared-memory stack continue in thread-local memory (stk_push/stk_pop): with a 3-entry
    stack nearly every walk spills, and the image must not change; with no spill room at all RT_ERR_LIMIT is flagged."""
    sc = load_case("TorusMesh_smooth")
    fs = sc.flatten()
    a, ca = hostsim.render(fs, abi)
    tiny = hostsim.build_variant("stack3", ["-DRT_STACK_DEPTH=3"])
    b, cb = hostsim.render(fs, abi, library=tiny)
    assert np.array_equal(a, b) and cb["error_flags"] == 0 and ca["bvh_nodes"] == cb["bvh_nodes"]
    none = hostsim.build_variant("stack1", ["-DRT_STACK_DEPTH=1", "-DRT_STACK_SPILL=1"])
    _, cc = hostsim.render(fs, abi, library=none)
    assert cc["error_flags"] & (1 << 5)


def test_lbvh_structure_is_a_valid_tree():
    """4-wide nodes (BvhNode4: 6 float4 of child boxes, int4 links, int4 flags): every triangle slot belongs to exactly one
    leaf (a leaf link covers 1..RT_LEAF_TRIS consecutive slots: ~(first | (count-1) << 28)), every node but the root has
    exactly one parent, child boxes contain their subtrees and leaf boxes their triangles."""
    sc = load_case("TorusMesh_flat")
    nodes, order = hostsim.bvh(sc.flatten())
    n, n4 = len(order), len(nodes)
    assert sorted(order.tolist()) == list(range(n)) and 0 < n4 < n - 1
    links = nodes[:, 24:28].copy().view(np.int32)
    NONE = 0x7fffffff
    seen_leaf, seen_node = np.zeros(n, int), np.zeros(n4, int)
    for c in links.reshape(-1):
        if c == NONE:
            continue
        if c < 0:
            first, count = (~c) & ((1 << 28) - 1), ((~c) >> 28) + 1
            assert 1 <= count <= 2
            seen_leaf[first:first + count] += 1
        else:
            seen_node[c] += 1
    assert (seen_leaf == 1).all() and seen_node[0] == 0 and (seen_node[1:] == 1).all()
    assert ((links != NONE).sum(axis=1) >= 2).all()
    # a child's box contains the boxes of that child's own children
    lo, hi = nodes[:, 0:12].reshape(n4, 3, 4), nodes[:, 12:24].reshape(n4, 3, 4)
    for i in range(n4):
        for k in range(4):
            c = links[i, k]
            if c != NONE and c >= 0:
                used = links[c] != NONE
                assert (lo[c][:, used].min(axis=1) >= lo[i][:, k] - 1e-6).all() and (hi[c][:, used].max(axis=1) <= hi[i][:, k] + 1e-6).all()


def test_cli_glue_matches_glue_py(tmp_path):
    """glue.py:17-26: strips sorted by int(stem) (10.npy after 9.npy), concatenated on axis 0, rot90, *255 truncated."""
    from python_raytracer_b200 import cli
    from PIL import Image
    rng = np.random.default_rng(3)
    full = rng.random((23, 7, 3))
    for n, part in enumerate(np.array_split(full, 11, axis=0)):
        np.save(tmp_path / f"{n}.npy", part)
    out = tmp_path / "o.png"
    assert cli.main(["glue", "--dir", str(tmp_path), "--outfile", str(out)]) == 0
    assert np.array_equal(np.array(Image.open(out)), to_png_array(full))


def test_tile_interleave_partitions_the_region():
    for (ncols, nrows, tw, th, world) in [(640, 480, 8, 8, 1), (37, 23, 8, 8, 3), (5, 3, 8, 8, 2), (64, 64, 16, 4, 8)]:
        seen = np.zeros((ncols, nrows), int)
        for rank in range(world):
            job = abi.RtRenderJob(col0=3, ncols=ncols, row0=2, nrows=nrows, tile_w=tw, tile_h=th, rank=rank, world=world)
            cols, rows, valid = hostsim.job_pixels(job)
            assert len(cols) % (tw * th) == 0
            np.add.at(seen, (cols[valid] - 3, rows[valid] - 2), 1)
        assert (seen == 1).all()


# ---- instancing of `ref` nodes and moving meshes ---------------------------------------------------------
def _instancing_scene(width=96, height=64):
    mesh = {"name": "t", "type": "mesh", "filepath": "./meshes/torus.obj", "scale": 0.8, "position": [0.0, 0.3, 0.0], "materials": []}
    ball = {"name": "b", "type": "sphere", "radius": 0.45, "position": [0.0, 0.3, 0.0], "materials": [1]}
    return {"resolution": [width, height], "AA": {"jitter": False, "samples": 1},
            "camera": {"position": [0, 3.0, 7.0], "lookAt": [0, 0.6, 0], "up": [0, 1, 0], "fov": 45.0}, "ambient": [0.1, 0.1, 0.1],
            "materials": [{"name": "a", "ID": 0, "diffuse": [0.8, 0.3, 0.2], "specular": [0.4, 0.4, 0.4], "hardness": 16},
                          {"name": "b", "ID": 1, "diffuse": [0.2, 0.6, 0.9], "specular": [0, 0, 0], "hardness": 0},
                          {"name": "c", "ID": 2, "diffuse": [0.9, 0.9, 0.3], "specular": [0.5, 0.5, 0.5], "hardness": 32, "type": "mirror", "tint": 0.5}],
            "lights": [{"name": "l", "type": "point", "colour": [1, 1, 1], "power": 0.9, "position": [3, 6, 5]},
                       {"name": "s", "type": "directional", "colour": [0.6, 0.6, 0.7], "power": 1.0, "direction": [-0.4, -1.0, -0.3]}],
            "objects": [{"name": "ground", "type": "plane", "normal": [0, 1, 0], "position": [0, -0.5, 0], "materials": [1, 2]},
                        {"name": "thing", "type": "node", "hierarchy_type": "union", "position": [-1.6, 0.8, 0.0], "rotation": [20, 30, 0],
                         "scale": [1, 1, 1], "materials": [0], "children": [mesh, ball]},
                        {"name": "copy1", "type": "node", "ref": "thing", "position": [1.7, 0.9, -0.5], "rotation": [-40, 10, 25],
                         "scale": [0.8, 1.1, 0.9], "materials": [2]},
                        {"name": "copy2", "type": "node", "ref": "thing", "position": [0.2, 2.2, -1.5], "rotation": [70, 0, 0],
                         "scale": [0.7, 0.7, 0.7], "materials": [1]}]}


def test_ref_nodes_are_instances_sharing_one_mesh_and_render_like_deep_copies():
    """scene_parser.py:190-203 deep-copies a whole subtree per `ref`; here the copies of a mesh share its arrays
    (geometry.Mesh.__deepcopy__) and `flatten` emits ONE RtMesh (one LBVH on the device) with a transform and materials per
    instance.  The image is bit-identical to the one rendered from real deep copies, and equals the oracle's."""
    rt.scene_parser.VERBOSE = False
    base = os.path.join(GOLDEN, "scenes")
    sc = rt.scene_parser.load_scene_dict(_instancing_scene(), base)
    m0, m1, m2 = (sc.objects[k].children[0] for k in (1, 2, 3))
    assert m0 is not m1 and m1.verts is m0.verts and m2.faces is m0.faces and m1.bounding_volume is m0.bounding_volume
    assert [m.name for m in m0.materials] == ["a"] and [m.name for m in m1.materials] == ["c"] and [m.name for m in m2.materials] == ["b"]
    fs = sc.flatten()
    assert fs.desc.n_meshes == 1 and fs.desc.n_objects == 4 + 3 * 2 and fs.desc.n_xforms == 3
    objs = C.cast(fs.desc.objects, C.POINTER(abi.RtObject))
    assert [objs[i].mesh for i in range(fs.desc.n_objects) if objs[i].kind == abi.KIND_MESH] == [0, 0, 0]
    rt.geometry.Mesh.SHARE_ON_COPY = False
    try:
        dup = rt.scene_parser.load_scene_dict(_instancing_scene(), base)
    finally:
        rt.geometry.Mesh.SHARE_ON_COPY = True
    fd = dup.flatten()
    assert fd.desc.n_meshes == 3 and dup.objects[2].children[0].verts is not dup.objects[1].children[0].verts
    a, ca = hostsim.render(fs, abi)
    b, cb = hostsim.render(fd, abi)
    assert np.array_equal(a, b) and ca["rays_shadow"] == cb["rays_shadow"]
    ref, _ = oracle.render(fs, abi)
    assert parity_stats(to_png_array(ref), to_png_array(a))[0] >= PARITY_FRACTION
    assert len(np.unique(to_png_array(a).reshape(-1, 3), axis=0)) > 50          # three visible instances, not an empty frame


def _moving_mesh_scene(mesh_motion):
    d = _instancing_scene(120, 80)
    d["motion"] = {"time": 1.0, "samples": 3, "final": 1}
    d["mesh_motion"] = mesh_motion
    d["objects"][1]["speed"] = [0.5, 0.0, 0.3]                      # a moving node: the speed reaches its mesh and its sphere
    d["objects"] = d["objects"][:2] + [{"name": "free", "type": "mesh", "filepath": "./meshes/torus.obj", "scale": 0.9, "flat_shaded": True,
                                        "position": [1.5, 0.8, 0.5], "materials": [0], "speed": [-0.8, 0.4, 0.0]}]
    return d


def test_moving_meshes_extension_matches_the_extended_oracle_and_is_off_by_default():
    """Reference behaviour (default): a mesh's `speed` is stored and never used (scene_parser.py:255, mesh.py:72-153), so the
    meshes stand still while the primitives beside them move.  With the extension key "mesh_motion": true they translate by
    speed * time like the primitives (simple_geometry.py:21-24): the device functions meet the mesh in its rest frame (ray
    origin shifted), the oracle moves every vertex - two formulations, same image within the tolerance."""
    rt.scene_parser.VERBOSE = False
    base = os.path.join(GOLDEN, "scenes")
    still = rt.scene_parser.load_scene_dict(_moving_mesh_scene(False), base)
    moving = rt.scene_parser.load_scene_dict(_moving_mesh_scene(True), base)
    assert still.mesh_motion is False and moving.mesh_motion is True and len(moving.vc.motion_times) == 4
    imgs = {}
    for name, sc in (("still", still), ("moving", moving)):
        fs = sc.flatten()
        assert fs.desc.mesh_motion == (1 if name == "moving" else 0)
        ref, c_ref = oracle.render(fs, abi)
        img, c_sim = hostsim.render(fs, abi)
        frac, dmax, _ = parity_stats(to_png_array(ref), to_png_array(img))
        assert frac >= PARITY_FRACTION, (name, frac, dmax)
        assert c_sim["error_flags"] == 0 and c_ref["rays_primary"] == c_sim["rays_primary"]
        plain, _ = hostsim.render(fs, abi, mesh_mode=3)               # brute-force faces + plain object loop: the BVHs only prune
        assert np.array_equal(img, plain)
        imgs[name] = to_png_array(img).astype(int)
    assert (np.abs(imgs["still"] - imgs["moving"]).max(axis=2) > 8).mean() > 0.01      # the motion is visible
    # switched off, the speeds on the meshes change nothing: same image as the scene without them
    d = _moving_mesh_scene(False)
    d["objects"][2].pop("speed")
    nospeed = rt.scene_parser.load_scene_dict(d, base)
    a, _ = hostsim.render(still.flatten(), abi)
    b, _ = hostsim.render(nospeed.flatten(), abi)
    assert np.array_equal(a, b)


def test_torch_extension_registers_the_boundary_ops():
    """The thin PyTorch extension (csrc/torch_ops.cpp): built in-tree, loads without a GPU, registers torch.ops.b200rt.*."""
    import torch
    from python_raytracer_b200 import torch_ext
    assert os.path.exists(torch_ext.LIB_PATH), "build with `python __graft_entry__.py`"
    ops = torch_ext.load()
    for name in ("render", "render_into", "finalize_u8", "assemble_tiles"):
        assert hasattr(ops, name)
    job = abi.RtRenderJob(col0=1, ncols=2, row0=3, nrows=4, tile_w=8, tile_h=8, rank=0, world=1, direct=1, mesh_mode=0, max_batch_rays=5)
    assert torch_ext.job_tensor(job).tolist() == [1, 2, 3, 4, 8, 8, 0, 1, 1, 0, 5]
    with pytest.raises(RuntimeError):             # argument checks happen before any device work
        ops.finalize_u8(torch.zeros(4, 4, 3))
