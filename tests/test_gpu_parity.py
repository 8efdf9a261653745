"""GPU parity tests (run on the B200 box with -m gpu).  Everything goes through the C ABI of
include/b200rt.h (ctypes, `python-raytracer_b200/_abi.py`): Scene.render -> rt_render_host etc.
Checker = the tier-1 oracle and the committed renders of the unmodified reference.

Tolerance (BASELINE.json north_star): uint8 images as main.py:31-33 writes them, per-channel
|delta| <= 1 on >= 99.9 % of pixels."""
import ctypes as C

import numpy as np
import pytest

import oracle
import python_raytracer_b200 as rt
from conftest import (ALL_CASES, FUZZ_CHAOTIC_SEEDS, FUZZ_GPU_SEEDS, PARITY_FRACTION, golden_png, load_bundled, load_case, load_fuzz,
                      parity_stats, residue_report, to_png_array, u8_cr, unstable_fraction)

pytestmark = pytest.mark.gpu
abi = rt._abi


@pytest.fixture(scope="module")
def lib():
    return abi.load_library()


@pytest.mark.parametrize("case", ALL_CASES)
def test_render_matches_reference_render(case):
    """Scene.render() vs the committed render of the UNMODIFIED reference for the same scene and samples."""
    sc = load_case(case)
    img = sc.render()
    assert img.dtype == np.float64 and img.shape == (sc.vc.width, sc.vc.height, 3)
    assert img.min() >= 0.0 and img.max() <= 1.0
    frac, dmax, d = parity_stats(to_png_array(img), golden_png(case))
    assert frac >= PARITY_FRACTION, (case, frac, dmax)
    # the oracle's ray tallies for the same frame (classes of SURVEY 8d)
    _, c_ref = oracle.render(sc.flatten(), abi)
    c = sc.last_counters
    assert c["rays_primary"] == c_ref["rays_primary"]
    for k in ("rays_shadow", "rays_reflect", "rays_refract", "shaded_hits"):
        assert abs(c[k] - c_ref[k]) <= max(8, 5e-4 * c_ref[k]), (k, c[k], c_ref[k])
    assert c["error_flags"] == 0


@pytest.mark.parametrize("seed", FUZZ_GPU_SEEDS)
def test_random_scene_matches_oracle(seed):
    """Fresh random scenes (tests/fuzz_scenes.py: nested CSG, `ref` copies, meshes in nodes, moving objects, textures, DOF /
    motion / jitter) against the oracle, which reproduces the reference float-exactly on 43 + 5 scenes of the same generator
    (oracle/fuzz_check.py, tests/golden Fuzz*).  Also: BVHs and child-bounds pruning leave the image bit-identical."""
    sc = load_fuzz(seed)
    img = sc.render()
    ref, c_ref = oracle.render(sc.flatten(), abi)
    frac, dmax, _ = parity_stats(to_png_array(img), to_png_array(ref))
    assert frac >= PARITY_FRACTION, (seed, frac, dmax)
    c = sc.last_counters
    assert c["rays_primary"] == c_ref["rays_primary"] and c["error_flags"] == 0
    for k in ("rays_shadow", "rays_reflect", "rays_refract", "shaded_hits"):
        assert abs(c[k] - c_ref[k]) <= max(8, 1e-3 * c_ref[k]), (k, c[k], c_ref[k])
    plain = sc.render_region(0, sc.vc.width, 0, sc.vc.height, mesh_mode=3)
    # identical, except where a far-away ray (ground plane hit at the horizon, t ~ 1e4) makes a primitive's float32 test report
    # a hit that is geometrically impossible (cancellation in the sphere discriminant): the conservative boxes skip that
    # primitive, the plain loop evaluates it.  Seen on 4 of 12288 pixels of seed 123, colour difference 5e-5.
    differ = np.abs(np.asarray(img, dtype=np.float32) - plain).max(axis=2) > 0
    assert differ.mean() <= 1e-3 and np.abs(np.asarray(img, dtype=np.float32) - plain).max() < 1e-3


def test_render_is_deterministic_and_strips_concatenate():
    sc = load_case("MirrorRefraction")
    a = sc.render()
    b = sc.render()
    assert np.array_equal(a, b)                       # queue order varies, sample slots do not
    strips = [sc.render(n, 7) for n in range(7)]      # render.nu:10-13 + glue.py:22
    assert [s.shape[0] for s in strips] == [len(x) for x in np.array_split(np.arange(900), 7)]
    assert np.array_equal(np.concatenate(strips, axis=0), a)


def test_more_tasks_than_columns_gives_empty_strips():
    sc = load_case("TorusMesh_flat", {"resolution": [4, 6]})
    parts = [sc.render(n, 6) for n in range(6)]
    assert [p.shape for p in parts] == [(1, 6, 3)] * 4 + [(0, 6, 3)] * 2


def test_small_batches_equal_one_batch():
    sc = load_case("DepthOfField")
    full = sc.render_region(0, 256, 0, 256).copy()
    h = sc.upload()
    out = np.empty((256, 256, 3), np.float32)
    job = sc.make_job(0, 256, 0, 256, max_batch_rays=100_000)
    ctr = abi.RtCounters()
    abi.check(sc._lib, sc._lib.rt_render_host(h, C.byref(job), out.ctypes.data, C.byref(ctr), None), "rt_render_host")
    assert np.array_equal(out, full)
    assert ctr.kernel_launches > 20


def test_batch_size_does_not_change_a_bit_with_recursive_materials():
    """Mirror / refraction chains of depth 10 (queued levels 1-2, k_tail from level 3): the frame must not depend on how it is
    cut into wavefront batches - the property that makes an N-rank render equal the 1-rank render (smaller batches per rank)."""
    sc = load_case("MirrorRefraction")
    full = sc.render_region(0, 900, 0, 512).copy()
    assert sc.last_counters["rays_refract"] > 0 and sc.last_counters["rays_reflect"] > 0
    h = sc.upload()
    for batch in (150_000, 20_000, 1_000):
        out = np.empty((900, 512, 3), np.float32)
        job = sc.make_job(0, 900, 0, 512, max_batch_rays=batch)
        ctr = abi.RtCounters()
        abi.check(sc._lib, sc._lib.rt_render_host(h, C.byref(job), out.ctypes.data, C.byref(ctr), None), "rt_render_host")
        assert np.array_equal(out, full), batch
    win = sc.render_region(300, 64, 200, 48)                       # a window rendered alone == the same pixels of the frame
    assert np.array_equal(win, full[300:364, 200:248])


def test_render_host_into_pinned_memory_leaves_in_chunks_and_equals_the_device_render(lib):
    """rt_render_host with a PINNED destination copies what every wavefront batch completes while the next batches render (whole
    tile columns of the reference layout / the batch's slots of the compact layout): same bytes as the pageable path and as the
    device render, also when batch boundaries fall inside tiles and tile columns."""
    import torch
    sc = load_case("MirrorRefraction")
    W, H = 900, 512
    ref = sc.render_region(0, W, 0, H).copy()                      # pageable numpy destination: one copy at the end
    h = sc.upload()
    for direct, batch in ((True, 37_000), (True, 0), (False, 53_000)):
        job = sc.make_job(0, W, 0, H, direct=direct, max_batch_rays=batch)
        nt, nf = C.c_int64(), C.c_int64()
        abi.check(lib, lib.rt_job_extent(C.byref(job), C.byref(nt), C.byref(nf)), "rt_job_extent")
        host = torch.full((nf.value,), -1.0, dtype=torch.float32).pin_memory()
        ctr = abi.RtCounters()
        abi.check(lib, lib.rt_render_host(h, C.byref(job), host.data_ptr(), C.byref(ctr), None), "rt_render_host")
        dev = torch.full((nf.value,), -2.0, dtype=torch.float32, device="cuda")
        abi.check(lib, lib.rt_render(h, C.byref(job), dev.data_ptr(), None, None), "rt_render")
        torch.cuda.synchronize()
        assert torch.equal(host, dev.cpu()), (direct, batch)
        if direct:
            assert np.array_equal(host.numpy().reshape(W, H, 3), ref)
        assert ctr.error_flags == 0 and ctr.rays_primary == W * H


@pytest.mark.parametrize("case,over", [("MirrorRefraction", {"resolution": [96, 64]}), ("TorusMesh_smooth", {"resolution": [64, 64]}),
                                       ("NovelScene1_small", {"resolution": [64, 32]}), ("MeshInNode", {"resolution": [64, 48]})])
def test_every_kernel_variant_in_many_small_batches_on_two_streams(case, over, lib):
    """The three kernel variants (lean / nodes without a mesh / nodes and meshes) with the frame cut into dozens of wavefront
    batches: both work-buffer sets, level 0 on the caller's stream and the deeper levels on the side stream, the chunked copy
    into pinned host memory - all must give the bytes of the one-batch render."""
    import torch
    sc = load_case(case, over)
    W, H = sc.vc.width, sc.vc.height
    ref = sc.render_region(0, W, 0, H).copy()
    h = sc.upload()
    for batch in (1500, 7000):
        job = sc.make_job(0, W, 0, H, max_batch_rays=batch)
        host = torch.full((W * H * 3,), -1.0, dtype=torch.float32).pin_memory()
        ctr = abi.RtCounters()
        abi.check(lib, lib.rt_render_host(h, C.byref(job), host.data_ptr(), C.byref(ctr), None), "rt_render_host")
        assert np.array_equal(host.numpy().reshape(W, H, 3), ref), (case, batch)
        assert ctr.error_flags == 0 and (batch > 1500 or ctr.kernel_launches >= 8)      # several batches


@pytest.mark.parametrize("world", [2, 3, 8])
def test_tile_sharded_render_assembles_to_the_single_rank_image(world, lib):
    """Multi-GPU split emulated on one device: every rank's compact tile buffer, concatenated as the
    NCCL gather delivers them, scattered by rt_assemble_tiles == the 1-rank image, bit for bit."""
    import torch
    sc = load_case("MirrorRefraction", {"resolution": [203, 117]})
    ref = sc.render_region(0, 203, 0, 117)
    h = sc.upload()
    bufs, offs, total = [], [], 0
    for rank in range(world):
        job = sc.make_job(0, 203, 0, 117, rank=rank, world=world, direct=False)
        nt, nf = C.c_int64(), C.c_int64()
        abi.check(lib, lib.rt_job_extent(C.byref(job), C.byref(nt), C.byref(nf)), "rt_job_extent")
        buf = torch.zeros(nf.value, dtype=torch.float32, device="cuda")
        sc.render_device(buf.data_ptr(), job)
        bufs.append(buf); offs.append(total); total += nf.value
    gathered = torch.cat(bufs)
    out = torch.empty(203 * 117 * 3, dtype=torch.float32, device="cuda")
    job = sc.make_job(0, 203, 0, 117, rank=0, world=world, direct=False)
    off_arr = (C.c_int64 * world)(*offs)
    abi.check(lib, lib.rt_assemble_tiles(C.byref(job), gathered.data_ptr(), off_arr, out.data_ptr(), None), "rt_assemble_tiles")
    torch.cuda.synchronize()
    assert np.array_equal(out.cpu().numpy().reshape(203, 117, 3), ref)


def test_finalize_u8_equals_main_py_conversion(lib):
    import torch
    sc = load_case("TwoSpheresPlane")
    img = sc.render_region(0, 640, 0, 480)
    d_img = torch.from_numpy(img).cuda()
    d_u8 = torch.empty((480, 640, 3), dtype=torch.uint8, device="cuda")
    abi.check(lib, lib.rt_finalize_u8(d_img.data_ptr(), 640, 480, d_u8.data_ptr(), None), "rt_finalize_u8")
    torch.cuda.synchronize()
    assert np.array_equal(d_u8.cpu().numpy(), to_png_array(img))


def test_render_u8_and_cli_equal_main_py_output(tmp_path):
    """Device-side rot90 + *255 truncation (rt_render_u8_host) == main.py:31-33 applied to render(); the CLI writes that PNG,
    and strips saved by `main --subimage` glue back to the same image (render.nu:10-15)."""
    from PIL import Image
    from python_raytracer_b200 import cli
    import os
    sc = load_case("MotionBlur")
    want = to_png_array(sc.render())
    assert np.array_equal(sc.render_u8(), want)
    assert np.array_equal(np.concatenate([sc.render_u8(n, 3) for n in range(3)], axis=1), want)
    scene_json = os.path.join(os.path.dirname(__file__), "golden", "scenes", "TwoSpheresPlane.json")
    png = tmp_path / "a.png"
    assert cli.main(["main", "--infile", scene_json, "--outfile", str(png)]) == 0
    assert np.array_equal(np.array(Image.open(png)), golden_png("TwoSpheresPlane")) or \
        parity_stats(np.array(Image.open(png)), golden_png("TwoSpheresPlane"))[0] >= PARITY_FRACTION
    os.makedirs(tmp_path / "temp")
    for n in range(4):
        assert cli.main(["main", "--infile", scene_json, "--outfile", str(tmp_path / "temp" / str(n)), "--subimage", str(n), "--tasks", "4"]) == 0
    assert cli.main(["glue", "--dir", str(tmp_path / "temp"), "--outfile", str(tmp_path / "g.png")]) == 0
    assert np.array_equal(np.array(Image.open(tmp_path / "g.png")), np.array(Image.open(png)))
    # the whole render.nu pipeline in one call: strips -> temp -> glue -> temp removed
    assert cli.main(["render", scene_json, "5", str(tmp_path / "r.png"), "--temp", str(tmp_path / "temp5")]) == 0
    assert np.array_equal(np.array(Image.open(tmp_path / "r.png")), np.array(Image.open(png)))
    assert not os.path.exists(tmp_path / "temp5")


def test_device_mesh_ingestion_is_bit_identical_to_the_host_loader(tmp_path):
    """rt_mesh_ingest (transform, face-ordered area-weighted vertex normals, extent, gate radius on the GPU) + rt_obj_load
    against the numpy loader that the golden fixtures pin to the reference (mesh.py:17-70)."""
    import os
    from python_raytracer_b200 import geometry, synthetic
    torus = os.path.join(os.path.dirname(__file__), "golden", "meshes", "torus.obj")
    verts, faces = synthetic.displaced_torus(256, 128, 7)
    big = tmp_path / "big.obj"
    with open(big, "w") as fh:
        fh.write("".join(f"v {x!r} {y!r} {z!r}\n" for x, y, z in verts.tolist()))
        fh.write("".join(f"f {a + 1} {b + 1} {c + 1}\n" for a, b, c in faces.tolist()))
    lonely = (np.concatenate([verts[:100], [[9.0, 9.0, 9.0]]]), faces[(faces < 100).all(axis=1)])   # a vertex no face uses
    for src, t, s in ((torus, (0.0, 0.0, 0.0), 1.0), (torus, (0.3, -1.25, 2.0), 0.37), (str(big), (0.1, 0.2, -0.3), 1.7),
                      ((verts, faces), (0.0, 1.5, 0.0), 1.0), (lonely, (0.0, 0.0, 0.0), 2.0)):
        for flat in (False, True):
            h = geometry.Mesh("m", "mesh", [], t, s, src, flat, None, ingest="host")
            d = geometry.Mesh("m", "mesh", [], t, s, src, flat, None, ingest="device")
            assert np.array_equal(h.faces, d.faces)
            assert h.verts.dtype == d.verts.dtype and np.array_equal(h.verts, d.verts)
            if flat:
                assert h.norms is None and d.norms is None
            else:
                assert np.array_equal(h.norms, d.norms, equal_nan=True)
                assert np.isnan(d.norms).any() == (src is lonely)
            assert type(h.bounding_volume) is type(d.bounding_volume)
            for k, v in vars(h.bounding_volume).items():
                if v is None:
                    continue
                assert np.array_equal(np.asarray(v, dtype=np.float64), np.asarray(getattr(d.bounding_volume, k), dtype=np.float64)), k
    # the whole scene through the parser
    a, b = load_case("TorusMesh_smooth"), load_case("TorusMesh_smooth", ingest="device")
    assert np.array_equal(a.render(), b.render())


def test_bvh_traversal_equals_brute_force_on_device():
    for case in ("TorusMesh_flat", "TorusMesh_smooth"):
        sc = load_case(case)
        a = sc.render_region(0, 256, 0, 256, mesh_mode=0).copy()
        nodes = sc.last_counters["bvh_nodes"]
        b = sc.render_region(0, 256, 0, 256, mesh_mode=1)
        assert np.array_equal(a, b)
        assert nodes > 0 and sc.last_counters["bvh_nodes"] == 0


@pytest.mark.parametrize("case", ["TwoSpheresPlane", "MotionBlur", "MirrorRefraction", "TorusMesh_smooth", "NovelScene1_small"])
def test_object_bvh_equals_linear_object_loop_on_device(case):
    """mesh_mode bit 1 = plain in-order loop over scene.objects (scene.py:87,160); the object BVH must not change a bit."""
    sc = load_case(case)
    W, H = sc.vc.width, sc.vc.height
    a = sc.render_region(0, W, 0, H, mesh_mode=0).copy()
    ca = dict(sc.last_counters)
    b = sc.render_region(0, W, 0, H, mesh_mode=2)
    cb = dict(sc.last_counters)
    assert np.array_equal(a, b)
    for k in ("rays_primary", "rays_shadow", "rays_reflect", "rays_refract", "shaded_hits"):
        assert ca[k] == cb[k], k


def test_gpu_lbvh_equals_host_restatement(lib):
    """Radix sort + PLOC clustering + 4-wide collapse on the GPU vs the same RT_HD functions run sequentially on the host."""
    import hostsim
    sc = load_case("TorusMesh_flat")
    h = sc.upload()
    nodes_ref, order_ref = hostsim.bvh(sc.flatten())
    n4, nt, nb = C.c_int64(), C.c_int64(), C.c_int64()
    abi.check(lib, lib.rt_scene_bvh_stats(h, 0, C.byref(n4), C.byref(nt), C.byref(nb)), "rt_scene_bvh_stats")
    assert n4.value == len(nodes_ref) and nt.value == len(order_ref)
    nodes = np.zeros_like(nodes_ref); order = np.zeros_like(order_ref)
    abi.check(lib, lib.rt_scene_bvh_download(h, 0, nodes.ctypes.data, order.ctypes.data), "rt_scene_bvh_download")
    assert np.array_equal(order, order_ref)
    assert np.array_equal(nodes.view(np.int32), nodes_ref.view(np.int32))     # boxes, links and flags bit for bit


def _random_rays(rng, n, centre, spread=6.0):
    o = centre + rng.normal(size=(n, 3)) * spread
    target = centre + rng.normal(size=(n, 3)) * 1.5
    d = target - o
    d *= rng.uniform(0.2, 3.0, size=(n, 1)) / np.linalg.norm(d, axis=1, keepdims=True)   # un-normalised, like shadow rays
    return o.astype(np.float32), d.astype(np.float32)


@pytest.mark.parametrize("case,centre", [("TwoSpheresPlane", (0, 1, -0.5)), ("DepthOfField", (0, 1, -2)),
                                         ("TorusMesh_smooth", (0, 0, 0)), ("NovelScene1_small", (0, 1, 0)),
                                         ("MirrorRefraction", (0, 1.5, 0))])
def test_cast_ray_and_first_hit_on_random_rays(case, centre):
    """Per-ray unit test: closest hit record and cast_ray colour for rays that are NOT camera rays."""
    sc = load_case(case)
    rng = np.random.default_rng(1234)
    n = 20000
    o, d = _random_rays(rng, n, np.array(centre, dtype=np.float64))
    times = rng.choice(np.asarray(sc.vc.motion_times, dtype=np.float64), size=n)
    ref = oracle.trace_rays(sc.flatten(), o, d, times)
    got = sc.trace_rays(o, d, times.astype(np.float32))
    hit_ref, hit_got = np.isfinite(ref["t"]), np.isfinite(got["t"])
    assert (hit_ref == hit_got).mean() >= 0.999
    both = hit_ref & hit_got
    same_geom = got["geom"][both] == ref["geom"][both]
    assert same_geom.mean() >= 0.999
    sel = np.flatnonzero(both)[same_geom]
    assert np.allclose(got["t"][sel], ref["t"][sel], rtol=2e-4, atol=2e-4)
    assert (got["mat"][sel] == ref["mat"][sel]).mean() >= 0.999
    close_n = np.abs(got["n"][sel] - ref["n"][sel]).max(axis=1) < 2e-3
    assert close_n.mean() >= 0.998
    close_rgb = np.abs(got["rgb"] - ref["rgb"]).max(axis=1) <= 1.0 / 255.0
    assert close_rgb.mean() >= 0.995, close_rgb.mean()


@pytest.mark.parametrize("case,centre", [("TwoSpheresPlane", (0, 1, -0.5)), ("TorusMesh_flat", (0, 0, 0)),
                                         ("NovelScene2_small", (0, 1, 0))])
def test_shadow_any_hit_on_random_rays(case, centre):
    sc = load_case(case)
    rng = np.random.default_rng(99)
    n = 20000
    o, d = _random_rays(rng, n, np.array(centre, dtype=np.float64))
    tmax = np.where(rng.random(n) < 0.5, 1.0, np.inf)
    ref = oracle.occluded(sc.flatten(), o, d, tmax)
    got = sc.occluded(o, d, tmax.astype(np.float32))
    assert (ref == got).mean() >= 0.9995


def test_missing_material_raises_index_error_like_the_reference():
    # a hit sphere without materials: `self.materials[0]` -> IndexError (simple_geometry.py:44)
    sc = load_case("TwoSpheresPlane", {"resolution": [32, 24]})
    sc.objects[1].materials = []
    sc.invalidate()
    with pytest.raises(IndexError):
        sc.render()


def test_synthetic_mesh_scene_sampled_parity_and_bvh_equivalence():
    """Config 5 shape at reduced size: displaced-sphere mesh + spheres (diffuse / mirror / refractive) + checker
    plane; full-image GPU LBVH == GPU brute force, and a pixel sample against the (brute-force) oracle."""
    from python_raytracer_b200 import synthetic
    sc = synthetic.make_scene(width=160, height=90, aa_samples=2, grid=(64, 32), n_spheres=16)
    a = sc.render_region(0, 160, 0, 90, mesh_mode=0).copy()
    b = sc.render_region(0, 160, 0, 90, mesh_mode=1)
    assert np.array_equal(a, b)
    c = sc.render_region(0, 160, 0, 90, mesh_mode=3)       # brute-force faces AND plain object loop: the reference's own loops
    assert np.array_equal(a, c)
    ref, _ = oracle.render(sc.flatten(), abi)
    frac, dmax, _ = parity_stats(to_png_array(ref), to_png_array(a))
    assert frac >= PARITY_FRACTION, (frac, dmax)


def test_config5_full_size_4k_64spp_1m_triangles():
    """BASELINE.json configs[4] at FULL size (3840x2160 x 64 spp, 1 048 576 triangles + 64 spheres: the bench workload),
    through size-independent properties: ray tallies, a sub-rectangle rendered alone == the same pixels of the frame
    (counter-based jitter, tile/batch independence), LBVH + object BVH == the reference's own loops (brute force over
    every face, in-order object list) on a 16x16-pixel window on the mesh, and 320 pixels (all 64 samples each; 200 of them
    on the highest-contrast edges of the frame) against the brute-force oracle."""
    from python_raytracer_b200 import synthetic
    sc = synthetic.make_scene(ingest="device")
    W, H = 3840, 2160
    assert (sc.vc.width, sc.vc.height, sc.samples) == (W, H, 64) and sc.objects[1].faces.shape[0] == 1048576
    img = sc.render_region(0, W, 0, H)
    c = dict(sc.last_counters)
    assert c["rays_primary"] == W * H * 64 and c["error_flags"] == 0
    assert c["rays_shadow"] == 3 * c["shaded_hits"] and c["rays_reflect"] > 0 and c["rays_refract"] > 0
    assert np.isfinite(img).all() and img.min() >= 0.0 and img.max() <= 1.0
    c0, r0, n = 1912, 1072, 16
    win = sc.render_region(c0, n, r0, n, mesh_mode=0).copy()
    assert np.array_equal(win, img[c0:c0 + n, r0:r0 + n])
    assert sc.last_counters["tri_tests"] > 0                      # the window is on the mesh
    brute = sc.render_region(c0, n, r0, n, mesh_mode=3)
    assert sc.last_counters["tri_tests"] >= 1048576 * n * n * 64
    assert np.array_equal(win, brute)
    # oracle sample (brute force over 1 M faces per ray: ~0.5 s per pixel on 16 cores): 120 uniformly random pixels + 200 drawn
    # from the 1 % of the frame with the largest local contrast (sphere rims, mesh silhouette, shadow and checker edges, the
    # reflections in the spheres) - the places where a residue would be.  All 64 samples of each pixel.
    u8 = u8_cr(img)
    grad = np.zeros((W, H))
    grad[1:-1, 1:-1] = np.maximum(np.abs(u8[2:, 1:-1] - u8[:-2, 1:-1]).max(axis=2), np.abs(u8[1:-1, 2:] - u8[1:-1, :-2]).max(axis=2))
    rng = np.random.default_rng(11)
    edge_px = np.argwhere(grad >= np.quantile(grad, 0.99))
    pick = edge_px[rng.choice(len(edge_px), 200, replace=False)]
    cols = np.concatenate([rng.integers(0, W, 120), pick[:, 0]])
    rows = np.concatenate([rng.integers(0, H, 120), pick[:, 1]])
    ref, _ = oracle.render_pixels(sc.flatten(), abi, cols, rows)
    a, b = u8_cr(img[cols, rows]), u8_cr(ref)
    bad = np.abs(a - b).max(axis=1) > 1
    print(f"config 5 oracle sample: random 120: {int(bad[:120].sum())} off, edge-biased 200: {int(bad[120:].sum())} off, max delta {np.abs(a - b).max()}")
    assert bad[:120].sum() <= 1, (int(bad[:120].sum()), np.abs(a - b).max())     # uniformly random pixels: the 99.9 % statistic
    assert bad[120:].sum() <= 4, (int(bad[120:].sum()), np.abs(a - b).max())     # edge pixels only: 98 % of them within 1/255


@pytest.mark.parametrize("case,over", [
    ("DepthOfField", {"resolution": [1920, 1080], "DOF": {"samples": 64}, "AA": {"jitter": True, "samples": 1}}),
    ("MotionBlur", {"resolution": [1920, 1080], "motion": {"time": 2.0, "samples": 63, "final": 1}}),
])
def test_config4_1080p_64spp_sampled_parity(case, over):
    """BASELINE.json configs[3]: thin-lens DOF / motion blur at 1920x1080 x 64 samples (132.7 M primary rays, several
    wavefront batches).  Full frame on the GPU; parity on 3000 random pixels (all 64 samples each) against the oracle."""
    sc = load_case(case, over)
    assert sc.vc.dof_samples * sc.samples * len(sc.vc.motion_times) == 64
    img = sc.render_region(0, 1920, 0, 1080)
    c = dict(sc.last_counters)
    assert c["rays_primary"] == 1920 * 1080 * 64 and c["error_flags"] == 0
    rng = np.random.default_rng(5)
    cols, rows = rng.integers(0, 1920, 3000), rng.integers(0, 1080, 3000)
    ref, _ = oracle.render_pixels(sc.flatten(), abi, cols, rows)
    a = (np.asarray(img[cols, rows], dtype=np.float64) * 255).astype(np.uint8).astype(int)
    b = (ref * 255).astype(np.uint8).astype(int)
    frac = float((np.abs(a - b).max(axis=1) <= 1).mean())
    assert frac >= PARITY_FRACTION, (case, frac)


# ---- round 2: parity holes named by the round-1 review -------------------------------------------------------------------
@pytest.mark.parametrize("name,n_px", [("NovelScene1", 20000), ("NovelScene2", 4000)])
def test_bundled_csg_scenes_at_shipped_settings(name, n_px):
    """The two bundled CSG scenes at their SHIPPED settings (scenes/NovelScene1.json:2-3: 2048x1024 x 32 spp, jitter;
    NovelScene2.json:2-5: 1024x512 x AA 2 x DOF 15 x 16 motion times = 480 samples): nesting depth 6, `ref` copies, textures,
    inverted boxes, moving CSG.  Full frame on the GPU; parity on random pixels (all samples each) against the oracle."""
    sc = load_bundled(name)
    W, H = sc.vc.width, sc.vc.height
    N = sc.vc.dof_samples * sc.samples * len(sc.vc.motion_times)
    assert (W, H, N) == ((2048, 1024, 32) if name == "NovelScene1" else (1024, 512, 480))
    img = sc.render_region(0, W, 0, H)
    c = dict(sc.last_counters)
    assert c["rays_primary"] == W * H * N and c["error_flags"] == 0
    rng = np.random.default_rng(5)
    cols, rows = rng.integers(0, W, n_px), rng.integers(0, H, n_px)
    ref, _ = oracle.render_pixels(sc.flatten(), abi, cols, rows)
    d = np.abs(u8_cr(img[cols, rows]) - u8_cr(ref)).max(axis=1)
    print(f"{name} shipped: {n_px} pixels, {(d > 1).sum()} off by more than 1/255, max {d.max()}")
    assert (d <= 1).mean() >= PARITY_FRACTION, (name, (d <= 1).mean(), d.max())


@pytest.mark.parametrize("name", ["DepthOfField", "NovelScene1", "NovelScene2"])
def test_statistical_match_with_the_authors_jittered_renders(name):
    """External evidence that does not pass through the glm / igl shims or the patched RNG: the AUTHOR's stored renders of the
    three scenes with unseeded jitter (`/root/reference/renders/{DepthOfField,NovelScene1,NovelScene2}.png`, SURVEY section 4:
    "statistical comparison only") against the GPU render of the same scene file at its shipped settings with this repo's own
    jitter stream.  Statistic: mean uint8 value of every 8x8-pixel block (fixtures: tests/golden/author_stats.npz, written by
    oracle/make_author_stats.py in the dev container).  The tier-1 oracle scores mean |delta| 0.04 - 0.06, max 1.2 - 1.5 of 255."""
    import os
    from conftest import GOLDEN
    want = np.load(os.path.join(GOLDEN, "author_stats.npz"))[name]
    sc = load_bundled(name)
    png = sc.render_u8()                                              # (H, W, 3) uint8, main.py:31-33 on the device
    H, W, _ = png.shape
    got = png.astype(np.float32).reshape(H // 8, 8, W // 8, 8, 3).mean(axis=(1, 3))
    assert got.shape == want.shape
    d = np.abs(got - want).max(axis=2)
    assert d.mean() <= 0.25 and (d <= 2.0).mean() >= 0.999 and d.max() <= 6.0, (name, float(d.mean()), float(d.max()), float((d <= 2.0).mean()))


@pytest.mark.parametrize("seed", FUZZ_CHAOTIC_SEEDS)
def test_chaotic_random_scenes_match_on_reference_stable_pixels(seed):
    """The 13 random scenes round 1 screened out (the reference's CSG-difference shadows are a coin toss on some surfaces,
    hierarchy.py:94-107 + simple_geometry.py:223; 6 of the 13 miss a plain 99.9 % threshold): every out-of-tolerance pixel is
    classified with the oracle's stability probe (conftest.residue_report); the statistic over the reference-STABLE pixels
    must meet the tolerance and no residue pixel may be unexplained."""
    sc = load_fuzz(seed)
    fs = sc.flatten()
    img = sc.render()
    ref, _ = oracle.render(fs, abi)
    rep = residue_report(fs, ref, img)
    print(f"seed {seed}: {rep}")
    assert rep["frac_stable"] >= PARITY_FRACTION and rep["unexplained"] == 0, rep
    assert unstable_fraction(fs, ref) <= 0.10      # the probe calls 0.5-6 % of these frames unstable, not everything


@pytest.mark.parametrize("case", ALL_CASES)
def test_residue_lies_on_edges_and_refraction_regions(case):
    """SURVEY 8d parity statistic: the pixels outside 1/255 are reference-unstable, or on silhouettes / material edges /
    shadow edges, or in mirror / refraction regions - never in the interior of a diffuse surface."""
    sc = load_case(case)
    fs = sc.flatten()
    img = sc.render()
    ref, _ = oracle.render(fs, abi)
    rep = residue_report(fs, ref, img)
    print(f"{case}: {rep}")
    assert rep["frac"] >= PARITY_FRACTION and rep["unexplained"] == 0, rep


def test_gpu_lbvh_equals_host_restatement_on_a_large_mesh_with_equal_keys(lib):
    """The GPU build (63-bit Morton keys, 8-pass radix sort over 33 tiles, PLOC clustering rounds with their prefix sums and index
    tie-breaks, depth-first slot assignment, 4-wide collapse) against the sequential host restatement on 135 168 triangles of which 4 096 are exact duplicates
    (equal centroids -> equal Morton keys -> the stable sort and the index tie-break decide) and 128 are degenerate."""
    import hostsim
    from python_raytracer_b200 import geometry, synthetic
    verts, faces = synthetic.displaced_torus(512, 128, 3)
    assert faces.shape[0] == 131072
    rng = np.random.default_rng(9)
    dup = faces[rng.choice(len(faces), 4096, replace=False)]
    dup[::2] = dup[::2][:, [1, 2, 0]]                                    # same triangle, rotated corner order: same centroid up to rounding
    degenerate = np.repeat(rng.integers(0, len(verts), (128, 1)), 3, axis=1).astype(faces.dtype)   # zero-area: three times one vertex
    faces = np.concatenate([faces, dup, degenerate])
    mesh = geometry.Mesh("m", "mesh", [rt.helperclasses.Material("d", (0, 0, 0), (0.8, 0.8, 0.8), 0, 0)], (0.0, 0.0, 0.0), 1.0,
                         (verts, faces), True, None)
    sc = load_case("TorusMesh_flat", {"resolution": [64, 64]})
    sc.objects[1] = mesh
    sc.invalidate()
    h = sc.upload()
    nodes_ref, order_ref = hostsim.bvh(sc.flatten())
    n4, nt, nb = C.c_int64(), C.c_int64(), C.c_int64()
    abi.check(lib, lib.rt_scene_bvh_stats(h, 0, C.byref(n4), C.byref(nt), C.byref(nb)), "rt_scene_bvh_stats")
    assert nt.value == len(faces) == len(order_ref) and n4.value == len(nodes_ref)
    nodes = np.zeros_like(nodes_ref); order = np.zeros_like(order_ref)
    abi.check(lib, lib.rt_scene_bvh_download(h, 0, nodes.ctypes.data, order.ctypes.data), "rt_scene_bvh_download")
    # the slots follow a depth-first walk of the PLOC tree built over the Morton order (rt_bvh.cuh), so the tree decides them; what
    # the sort contributes - stable order of the 63-bit keys, ties included - shows in the tree: one swapped pair of tied leaves
    # gives another neighbourhood and another tree
    assert sorted(order_ref.tolist()) == list(range(len(faces))), "host leaf order is not a permutation"
    assert sorted(order.tolist()) == list(range(len(faces))), "GPU leaf order is not a permutation"
    wrong = np.flatnonzero(order != order_ref)
    assert len(wrong) == 0, f"{len(wrong)} of {len(order)} leaf slots differ; first: {wrong[:8]}, gpu {order[wrong[:8]]}, host {order_ref[wrong[:8]]}"
    assert np.array_equal(nodes.view(np.int32), nodes_ref.view(np.int32))
    img = sc.render_region(0, 64, 0, 64)                                 # and the tree is usable
    assert sc.last_counters["error_flags"] == 0 and img.max() > 0


def test_more_top_level_meshes_than_the_parameter_bank_holds():
    """11 top-level meshes (> RT_CONST_MESH = 8): the scene runs on the general kernels (variant 2), mesh headers from global
    memory; against the oracle, BVH == brute force, several batches == one."""
    from conftest import many_meshes_scene
    sc = many_meshes_scene()
    W, H = sc.vc.width, sc.vc.height
    img = sc.render_region(0, W, 0, H).copy()
    ref, c_ref = oracle.render(sc.flatten(), abi)
    frac, dmax, _ = parity_stats(to_png_array(img), to_png_array(ref))
    assert frac >= PARITY_FRACTION, (frac, dmax)
    c = sc.last_counters
    assert c["error_flags"] == 0 and c["rays_primary"] == c_ref["rays_primary"] and c["tri_tests"] > 0
    assert np.array_equal(sc.render_region(0, W, 0, H, mesh_mode=3), img)
    h = sc.upload()
    out = np.empty((W, H, 3), np.float32)
    job = sc.make_job(0, W, 0, H, max_batch_rays=1000)
    abi.check(sc._lib, sc._lib.rt_render_host(h, C.byref(job), out.ctypes.data, None, None), "rt_render_host")
    assert np.array_equal(out, img)


def test_ref_instances_share_one_lbvh_and_moving_meshes_extension():
    """SURVEY 8f rank 3 on the device: (a) the three instances of a `ref`-copied node share ONE RtMesh / LBVH and render exactly
    like real deep copies; (b) with "mesh_motion": true the meshes translate by speed * time (ray taken into the mesh's rest
    frame; the oracle moves the vertices), off by default = reference behaviour (meshes stand still)."""
    import os
    from conftest import GOLDEN
    from test_host_logic import _instancing_scene, _moving_mesh_scene
    base = os.path.join(GOLDEN, "scenes")
    sc = rt.scene_parser.load_scene_dict(_instancing_scene(192, 128), base)
    assert sc.flatten().desc.n_meshes == 1
    img = sc.render()
    rt.geometry.Mesh.SHARE_ON_COPY = False
    try:
        dup = rt.scene_parser.load_scene_dict(_instancing_scene(192, 128), base)
    finally:
        rt.geometry.Mesh.SHARE_ON_COPY = True
    assert dup.flatten().desc.n_meshes == 3
    assert np.array_equal(img, dup.render())
    ref, _ = oracle.render(sc.flatten(), abi)
    assert parity_stats(to_png_array(img), to_png_array(ref))[0] >= PARITY_FRACTION
    for on in (False, True):
        mv = rt.scene_parser.load_scene_dict(_moving_mesh_scene(on), base)
        got = mv.render()
        assert mv.last_counters["error_flags"] == 0
        want, c_ref = oracle.render(mv.flatten(), abi)
        frac, dmax, _ = parity_stats(to_png_array(got), to_png_array(want))
        assert frac >= PARITY_FRACTION, (on, frac, dmax)
        assert np.array_equal(np.asarray(got, dtype=np.float32), mv.render_region(0, mv.vc.width, 0, mv.vc.height, mesh_mode=3))


def test_torch_ops_boundary_equals_the_ctypes_boundary():
    """torch.ops.b200rt.* (thin PyTorch extension over the C ABI): tensors on the current stream, same pixels."""
    import torch
    from python_raytracer_b200 import torch_ext
    ops = torch_ext.load()
    sc = load_case("MirrorRefraction", {"resolution": [225, 128]})
    want = sc.render()
    t = sc.render_tensor()
    assert t.is_cuda and t.dtype == torch.float32 and tuple(t.shape) == (225, 128, 3)
    assert np.array_equal(t.cpu().numpy().astype(np.float64), want)
    strips = torch.cat([sc.render_tensor(n, 4) for n in range(4)], dim=0)
    assert torch.equal(strips, t)
    assert np.array_equal(ops.finalize_u8(t).cpu().numpy(), to_png_array(want))
    with torch.cuda.stream(torch.cuda.Stream()):                      # a non-default current stream
        t2 = sc.render_tensor()
    torch.cuda.synchronize()
    assert torch.equal(t2, t)
    # compact tile buffers of 3 emulated ranks -> render_into + assemble_tiles
    h = torch_ext.scene_handle(sc)
    bufs, offs, total = [], [], 0
    for rank in range(3):
        job = sc.make_job(0, 225, 0, 128, rank=rank, world=3, direct=False)
        nt, nf = C.c_int64(), C.c_int64()
        abi.check(sc._lib, sc._lib.rt_job_extent(C.byref(job), C.byref(nt), C.byref(nf)), "rt_job_extent")
        buf = torch.zeros(nf.value, dtype=torch.float32, device="cuda")
        ops.render_into(buf, h, torch_ext.job_tensor(job))
        bufs.append(buf); offs.append(total); total += nf.value
    job = sc.make_job(0, 225, 0, 128, rank=0, world=3, direct=False)
    img = ops.assemble_tiles(torch.cat(bufs), torch.tensor(offs, dtype=torch.int64), torch_ext.job_tensor(job))
    assert torch.equal(img, t)
    with pytest.raises(RuntimeError):
        ops.render_into(torch.zeros(10, device="cuda"), h, torch_ext.job_tensor(job))
