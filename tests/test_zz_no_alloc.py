"""rt_render_static (include/b200rt.h): the render with its wavefront scratch in the library's static device array, for callers
that cannot call cudaMalloc.  Called through the C ABI (ctypes) and compared with rt_render, the existing entry point, on two
calls in a row (the array keeps the previous call's contents).  Tolerance: the two entry points run the same kernels on the same
rays in smaller batches, and the frame does not depend on the batch size (test_batch_size_does_not_change_a_bit_...), so the
images must be EQUAL bit for bit and so must the ray tallies.

Collected last on purpose (file name)."""
import ctypes as C

import numpy as np
import pytest

import python_raytracer_b200 as rt
from conftest import load_case

pytestmark = pytest.mark.gpu
abi = rt._abi

RT_STATIC_MAX_RAYS = 65536       # include/b200rt.h
RT_STATIC_MAX_LIGHTS = 8
RT_ERR_LIMIT = 5


def _render_both(sc, ncols, nrows):
    import torch
    h = sc.upload()
    job = sc.make_job(0, ncols, 0, nrows)
    ref = torch.zeros((ncols, nrows, 3), dtype=torch.float32, device="cuda")
    c_ref = abi.RtCounters()
    abi.check(sc._lib, sc._lib.rt_render(h, C.byref(job), ref.data_ptr(), C.byref(c_ref), None), "rt_render")
    outs = []
    for call in range(2):                                       # two calls in a row share the static array
        out = torch.full((ncols, nrows, 3), -1.0, dtype=torch.float32, device="cuda")
        torch.cuda.synchronize()
        before = torch.cuda.mem_get_info()[0]
        ctr = abi.RtCounters()
        abi.check(sc._lib, sc._lib.rt_render_static(h, C.byref(job), out.data_ptr(), C.byref(ctr), None), "rt_render_static")
        torch.cuda.synchronize()
        assert torch.cuda.mem_get_info()[0] >= before, "rt_render_static took device memory"
        outs.append((out.cpu().numpy(), ctr))
    return ref.cpu().numpy(), c_ref, outs


def test_static_scratch_render_equals_rt_render_recursive_materials():
    """Mirror / refraction chains (queued levels + k_tail) on 900 x 512: 460 800 rays = 8 static batches."""
    sc = load_case("MirrorRefraction")
    ref, c_ref, outs = _render_both(sc, 900, 512)
    assert c_ref.rays_reflect > 0 and c_ref.rays_refract > 0 and c_ref.error_flags == 0
    for out, ctr in outs:
        assert np.array_equal(out, ref)
        for k in ("rays_primary", "rays_shadow", "rays_reflect", "rays_refract", "shaded_hits", "error_flags"):
            assert getattr(ctr, k) == getattr(c_ref, k), k
        assert ctr.kernel_launches > c_ref.kernel_launches       # more, smaller batches


def test_static_scratch_render_equals_rt_render_mesh_csg_many_samples():
    """A mesh scene through the LBVH (shared-memory stack) and a CSG scene with 24 samples per pixel (DOF x AA x motion)."""
    for case in ("TorusMesh_smooth", "NovelScene2_small"):
        sc = load_case(case)
        ref, c_ref, outs = _render_both(sc, sc.vc.width, sc.vc.height)
        for out, ctr in outs:
            assert np.array_equal(out, ref), case
            assert ctr.rays_primary == c_ref.rays_primary and ctr.rays_shadow == c_ref.rays_shadow, case


def test_static_scratch_rejects_what_does_not_fit():
    import torch
    sc = load_case("DepthOfField", {"resolution": [8, 8], "AA": {"jitter": True, "samples": 300}, "DOF": {"samples": 300, "focal_length": 5.0, "aperture": 0.1}})
    spp = sc.vc.dof_samples * sc.samples * len(sc.vc.motion_times)
    assert spp > RT_STATIC_MAX_RAYS
    h = sc.upload()
    job = sc.make_job(0, 8, 0, 8)
    out = torch.zeros((8, 8, 3), dtype=torch.float32, device="cuda")
    rc = sc._lib.rt_render_static(h, C.byref(job), out.data_ptr(), None, None)
    assert rc == RT_ERR_LIMIT
    assert b"samples per pixel" in sc._lib.rt_last_error()
