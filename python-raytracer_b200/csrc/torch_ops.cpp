// torch_ops.cpp - the thin PyTorch C++ extension of the boundary (north_star: "hands flattened SoA buffers to
// hand-written sm_100a CUDA kernels through a thin PyTorch C++/CUDA extension").  It registers the entry points of the C ABI
// (include/b200rt.h, implemented in libb200rt.so next to this file) as torch.ops.b200rt.*: tensors in, tensors out, the
// CURRENT torch CUDA stream, errors as exceptions.  No kernels live here; everything is forwarded to the C ABI.
//
//   torch.ops.b200rt.render(scene, job)          -> float32 (ncols, nrows, 3) on the scene's device (Scene.render, scene.py:35-79)
//   torch.ops.b200rt.render_into(out, scene, job) : same into a caller-owned buffer (compact tile buffer / shared frame included)
//   torch.ops.b200rt.finalize_u8(image)          -> uint8 (nrows, ncols, 3)  (main.py:31-33)
//   torch.ops.b200rt.assemble_tiles(gathered, rank_offsets, job) -> float32 (ncols, nrows, 3)  (glue.py:19-22)
// `scene` is the RtScene* of rt_scene_create as an int; `job` is the RtRenderJob as an int64 tensor of 11 values on the CPU:
// [col0, ncols, row0, nrows, tile_w, tile_h, rank, world, direct, mesh_mode, max_batch_rays].
#include <ATen/ATen.h>
#include <c10/cuda/CUDAGuard.h>
#include <c10/cuda/CUDAStream.h>
#include <torch/library.h>

#include "../../include/b200rt.h"

namespace {

RtRenderJob job_of(const at::Tensor& j) {
    TORCH_CHECK(j.device().is_cpu() && j.scalar_type() == at::kLong && j.numel() == 11, "job: int64 CPU tensor of 11 values");
    const int64_t* v = j.data_ptr<int64_t>();
    RtRenderJob job;
    job.col0 = (int32_t)v[0]; job.ncols = (int32_t)v[1]; job.row0 = (int32_t)v[2]; job.nrows = (int32_t)v[3];
    job.tile_w = (int32_t)v[4]; job.tile_h = (int32_t)v[5]; job.rank = (int32_t)v[6]; job.world = (int32_t)v[7];
    job.direct = (int32_t)v[8]; job.mesh_mode = (int32_t)v[9]; job.max_batch_rays = v[10];
    return job;
}

void check(int rc, const char* what) {
    TORCH_CHECK(rc == RT_OK, what, " failed with status ", rc, ": ", rt_last_error());
}

void render_into(at::Tensor out, int64_t scene, const at::Tensor& job_t) {
    TORCH_CHECK(out.is_cuda() && out.scalar_type() == at::kFloat && out.is_contiguous(), "out: contiguous float32 CUDA tensor");
    const RtRenderJob job = job_of(job_t);
    int64_t n_floats = 0;
    check(rt_job_extent(&job, nullptr, &n_floats), "rt_job_extent");
    TORCH_CHECK(out.numel() >= n_floats, "out holds ", out.numel(), " floats, the job writes ", n_floats);
    c10::cuda::CUDAGuard guard(out.device());
    check(rt_render(reinterpret_cast<RtScene*>(scene), &job, out.data_ptr<float>(), nullptr,
                    c10::cuda::getCurrentCUDAStream().stream()), "rt_render");
}

at::Tensor render(int64_t scene, const at::Tensor& job_t) {
    const RtRenderJob job = job_of(job_t);
    TORCH_CHECK(job.world == 1 && job.direct == 1, "render returns the reference layout: world 1, direct 1 (use render_into otherwise)");
    at::Tensor out = at::empty({job.ncols, job.nrows, 3}, at::TensorOptions().dtype(at::kFloat).device(at::kCUDA));
    render_into(out, scene, job_t);
    return out;
}

at::Tensor finalize_u8(const at::Tensor& image) {
    TORCH_CHECK(image.is_cuda() && image.scalar_type() == at::kFloat && image.is_contiguous() && image.dim() == 3 && image.size(2) == 3,
                "image: contiguous float32 CUDA tensor (ncols, nrows, 3)");
    const int64_t ncols = image.size(0), nrows = image.size(1);
    at::Tensor out = at::empty({nrows, ncols, 3}, image.options().dtype(at::kByte));
    c10::cuda::CUDAGuard guard(image.device());
    check(rt_finalize_u8(image.data_ptr<float>(), (int32_t)ncols, (int32_t)nrows, out.data_ptr<uint8_t>(),
                         c10::cuda::getCurrentCUDAStream().stream()), "rt_finalize_u8");
    return out;
}

at::Tensor assemble_tiles(const at::Tensor& gathered, const at::Tensor& rank_offsets, const at::Tensor& job_t) {
    TORCH_CHECK(gathered.is_cuda() && gathered.scalar_type() == at::kFloat && gathered.is_contiguous(), "gathered: contiguous float32 CUDA tensor");
    TORCH_CHECK(rank_offsets.device().is_cpu() && rank_offsets.scalar_type() == at::kLong, "rank_offsets: int64 CPU tensor");
    const RtRenderJob job = job_of(job_t);
    TORCH_CHECK(rank_offsets.numel() == job.world, "one offset per rank");
    at::Tensor out = at::empty({job.ncols, job.nrows, 3}, gathered.options());
    c10::cuda::CUDAGuard guard(gathered.device());
    check(rt_assemble_tiles(&job, gathered.data_ptr<float>(), rank_offsets.data_ptr<int64_t>(), out.data_ptr<float>(),
                            c10::cuda::getCurrentCUDAStream().stream()), "rt_assemble_tiles");
    return out;
}

}  // namespace

TORCH_LIBRARY(b200rt, m) {
    m.def("render(int scene, Tensor job) -> Tensor");
    m.def("render_into(Tensor(a!) out, int scene, Tensor job) -> ()");
    m.def("finalize_u8(Tensor image) -> Tensor");
    m.def("assemble_tiles(Tensor gathered, Tensor rank_offsets, Tensor job) -> Tensor");
}
TORCH_LIBRARY_IMPL(b200rt, CompositeExplicitAutograd, m) {
    m.impl("render", render);
    m.impl("render_into", render_into);
    m.impl("finalize_u8", finalize_u8);
    m.impl("assemble_tiles", assemble_tiles);
}
