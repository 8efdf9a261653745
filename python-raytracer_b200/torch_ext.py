"""torch.ops.b200rt.* : the thin PyTorch C++ extension over the C ABI (csrc/torch_ops.cpp -> csrc/libb200rt_torch.so).

Tensors in, tensors out, the current torch CUDA stream; everything is forwarded to `libb200rt.so`.  There is no fallback:
`load()` raises when the extension has not been built (`python __graft_entry__.py`)."""
import os

import torch

from . import _abi

LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "csrc", "libb200rt_torch.so")
_loaded = False


def load():
    global _loaded
    if not _loaded:
        if not os.path.exists(LIB_PATH):
            raise _abi.B200RTError(f"{LIB_PATH} not found: build it with `python __graft_entry__.py`; there is no fallback")
        _abi.load_library()                       # libb200rt.so first (same handle the ctypes side uses)
        torch.ops.load_library(LIB_PATH)
        _loaded = True
    return torch.ops.b200rt


def job_tensor(job):
    """RtRenderJob (ctypes) -> the int64 CPU tensor the ops take."""
    return torch.tensor([job.col0, job.ncols, job.row0, job.nrows, job.tile_w, job.tile_h, job.rank, job.world, job.direct,
                         job.mesh_mode, job.max_batch_rays], dtype=torch.int64)


def scene_handle(scene):
    """RtScene* of an uploaded `Scene` as the int the ops take."""
    return int(scene.upload().value)
