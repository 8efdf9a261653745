"""Multi-GPU split of one frame: interleaved image tiles per rank, one gather to rank 0.

B200 analogue of the reference's process fan-out (`render.nu:10-15`: N x `main.py --subimage n
--tasks N`, then `glue.py` concatenates the strips).  One process per GPU (torch.distributed, NCCL
over NVLink/NVSwitch); each rank renders the tiles `tile % world == rank` of the frame into a compact
device buffer (`rt_render`, no data-path collective during rendering), rank 0 receives the buffers
(`dist.gather`) and scatters them into the reference layout (`rt_assemble_tiles`)."""
import ctypes as C

import torch
import torch.distributed as dist

from . import _abi


class TileSharder:
    def __init__(self, scene, rank, world, device=None, group=None):
        self.scene, self.rank, self.world, self.group = scene, rank, world, group
        if device is None:
            device = torch.device("cuda", torch.cuda.current_device())
        self.device = torch.device(device)
        self.lib = _abi.load_library()
        W, H = scene.vc.width, scene.vc.height
        self.W, self.H = W, H
        self.extents = []
        for r in range(world):
            job = scene.make_job(0, W, 0, H, rank=r, world=world, direct=False)
            nt, nf = C.c_int64(), C.c_int64()
            _abi.check(self.lib, self.lib.rt_job_extent(C.byref(job), C.byref(nt), C.byref(nf)), "rt_job_extent")
            self.extents.append(nf.value)
        self.pad = max(self.extents)                       # dist.gather wants equal sizes: pad to the largest rank buffer
        self.job = scene.make_job(0, W, 0, H, rank=rank, world=world, direct=(world == 1))
        n_local = W * H * 3 if world == 1 else self.pad
        self.local = torch.zeros(n_local, dtype=torch.float32, device=self.device)
        self.image = None
        self.gathered = None
        self.offsets = (C.c_int64 * world)(*[r * self.pad for r in range(world)])
        if rank == 0 and world > 1:
            self.gathered = torch.empty(world * self.pad, dtype=torch.float32, device=self.device)
            if self.device.type == "cuda":
                self.image = torch.empty(W * H * 3, dtype=torch.float32, device=self.device)
        elif world == 1:
            self.image = self.local

    def gather(self):
        """Fan-in of the compact tile buffers to rank 0 (glue.py:19: the N.npy files).  Device-agnostic
        (NCCL for CUDA tensors, gloo for CPU tensors in the tests)."""
        if self.world == 1:
            return self.local
        if self.rank == 0:
            chunks = list(self.gathered.view(self.world, self.pad).unbind(0))
            dist.gather(self.local, chunks, dst=0, group=self.group)
            return self.gathered
        dist.gather(self.local, None, dst=0, group=self.group)
        return None

    def assemble(self, stream=None):
        """rank 0: scatter the gathered tiles into the reference layout (glue.py:22)."""
        asm = self.scene.make_job(0, self.W, 0, self.H, rank=0, world=self.world, direct=False)
        _abi.check(self.lib, self.lib.rt_assemble_tiles(C.byref(asm), self.gathered.data_ptr(), self.offsets,
                                                         self.image.data_ptr(), stream), "rt_assemble_tiles")
        return self.image

    def render(self, counters=None):
        """One frame.  Returns the (W*H*3,) device image on rank 0 (reference layout), None elsewhere."""
        stream = torch.cuda.current_stream().cuda_stream
        self.scene.render_device(self.local.data_ptr(), self.job, counters=counters, stream=stream)
        if self.world == 1:
            return self.image
        self.gather()
        return self.assemble(stream) if self.rank == 0 else None
