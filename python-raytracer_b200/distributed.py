"""Multi-GPU split of one frame: interleaved image tiles per rank, fan-in on rank 0.

B200 analogue of the reference's process fan-out (`render.nu:10-15`: N x `main.py --subimage n --tasks N`, then `glue.py`
concatenates the strips).  One process per GPU (torch.distributed, NCCL over NVLink/NVSwitch); each rank renders the tiles
`tile % world == rank` of the frame (`rt_render`, no data-path collective during rendering).  Two fan-ins:

* ``"shared"`` (default on CUDA): rank 0 owns the frame buffer (`rt_shared_alloc`), the other ranks map it through NVLink peer
  access (CUDA IPC handle broadcast once) and the LAST KERNEL of their render (`k_resolve`) stores every finished pixel straight
  into rank 0's HBM in the reference layout (`RtRenderJob.direct = RT_DIRECT_SHARED_FRAME`).  No gather, no scatter kernel; one
  4-byte NCCL all-reduce per frame orders rank 0's readers after everybody's writes.
* ``"gather"``: every rank renders a compact tile buffer, `dist.gather` to rank 0 (NCCL; gloo for the CPU tests of the host
  logic), `rt_assemble_tiles` scatters them into the reference layout (glue.py:19-22).
"""
import ctypes as C

import torch
import torch.distributed as dist

from . import _abi


class TileSharder:
    def __init__(self, scene, rank, world, device=None, group=None, fan_in=None):
        self.scene, self.rank, self.world, self.group = scene, rank, world, group
        if device is None:
            device = torch.device("cuda", torch.cuda.current_device())
        self.device = torch.device(device)
        self.lib = _abi.load_library()
        W, H = scene.vc.width, scene.vc.height
        self.W, self.H = W, H
        if fan_in is None:
            fan_in = "shared" if (self.device.type == "cuda" and world > 1) else "gather"
        self.fan_in = fan_in if world > 1 else "single"
        self.image = None
        self.gathered = None
        self._shared_ptr = None          # rank 0: the allocation; others: the peer mapping
        self.extra_launches_per_frame = 0
        self.ops = None
        if self.device.type == "cuda":   # the render / scatter calls go through torch.ops.b200rt.* (current stream, tensors)
            from . import torch_ext
            self.ops = torch_ext.load()
            self._torch_ext = torch_ext
        if self.fan_in == "shared":
            self._init_shared()
            return
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
        self.offsets = (C.c_int64 * world)(*[r * self.pad for r in range(world)])
        if rank == 0 and world > 1:
            self.gathered = torch.empty(world * self.pad, dtype=torch.float32, device=self.device)
            if self.device.type == "cuda":
                self.image = torch.empty(W * H * 3, dtype=torch.float32, device=self.device)
            self.extra_launches_per_frame = 1              # k_assemble
        elif world == 1:
            self.image = self.local

    # ------------------------------------------------------------------ shared frame (NVLink peer stores)
    def _init_shared(self):
        W, H = self.W, self.H
        # TWO frames, used alternately: while rank 0 still reads frame k (device-to-host copy, conversion) the other ranks may
        # already store pixels of frame k+1.  Frame k+2 re-uses k's buffer: a rank gets there only through the all-reduce of
        # frame k+1, which rank 0 enters after - in stream order - everything it queued behind frame k's render.  So whatever
        # the caller does with the returned image must be queued on the current stream before the next render() call.
        self._frame_bytes = (W * H * 3 * 4 + 255) // 256 * 256
        self._frame_no = 0
        nbytes = 2 * self._frame_bytes
        handle = _abi.RtIpcHandle()
        ptr = C.c_void_p()
        if self.rank == 0:
            _abi.check(self.lib, self.lib.rt_shared_alloc(nbytes, C.byref(ptr), C.byref(handle)), "rt_shared_alloc")
        blob = torch.tensor(list(bytes(handle.bytes)), dtype=torch.uint8, device=self.device)
        dist.broadcast(blob, src=0, group=self.group)
        if self.rank != 0:
            C.memmove(handle.bytes, bytes(blob.cpu().tolist()), 64)
            _abi.check(self.lib, self.lib.rt_shared_open(C.byref(handle), C.byref(ptr)), "rt_shared_open")
        self._shared_ptr = ptr
        self.job = self.scene.make_job(0, W, 0, H, rank=self.rank, world=self.world, direct=_abi.DIRECT_SHARED_FRAME)
        self._flag = torch.zeros(1, dtype=torch.int32, device=self.device)
        # tensor views of the two frames (no copy): rank 0's own allocation / the peer mapping of it on the other ranks
        self._images = [_device_tensor_from_ptr(ptr.value + k * self._frame_bytes, W * H * 3, self.device) for k in (0, 1)]
        self._job_t = self._torch_ext.job_tensor(self.job)
        if self.rank == 0:
            self.image = self._images[0]

    # ------------------------------------------------------------------ frame to HOST memory over every GPU's own PCIe link
    def _init_host_frame(self):
        """One frame in POSIX shared memory, page-locked in every rank's process (cudaHostRegister): rank 0 owns the result, every
        rank can copy into it."""
        from multiprocessing import shared_memory
        nbytes = self.W * self.H * 3 * 4
        name = [None]
        if self.rank == 0:
            self._shm = shared_memory.SharedMemory(create=True, size=nbytes)
            name[0] = self._shm.name
        dist.broadcast_object_list(name, src=0, group=self.group)
        if self.rank != 0:
            self._shm = shared_memory.SharedMemory(name=name[0])
            try:                                             # rank 0 unlinks the segment: the attachers' resource trackers must not try again
                from multiprocessing import resource_tracker
                resource_tracker.unregister(self._shm._name, "shared_memory")
            except Exception:
                pass
        self.host_image = torch.frombuffer(self._shm.buf, dtype=torch.float32, count=self.W * self.H * 3)
        rc = torch.cuda.cudart().cudaHostRegister(self.host_image.data_ptr(), nbytes, 1)          # cudaHostRegisterPortable
        if int(rc) != 0:
            raise _abi.B200RTError(f"cudaHostRegister of the shared host frame failed: {rc}")
        n = self.W * self.H * 3
        per = (n // self.world + 63) // 64 * 64                      # contiguous slices of the reference layout, 256-byte aligned
        self._host_slice = (min(n, self.rank * per), min(n, (self.rank + 1) * per) if self.rank < self.world - 1 else n)

    def render_to_host(self):
        """One frame, delivered to HOST memory: like render(), then every rank copies ITS slice of the finished frame (rank 0's
        buffer, read through the NVLink peer mapping) to the shared page-locked host frame over its own PCIe link - N copies of
        1/N of the frame in parallel instead of one copy of all of it from rank 0 (99.5 MB at 4K: 1.4 ms at PCIe 5 speed).
        A second 4-byte all-reduce orders rank 0's reader after everybody's copy.  Returns the host tensor on rank 0 (valid once
        the current stream is synchronised), None elsewhere.  'shared' fan-in only."""
        assert self.fan_in == "shared"
        if not hasattr(self, "host_image"):
            self._init_host_frame()
        self.render()
        k = (self._frame_no - 1) & 1
        f0, f1 = self._host_slice
        if f1 > f0:
            self.host_image[f0:f1].copy_(self._images[k][f0:f1], non_blocking=True)
        dist.all_reduce(self._flag, group=self.group)
        return self.host_image if self.rank == 0 else None

    def close(self):
        if getattr(self, "_shm", None) is not None:
            torch.cuda.synchronize()
            if self.world > 1:
                dist.barrier(group=self.group)
            torch.cuda.cudart().cudaHostUnregister(self.host_image.data_ptr())
            self.host_image = None
            self._shm.close()
            if self.rank == 0:
                self._shm.unlink()
            self._shm = None
        if self._shared_ptr is not None and self._shared_ptr.value:
            torch.cuda.synchronize()
            if self.world > 1:
                dist.barrier(group=self.group)             # nobody unmaps / frees while another rank may still store
            self.image = self._images = None
            if self.rank == 0:
                self.lib.rt_shared_free(self._shared_ptr)
            else:
                self.lib.rt_shared_close(self._shared_ptr)
            self._shared_ptr = None

    # ------------------------------------------------------------------ gather (NCCL / gloo)
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
        if self.fan_in == "shared":
            k = self._frame_no & 1
            self._frame_no += 1
            if counters is None:
                self.ops.render_into(self._images[k], self._torch_ext.scene_handle(self.scene), self._job_t)
            else:       # the tallies come back through the C ABI call (it synchronises the stream)
                self.scene.render_device(C.c_void_p(self._shared_ptr.value + k * self._frame_bytes), self.job, counters=counters, stream=stream)
            # stream-ordered on every rank after its own stores; rank 0's later work (copies, conversions) waits for all of them
            dist.all_reduce(self._flag, group=self.group)
            if self.rank == 0:
                self.image = self._images[k]
            return self.image if self.rank == 0 else None
        if counters is None and self.ops is not None:
            if not hasattr(self, "_job_t"):
                self._job_t = self._torch_ext.job_tensor(self.job)
            self.ops.render_into(self.local, self._torch_ext.scene_handle(self.scene), self._job_t)
        else:           # the tallies come back through the C ABI call (it synchronises the stream)
            self.scene.render_device(self.local.data_ptr(), self.job, counters=counters, stream=stream)
        if self.world == 1:
            return self.image
        self.gather()
        return self.assemble(stream) if self.rank == 0 else None


def _device_tensor_from_ptr(ptr, n_floats, device):
    """A float32 torch tensor over device memory owned by the library (no copy), through __cuda_array_interface__."""
    class _Mem:
        __cuda_array_interface__ = {"shape": (n_floats,), "typestr": "<f4", "data": (int(ptr), False), "version": 3, "strides": None}
    return torch.as_tensor(_Mem(), device=device)
