"""N>1 host logic on CPU: world_size-2 gloo processes run the tile sharder's extents / padding / gather,
and the gathered buffers are scattered with the library's own pixel mapping (tests/hostsim exports the
RT_HD `pixel_of` used by k_assemble)."""
import os
import socket
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import ROOT, load_case


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, W, H, result_path):
    for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests", "hostsim"), os.path.join(ROOT, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import hostsim
    from python_raytracer_b200.distributed import TileSharder
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sc = load_case("TwoSpheresPlane", {"resolution": [W, H]})
    sh = TileSharder(sc, rank, world, device="cpu")
    # what the render kernels would leave in this rank's compact buffer: here the pixel's own coordinates
    cols, rows, valid = hostsim.job_pixels(sh.job)
    buf = np.zeros((sh.pad // 3, 3), np.float32)
    buf[: len(cols)][valid] = np.stack([cols[valid], rows[valid], np.full(valid.sum(), rank)], axis=1)
    sh.local.copy_(torch.from_numpy(buf.reshape(-1)))
    g = sh.gather()
    if rank == 0:
        g = g.numpy().reshape(world, sh.pad // 3, 3)
        out = np.full((W, H, 3), -1.0, np.float32)
        for r in range(world):
            job = sc.make_job(0, W, 0, H, rank=r, world=world, direct=False)
            c, rr, v = hostsim.job_pixels(job)
            assert len(c) * 3 == sh.extents[r] <= sh.pad
            out[c[v], rr[v]] = g[r, : len(c)][v]
        np.save(result_path, out)
    dist.barrier()
    dist.destroy_process_group()


def test_gloo_world2_gather_reassembles_the_frame(tmp_path):
    W, H, world = 37, 21, 2
    path = str(tmp_path / "out.npy")
    mp.spawn(_worker, args=(world, _free_port(), W, H, path), nprocs=world, join=True)
    out = np.load(path)
    cc, rr = np.meshgrid(np.arange(W), np.arange(H), indexing="ij")
    assert np.array_equal(out[..., 0], cc) and np.array_equal(out[..., 1], rr)
    tile = (cc // 8) * ((H + 7) // 8) + rr // 8          # column-major tile index, tile % world == rank
    assert np.array_equal(out[..., 2], tile % world)
