"""torchrun --nproc-per-node N tests/run_multigpu_check.py : N-rank tile-sharded render gathered over NCCL
must equal the single-rank render bit for bit.  Prints one JSON line on rank 0."""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import load_case  # noqa: E402
from python_raytracer_b200.distributed import TileSharder  # noqa: E402
from python_raytracer_b200 import synthetic  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
res = {}
scenes = {"MirrorRefraction": load_case("MirrorRefraction"),
          "synthetic_640x360x4": synthetic.make_scene(width=640, height=360, aa_samples=4, grid=(256, 128))}
for name, sc in scenes.items():
    for fan_in in ("shared", "gather"):      # NVLink peer stores from k_resolve into rank 0's frame / NCCL gather + scatter kernel
        sh = TileSharder(sc, rank, world, fan_in=fan_in)
        for frame in range(3):               # the shared frame is double-buffered: three frames touch both buffers
            img = sh.render()
            torch.cuda.synchronize()
            if rank == 0:
                got = img.cpu().numpy().reshape(sc.vc.width, sc.vc.height, 3)
                ref = sc.render_region(0, sc.vc.width, 0, sc.vc.height)
                ok = bool(np.array_equal(got, ref))
                prev = res.get(f"{name}/{fan_in}", {"bit_identical": True, "max_abs_diff": 0.0})
                res[f"{name}/{fan_in}"] = {"bit_identical": ok and prev["bit_identical"],
                                           "max_abs_diff": max(prev["max_abs_diff"], float(np.abs(got - ref).max()))}
            dist.barrier()
        if fan_in == "shared":               # the same frames delivered to the shared page-locked HOST frame, one slice per rank over its own PCIe link
            for frame in range(3):
                host = sh.render_to_host()
                torch.cuda.synchronize()
                if rank == 0:
                    got = host.numpy().reshape(sc.vc.width, sc.vc.height, 3)
                    ref = sc.render_region(0, sc.vc.width, 0, sc.vc.height)
                    prev = res.get(f"{name}/host", {"bit_identical": True, "max_abs_diff": 0.0})
                    res[f"{name}/host"] = {"bit_identical": bool(np.array_equal(got, ref)) and prev["bit_identical"],
                                           "max_abs_diff": max(prev["max_abs_diff"], float(np.abs(got - ref).max()))}
                dist.barrier()
        sh.close()
if rank == 0:
    print(json.dumps({"world": world, "results": res}))
    assert all(v["bit_identical"] for v in res.values())
dist.destroy_process_group()
