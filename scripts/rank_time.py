"""Device time of ONE rank's share of the bench frame (tiles rank, rank + world, ... rendered into the compact tile buffer), on one
GPU: what a rank of an N-GPU run does between two barriers, without the fan-in.  usage: rank_time.py [--world 8] [--batch-rays N ...]"""
import argparse, ctypes as C, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import torch
import python_raytracer_b200 as rt
from python_raytracer_b200 import synthetic
ap = argparse.ArgumentParser()
ap.add_argument("--world", type=int, default=8)
ap.add_argument("--batch-rays", type=int, nargs="*", default=[0])
ap.add_argument("--frames", type=int, default=12)
a = ap.parse_args()
abi = rt._abi
sc = synthetic.make_scene(ingest="device")
W, H = sc.vc.width, sc.vc.height
h = sc.upload()
for br in a.batch_rays:
    job = sc.make_job(0, W, 0, H, rank=0, world=a.world, direct=False, max_batch_rays=br)
    nt, nf = C.c_int64(), C.c_int64()
    abi.check(sc._lib, sc._lib.rt_job_extent(C.byref(job), C.byref(nt), C.byref(nf)), "extent")
    out = torch.empty(nf.value, dtype=torch.float32, device="cuda")
    for _ in range(3):
        sc.render_device(out.data_ptr(), job)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.frames):
        sc.render_device(out.data_ptr(), job)
    e1.record(); torch.cuda.synchronize()
    print(f"world {a.world} batch_rays {br}: {e0.elapsed_time(e1) / a.frames:.3f} ms/frame (ideal {115.4 / a.world:.2f})", flush=True)
