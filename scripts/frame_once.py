"""One frame of a bench workload between cudaProfilerStart/Stop (for `ncu --profile-from-start off`), after warm-up frames.
usage: python scripts/frame_once.py [--width W --height H --spp S] [--scene NovelScene1|NovelScene2] [--frames N]"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import torch  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--width", type=int, default=3840)
ap.add_argument("--height", type=int, default=2160)
ap.add_argument("--spp", type=int, default=64)
ap.add_argument("--scene", default="synthetic")
ap.add_argument("--frames", type=int, default=1)
ap.add_argument("--warm", type=int, default=2)
a = ap.parse_args()
import python_raytracer_b200 as rt  # noqa: E402
from python_raytracer_b200 import synthetic  # noqa: E402
rt.scene_parser.VERBOSE = False
if a.scene == "synthetic":
    sc = synthetic.make_scene(width=a.width, height=a.height, aa_samples=a.spp, ingest="device")
else:
    from conftest import load_bundled
    sc = load_bundled(a.scene)
W, H = sc.vc.width, sc.vc.height
out = torch.empty(W * H * 3, dtype=torch.float32, device="cuda")
job = sc.make_job(0, W, 0, H)
for _ in range(a.warm):
    sc.render_device(out.data_ptr(), job)
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStart()
for _ in range(a.frames):
    sc.render_device(out.data_ptr(), job)
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStop()
print("ok", W, H, float(out.sum()))
