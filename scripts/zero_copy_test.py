"""Experiment: k_resolve writing the frame straight into pinned HOST memory (UVA pointer as d_out) instead of device buffer + D2H."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import torch
import python_raytracer_b200 as rt
from python_raytracer_b200 import synthetic
sc = synthetic.make_scene(ingest="device")
W, H = sc.vc.width, sc.vc.height
job = sc.make_job(0, W, 0, H)
dev = torch.empty(W * H * 3, dtype=torch.float32, device="cuda")
host = torch.empty(W * H * 3, dtype=torch.float32).pin_memory()
host2 = torch.empty(W * H * 3, dtype=torch.float32).pin_memory()
def timed(fn, n=8):
    for _ in range(2): fn()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n): fn()
    torch.cuda.synchronize(); return 1e3 * (time.perf_counter() - t0) / n
def a():
    sc.render_device(dev.data_ptr(), job); host.copy_(dev, non_blocking=True); torch.cuda.current_stream().synchronize()
def b():
    sc.render_device(host2.data_ptr(), job); torch.cuda.current_stream().synchronize()
def c():
    sc.render_device(dev.data_ptr(), job); torch.cuda.current_stream().synchronize()
print("device only", round(timed(c), 3), "device + D2H", round(timed(a), 3), "resolve into pinned host", round(timed(b), 3))
print("identical:", bool(torch.equal(host, host2)))
