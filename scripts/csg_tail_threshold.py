"""NovelScene1 at growing sizes with k_tail from level 1 vs never (B200RT_TAIL_LEVEL is read once per process: run per value)."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "oracle")]
import torch
from conftest import load_case
for w, h in ((256, 128), (512, 256), (1024, 512), (2048, 1024)):
    sc = load_case("NovelScene1_small", {"resolution": [w, h], "AA": {"jitter": True, "samples": 32}})
    sc.render_u8()
    t = []
    for _ in range(5):
        torch.cuda.synchronize(); t0 = time.perf_counter(); sc.render_u8(); torch.cuda.synchronize(); t.append(time.perf_counter() - t0)
    print(os.environ.get("B200RT_TAIL_LEVEL"), w, h, w * h * 32, round(1e3 * sorted(t)[2], 3), "ms", flush=True)
