"""ms/frame and Mrays/s of every bundled-scene case (tests/golden/MANIFEST.json) + the two 1080p x 64 configs on the GPU,
next to the CPU seconds the UNMODIFIED reference took for the same case in the dev container (recorded in the manifest by
oracle/make_golden.py).  Prints one JSON line per case; run under gpurun."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "oracle")]
import torch
from conftest import ALL_CASES, load_case, manifest
RAYS = ("rays_primary", "rays_shadow", "rays_reflect", "rays_refract")
cases = [(c, None) for c in ALL_CASES] + [
    ("DepthOfField", {"resolution": [1920, 1080], "DOF": {"samples": 64}, "AA": {"jitter": True, "samples": 1}}),
    ("MotionBlur", {"resolution": [1920, 1080], "motion": {"time": 2.0, "samples": 63, "final": 1}}),
    # the two CSG showcase scenes at the values their files ship with (scenes/NovelScene1.json:2-4, NovelScene2.json:2-6)
    ("NovelScene1_small", {"resolution": [2048, 1024], "AA": {"jitter": True, "samples": 32}}),
    ("NovelScene2_small", {"resolution": [1024, 512], "AA": {"jitter": True, "samples": 2}, "DOF": {"samples": 15},
                           "motion": {"time": 1.0, "final": 1, "samples": 15}})]
if len(sys.argv) > 1:     # e.g. `Novel`: only the big versions of the named cases
    cases = [c for c in cases if any(a in c[0] for a in sys.argv[1:]) and c[1]]
man = manifest()
for case, over in cases:
    sc = load_case(case, over)
    sc.render_u8()                                   # upload + LBVH + first frame
    t = []
    for _ in range(5):
        torch.cuda.synchronize(); t0 = time.perf_counter(); sc.render_u8(); torch.cuda.synchronize(); t.append(time.perf_counter() - t0)
    ms = 1e3 * sorted(t)[len(t) // 2]
    c = sc.last_counters
    rays = sum(c[k] for k in RAYS)
    out = {"case": case + (f"@{sc.vc.width}x{sc.vc.height}x{sc.samples * sc.vc.dof_samples * len(sc.vc.motion_times)}" if over else ""), "W": sc.vc.width, "H": sc.vc.height, "rays": rays,
           "gpu_ms_e2e_u8": round(ms, 3), "gpu_mrays_s": round(rays / ms / 1e3, 1), "launches": c["kernel_launches"]}
    if not over:
        out["reference_python_cpu_seconds_dev_container"] = man[case].get("reference_cpu_seconds")
        if out["reference_python_cpu_seconds_dev_container"]:
            out["reference_mrays_s_per_core"] = round(rays / out["reference_python_cpu_seconds_dev_container"] / 1e6, 5)
    print(json.dumps(out), flush=True)
