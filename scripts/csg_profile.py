"""Per-kernel-class device time and tallies of the bundled CSG scenes at their SHIPPED settings (rt_scene_profile event pairs).
usage: python scripts/csg_profile.py [NovelScene1 NovelScene2 ...]   (under gpurun); one JSON line per scene."""
import ctypes as C, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import torch
from conftest import load_bundled
import python_raytracer_b200 as rt
abi = rt._abi
for name in (sys.argv[1:] or ["NovelScene1", "NovelScene2"]):
    sc = load_bundled(name)
    W, H = sc.vc.width, sc.vc.height
    out = torch.empty(W * H * 3, dtype=torch.float32, device="cuda")
    job = sc.make_job(0, W, 0, H)
    ctr = abi.RtCounters()
    for _ in range(2):
        sc.render_device(out.data_ptr(), job, ctr)
    torch.cuda.synchronize()
    t = []
    for _ in range(3):
        t0 = time.perf_counter(); sc.render_device(out.data_ptr(), job); torch.cuda.synchronize(); t.append(1e3 * (time.perf_counter() - t0))
    lib, h = sc._lib, sc.upload()
    abi.check(lib, lib.rt_scene_profile(h, 1), "profile")
    n = 3
    for _ in range(n):
        sc.render_device(out.data_ptr(), job)
    prof = abi.RtProfile()
    abi.check(lib, lib.rt_scene_profile_read(h, C.byref(prof)), "read")
    abi.check(lib, lib.rt_scene_profile(h, 0), "profile")
    ms = {k: round(getattr(prof, k) / n, 3) for k in ("ms_hit", "ms_shadow", "ms_shade", "ms_resolve", "ms_l1_hit", "ms_l1_shadow", "ms_l1_shade", "ms_deep", "ms_tail")}
    tal = {cls: {k: getattr(getattr(prof, cls), k) for k in ("rays_primary", "rays_shadow", "rays_reflect", "rays_refract", "bvh_nodes", "tri_tests", "prim_tests", "shaded_hits", "queued_rays")}
           for cls in ("hit", "shadow", "shade", "secondary")}
    print(json.dumps({"scene": name, "W": W, "H": H, "ms_frame_wall_median": round(sorted(t)[1], 3), "ms_per_class": ms,
                      "rays": {k: getattr(ctr, k) for k in ("rays_primary", "rays_shadow", "rays_reflect", "rays_refract")}, "tallies": tal}), flush=True)
