#!/usr/bin/env python
"""bench.py - one JSON line per run (see DESIGN.md "Measurement").

  python bench.py [--gpus N --steps K --warmup W]            our arm: B200 kernels through the C ABI
  python bench.py --impl reference [...]                     the reference's CPU path (oracle port), host cores

Workload (config.workload): the synthetic 1M-triangle mesh + 64 spheres scene at 3840x2160x64 spp
(BASELINE.json configs[4], the configuration the metric is quoted on).  A "step" is one frame.
N > 1: image tiles interleaved over ranks, NCCL gather to rank 0 (strong scaling: the frame is fixed).
"""
import argparse
import ctypes as C
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

import numpy as np  # noqa: E402

METRIC = "Mrays/s (primary+secondary)"
RAY_KEYS = ("rays_primary", "rays_shadow", "rays_reflect", "rays_refract")


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--width", type=int, default=3840)
    ap.add_argument("--height", type=int, default=2160)
    ap.add_argument("--spp", type=int, default=64)
    ap.add_argument("--grid", type=int, nargs=2, default=[1024, 512], help="mesh quad grid (2*a*b triangles)")
    ap.add_argument("--spheres", type=int, default=64)
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="target CPU time of the cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--batch-rays", type=int, default=0, help="rays per wavefront batch (0 = library default)")
    ap.add_argument("--mode-bits", type=int, default=0, help="RtRenderJob.mesh_mode test-aid bits OR-ed into every job (tuning experiments)")
    return ap.parse_args()


def workload_config(a):
    tris = 2 * a.grid[0] * a.grid[1]
    return {"workload": f"synthetic {tris}-triangle mesh + {a.spheres} spheres, {a.width}x{a.height}x{a.spp} spp, jitter on",
            "width": a.width, "height": a.height, "spp": a.spp, "triangles": tris, "spheres": a.spheres,
            "primary_rays_per_frame": a.width * a.height * a.spp}


def make_scene(a, ingest="host"):
    from python_raytracer_b200 import synthetic
    return synthetic.make_scene(width=a.width, height=a.height, aa_samples=a.spp, grid=tuple(a.grid), n_spheres=a.spheres,
                                ingest=ingest)


# ----------------------------------------------------------------------------- CPU arm (oracle port)
def cpu_sample(sc, target_seconds, threads=0):
    """Time the oracle (brute-force restatement of the reference) on a bounded, evenly spread pixel sample of
    the SAME frame (all spp of each sampled pixel); returns dict(value Mrays/s, cores, sample, seconds)."""
    import oracle
    import python_raytracer_b200 as rt
    fs = sc.flatten()
    W, H = sc.vc.width, sc.vc.height
    rng = np.random.default_rng(7)
    cores = oracle.lib().oracle_num_threads() if threads == 0 else threads

    def run(n):
        cols = rng.integers(0, W, size=n).astype(np.int32)
        rows = rng.integers(0, H, size=n).astype(np.int32)
        t0 = time.perf_counter()
        _, ctr = oracle.render_pixels(fs, rt._abi, cols, rows, threads)
        return time.perf_counter() - t0, ctr

    n = max(cores, 8)
    dt, ctr = run(n)
    total_dt, total_rays, total_px = dt, sum(ctr[k] for k in RAY_KEYS), n
    while total_dt < target_seconds and total_px < W * H:
        rate = total_px / max(total_dt, 1e-6)
        n = int(max(cores, min(rate * (target_seconds - total_dt) * 0.9, W * H - total_px)))
        dt, ctr = run(n)
        total_dt += dt; total_rays += sum(ctr[k] for k in RAY_KEYS); total_px += n
        if dt < 0.05 * target_seconds and n == cores:
            break
    return {"value": total_rays / total_dt / 1e6, "unit": "Mrays/s", "cores": int(cores), "kind": "port",
            "sample": f"{total_px} random pixels x all {sc.samples * sc.vc.dof_samples * len(sc.vc.motion_times)} samples of the "
                      f"{W}x{H} frame, brute force over every face like mesh.py:77 (no BVH), OpenMP over pixels",
            "seconds": round(total_dt, 2), "rays": int(total_rays), "pixels": int(total_px)}


def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import python_raytracer_b200 as rt
    rt.scene_parser.VERBOSE = False
    sc = make_scene(a)
    per_step = max(1.0, min(20.0, 60.0 / max(1, a.steps + a.warmup)))
    for _ in range(a.warmup):
        cpu_sample(sc, per_step / 4)
    vals, last = [], None
    for _ in range(a.steps):
        last = cpu_sample(sc, per_step)
        vals.append(last)
    rays = sum(v["rays"] for v in vals); secs = sum(v["seconds"] for v in vals)
    value = rays / secs / 1e6
    frame_rays_est = None
    out = {"impl": "reference", "metric": METRIC, "value": value, "unit": "Mrays/s", "n_gpus": a.gpus, "steps": a.steps,
           "warmup": a.warmup, "ms_per_step": 1e3 * secs / a.steps, "higher_is_better": True, "scaling": "strong",
           "vs_baseline": None, "dtype": "f32 vectors / f64 scalars", "data": "synthetic",
           "config": workload_config(a),
           "cpu_baseline": {**{k: last[k] for k in ("unit", "cores", "kind", "sample")}, "value": value},
           "e2e": {"value": value, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
           "note": "the Python reference cannot travel to the GPU box; this arm times its C restatement (oracle/rt_oracle.c, "
                   "bit-exact vs the reference on every bundled scene) on all host cores; each step is a bounded pixel sample"}
    print(json.dumps(out), flush=True)


# ----------------------------------------------------------------------------- GPU arm
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(gpu_index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                       "-lms", "100"], stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.p is None:
            return out
        time.sleep(0.15)
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush(); self.f.seek(0)
        sm, mx, reasons, power = [], [], set(), []
        for line in self.f.read().splitlines():
            parts = [x.strip() for x in line.split(",")]
            if len(parts) < 9:
                continue
            try:
                sm.append(float(parts[1])); mx.append(float(parts[2])); power.append(float(parts[3]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), parts[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        os.unlink(self.f.name)
        if sm:
            out = {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm),
                   "power_w_max": max(power)}
        return out


def traffic_of(kernel, a):
    """dram__bytes_read+write of one launch of `kernel` from the committed ncu --set full capture of THIS workload
    (profiles/traffic.json, written by profiles/summarize.py), or None when no capture matches."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as fh:
            t = json.load(fh)
        return t.get(f"{kernel}@{a.width}x{a.height}x{a.spp}")
    except Exception:
        return None


def run_ours(a):
    import torch
    import torch.distributed as dist
    import python_raytracer_b200 as rt
    from python_raytracer_b200 import _abi
    from python_raytracer_b200.distributed import TileSharder
    rt.scene_parser.VERBOSE = False
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py (our arm) needs a CUDA device; there is no CPU fallback"
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    assert world == a.gpus or world == 1, (world, a.gpus)

    lib = _abi.load_library()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    sc = make_scene(a, ingest="device")          # mesh transform / vertex normals / extent on the GPU (rt_mesh_ingest)
    t_scene = time.perf_counter() - t0
    t_ingest = {}
    if rank == 0:                                # the per-mesh load-time work alone (mesh.py:17-70), both loaders, same arrays
        from python_raytracer_b200 import geometry, synthetic
        vf = synthetic.displaced_torus(a.grid[0], a.grid[1], 1234)
        for how in ("device", "host"):
            t0 = time.perf_counter()
            geometry.Mesh("m", "mesh", [], (0.0, 0.0, 0.0), 1.0, vf, False, None, ingest=how)
            t_ingest[f"mesh_ingest_{how}"] = 1e3 * (time.perf_counter() - t0)
    t0 = time.perf_counter()
    fs = sc.flatten()
    h = C.c_void_p()
    _abi.check(lib, lib.rt_scene_create(C.byref(fs.desc), C.byref(h)), "rt_scene_create")
    torch.cuda.synchronize()
    t_upload = time.perf_counter() - t0
    t0 = time.perf_counter()
    _abi.check(lib, lib.rt_scene_build_bvh(h, None), "rt_scene_build_bvh")
    torch.cuda.synchronize()
    t_bvh = time.perf_counter() - t0
    sc._lib, sc._handle = lib, h
    sc.max_batch_rays = a.batch_rays
    sc.debug_mode_bits = a.mode_bits

    sharder = TileSharder(sc, rank, world)
    W, H = a.width, a.height

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- warm-up (also yields the per-frame ray tallies: deterministic for a fixed scene)
    ctr = _abi.RtCounters()
    sharder.render(counters=ctr)
    assert ctr.error_flags == 0, f"kernels flagged errors 0x{ctr.error_flags:x} (traversal stack / material limits)"
    for _ in range(max(0, a.warmup - 1)):
        sharder.render()
    tallies = torch.tensor([getattr(ctr, k) for k in RAY_KEYS] + [ctr.kernel_launches, ctr.bvh_nodes, ctr.tri_tests, ctr.queued_rays,
                                                                  ctr.mesh_hits_smooth, ctr.shaded_hits],
                           dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tallies)
    tl = [int(x) for x in tallies.tolist()]
    rays_frame = sum(tl[:4])
    launches_frame = tl[4] + (1 if world > 1 else 0)

    # ---- timed region 1: inputs resident in HBM, image left on the device (rank 0)
    sampler = ClockSampler(local_rank) if rank == 0 else None
    _abi.check(lib, lib.rt_scene_profile(h, 1), "rt_scene_profile")
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        sharder.render()
    e1.record()
    barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_total = float(ms.item())
    prof = _abi.RtProfile()
    _abi.check(lib, lib.rt_scene_profile_read(h, C.byref(prof)), "rt_scene_profile_read")
    _abi.check(lib, lib.rt_scene_profile(h, 0), "rt_scene_profile")

    # ---- timed region 2 (e2e): the public call with HOST output; D2H of the frame inside the timed region
    host_img = torch.empty(W * H * 3, dtype=torch.float32).pin_memory() if rank == 0 else None
    e2e_steps = a.steps

    def e2e_frame():
        img = sharder.render()
        if rank == 0:
            host_img.copy_(img, non_blocking=True)
        # every rank's caller waits for its call to finish (rank 0: owns the pixels) before the next frame starts; with only
        # rank 0 waiting, the other ranks queued whole frames ahead and the two-rank e2e time wandered between 65 and 93 ms
        torch.cuda.current_stream().synchronize()

    for _ in range(min(2, a.warmup)):       # warm-up of THIS path: first copy into the fresh pinned buffer, first per-frame waits
        e2e_frame()
    barrier()
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    f0.record()
    for _ in range(e2e_steps):
        e2e_frame()
    f1.record()
    barrier()
    ms2 = torch.tensor([f0.elapsed_time(f1)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(ms2, op=dist.ReduceOp.MAX)
    ms2_total = float(ms2.item())
    clocks = sampler.stop() if sampler else None

    if rank == 0:
        import json as _json
        peaks = {}
        pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
        if os.path.exists(pk):
            with open(pk) as fh:
                peaks = _json.load(fh)
        hbm_peak, peak_src = (peaks.get("hbm_gbs"), "measured (MEASURED_PEAKS.json)") if peaks.get("hbm_gbs") else (6650.0, "fallback")
        # roofline of the dominant kernel: algorithmic bytes per launch / mean launch time (DESIGN.md "Measurement").
        # Bytes per unit: 128 B per 4-wide BVH node fetched, 48 B per triangle tested, 64 B of vertex ids + normals per
        # smooth-shaded mesh hit, 48 B hit record per ray (k_hit writes, k_shade reads), 16 B hit position per shadow
        # ray + 1 B flag, 16 B per finished sample, 80 B per queued secondary ray.
        primary_rays_rank = W * H * a.spp / world
        n_lights = len(sc.lights)
        kinds = {
            "k_hit": (prof.ms_hit, prof.n_hit, prof.hit,
                      lambda c: 128 * c.bvh_nodes + 48 * c.tri_tests + 64 * c.mesh_hits_smooth + 48 * primary_rays_rank),
            "k_shadow": (prof.ms_shadow, prof.n_shadow, prof.shadow,
                         lambda c: 128 * c.bvh_nodes + 48 * c.tri_tests + 16 * primary_rays_rank * n_lights + c.rays_shadow),
            "k_shade": (prof.ms_shade, prof.n_shade, prof.shade,
                        lambda c: (48 + n_lights) * c.shaded_hits + 16 * primary_rays_rank + 80 * c.queued_rays),
        }
        dom = max(kinds, key=lambda k: kinds[k][0])
        ms_dom_total, n_launch, pc, bytes_of = kinds[dom]
        n_launch = max(1, int(n_launch))
        launches_per_frame = max(1, n_launch // max(1, a.steps))
        alg_bytes_frame = bytes_of(pc)
        ms_dom = ms_dom_total / n_launch
        achieved = alg_bytes_frame / launches_per_frame / (ms_dom * 1e-3) / 1e9 if ms_dom > 0 else None
        ms_all = prof.ms_hit + prof.ms_shadow + prof.ms_shade + prof.ms_secondary + prof.ms_resolve
        kernel_share = ms_dom_total / max(1e-9, ms_all)
        value = rays_frame * a.steps / (ms_total * 1e-3) / 1e6
        e2e_value = rays_frame * e2e_steps / (ms2_total * 1e-3) / 1e6
        out = {"metric": METRIC, "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
               "ms_per_step": ms_total / a.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
               "dtype": "f32", "data": "synthetic",
               "config": {**workload_config(a), "tiles": f"{sc.tile[0]}x{sc.tile[1]} interleaved over ranks",
                          "l2_policy": "working set (BVH nodes+triangles+sample/queue buffers) exceeds the 126 MB L2; no flush between frames"},
               "rays_per_frame": dict(zip(RAY_KEYS, tl[:4])),
               "e2e": {"value": e2e_value, "unit": "Mrays/s", "ms_per_step": ms2_total / e2e_steps,
                       "h2d_bytes_per_step": C.sizeof(_abi.RtRenderJob), "d2h_bytes_per_step": W * H * 3 * 4,
                       "what": "frame rendered through the C ABI and copied to pinned HOST memory every step (scene resident: "
                               "uploaded once like the reference's load_scene)"},
               "gpu_launches": launches_frame * a.steps,
               "roofline": {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": hbm_peak, "unit": "GB/s",
                            "frac": (achieved / hbm_peak) if achieved else None, "traffic": traffic_of(dom, a), "peak_source": peak_src,
                            "note": "algorithmic bytes = BVH nodes/triangles the kernel fetches (tallies x record size); the 93 MB BVH "
                                    "is L1/L2 resident, so DRAM traffic is far below them and frac can exceed 1: the kernel is bound by "
                                    "L1 register write-back (85 % of peak wavefronts) and issue slots (75 %) (profiles/r01w_final_4k64.txt), not by HBM",
                            "algorithmic_bytes_per_launch": alg_bytes_frame / launches_per_frame,
                            "launch_ms": ms_dom, "launches_per_frame": launches_per_frame,
                            "kernel_share_of_step": kernel_share,
                            "counts_rank0_per_frame": {"bvh_nodes": int(pc.bvh_nodes), "tri_tests": int(pc.tri_tests),
                                                       "queued_rays": int(pc.queued_rays)}},
               "tallies_rank0_per_frame": {k: {n: v for n, v in c.as_dict().items() if v and n not in ("kernel_launches", "error_flags")}
                                           for k, c in (("k_hit", prof.hit), ("k_shadow", prof.shadow), ("k_shade", prof.shade),
                                                        ("levels>=1", prof.secondary))},
               "kernel_ms_per_frame": {"k_hit": prof.ms_hit / a.steps, "k_shadow": prof.ms_shadow / a.steps,
                                       "k_shade": prof.ms_shade / a.steps, "levels>=1 (hit+shadow+shade)": prof.ms_secondary / a.steps,
                                       "k_resolve": prof.ms_resolve / a.steps},
               "setup_ms": {"scene_python": 1e3 * t_scene, **t_ingest, "upload": 1e3 * t_upload, "lbvh_build": 1e3 * t_bvh},
               "clocks": clocks}
        if not a.no_cpu_baseline and world == 1:
            out["cpu_baseline"] = cpu_sample(sc, a.cpu_seconds)
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)
