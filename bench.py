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
    ap.add_argument("--fan-in", default=None, choices=["shared", "gather"],
                    help="N > 1: 'shared' = every rank's k_resolve stores its tiles into rank 0's frame over NVLink peer access (default); "
                         "'gather' = NCCL gather of compact tile buffers + scatter kernel")
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


def frame_metrics(a):
    """Per-frame ncu sums per kernel for THIS workload (profiles/frame_metrics.json, written by profiles/summarize.py
    --frame from a one-frame `ncu --metrics smsp__inst_executed.sum,l1tex__data_pipe_lsu_wavefronts.sum,dram__bytes_*`
    capture at N=1): {kernel: {inst, lsu_wavefronts, dram_bytes, launches}} or {} when no capture matches."""
    try:
        with open(os.path.join(ROOT, "profiles", "frame_metrics.json")) as fh:
            return json.load(fh).get(f"{a.width}x{a.height}x{a.spp}", {})
    except Exception:
        return {}


# reference-formulation flop counts (FMA = 2): the whole-frame FP32 figure reported beside the issue-slot roofline
FLOPS = {"raygen": 60, "node4": 4 * 30, "tri": 110, "prim": 35, "light": 45}


def run_ours(a):
    import hashlib
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

    # roofline denominators measured here and now on this GPU (hand-written micro-benchmarks, csrc/rt_peaks.cuh)
    peaks = _abi.RtPeaks()
    _abi.check(lib, lib.rt_measure_peaks(C.byref(peaks), torch.cuda.current_stream().cuda_stream), "rt_measure_peaks")

    sharder = TileSharder(sc, rank, world, fan_in=a.fan_in)
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
                                                                  ctr.mesh_hits_smooth, ctr.shaded_hits, ctr.prim_tests],
                           dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tallies)
    tl = [int(x) for x in tallies.tolist()]
    rays_frame = sum(tl[:4])
    launches_frame = tl[4] + sharder.extra_launches_per_frame          # + rank 0's k_assemble in 'gather' mode

    # ---- timed region 1: inputs resident in HBM, image left on the device (rank 0).  Profiling instrumentation is OFF.
    sampler = ClockSampler(local_rank) if rank == 0 else None
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

    # ---- timed region 2 (e2e): the public call with HOST output; D2H of the frame inside the timed region
    host_img = torch.empty(W * H * 3, dtype=torch.float32).pin_memory() if rank == 0 else None
    host_u8 = torch.empty(W * H * 3, dtype=torch.uint8).pin_memory() if rank == 0 else None
    dev_u8 = torch.empty(W * H * 3, dtype=torch.uint8, device="cuda") if rank == 0 else None
    e2e_steps = a.steps

    job_all = sc.make_job(0, W, 0, H)

    def e2e_frame(u8=False):
        if world == 1 and not u8:
            # the reference-facing C-ABI call with a HOST buffer: rt_render_host (pinned destination: the part of the frame a batch
            # completes is copied while the next batches render); synchronous
            _abi.check(lib, lib.rt_render_host(h, C.byref(job_all), host_img.data_ptr(), None, torch.cuda.current_stream().cuda_stream), "rt_render_host")
            return
        if world > 1 and not u8 and sharder.fan_in == "shared":
            # every rank copies its slice of the finished frame (rank 0's HBM, read over NVLink) to one page-locked host frame in
            # shared memory over its OWN PCIe link; rank 0 owns the result
            sharder.render_to_host()
            torch.cuda.current_stream().synchronize()
            return
        img = sharder.render()
        if rank == 0:
            if u8:   # main.py:31-33 on the device (rt_finalize_u8): a quarter of the bytes cross PCIe
                _abi.check(lib, lib.rt_finalize_u8(img.data_ptr(), W, H, dev_u8.data_ptr(), torch.cuda.current_stream().cuda_stream), "rt_finalize_u8")
                host_u8.copy_(dev_u8, non_blocking=True)
            else:
                host_img.copy_(img, non_blocking=True)
        # every rank's caller waits for its call to finish (rank 0: owns the pixels) before the next frame starts
        torch.cuda.current_stream().synchronize()

    def timed_e2e(u8):
        for _ in range(min(2, a.warmup)):       # warm-up of THIS path: first copy into the fresh pinned buffer, first per-frame waits
            e2e_frame(u8)
        barrier()
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        f0.record()
        for _ in range(e2e_steps):
            e2e_frame(u8)
        f1.record()
        barrier()
        t = torch.tensor([f0.elapsed_time(f1)], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    ms2_total = timed_e2e(False)
    clocks = sampler.stop() if sampler else None
    ms3_total = timed_e2e(True)

    # ---- untimed: per-kernel split (CUDA-event pairs around every launch) and per-kernel tallies, from a separate pass
    prof_frames = max(1, min(3, a.steps))
    _abi.check(lib, lib.rt_scene_profile(h, 1), "rt_scene_profile")
    for _ in range(prof_frames):
        sharder.render()
    torch.cuda.synchronize()
    prof = _abi.RtProfile()
    _abi.check(lib, lib.rt_scene_profile_read(h, C.byref(prof)), "rt_scene_profile_read")
    _abi.check(lib, lib.rt_scene_profile(h, 0), "rt_scene_profile")

    # ---- untimed: the frame's checksum; N > 1: rank 0 also renders the whole frame alone and the two must be bit-identical
    image_sha = None
    equals_single = None
    if world > 1 and sharder.fan_in == "shared" and rank == 0:
        host_img = sharder.host_image                       # where the e2e frames of this run landed
    if rank == 0:
        image_sha = hashlib.sha256(host_img.numpy().tobytes()).hexdigest()
    if world > 1:
        if rank == 0:
            solo = TileSharder(sc, 0, 1)
            ref = solo.render().cpu()
            equals_single = bool(torch.equal(ref, host_img))
            del solo, ref
        dist.barrier()

    if rank == 0:
        hw = {}
        pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
        if os.path.exists(pk):
            with open(pk) as fh:
                hw = json.load(fh)
        hbm_peak, hbm_src = (hw.get("hbm_gbs"), "measured (MEASURED_PEAKS.json)") if hw.get("hbm_gbs") else (6650.0, "fallback")
        per_frame = lambda x: x / prof_frames                                   # noqa: E731
        kernel_ms = {"k_hit": per_frame(prof.ms_hit), "k_shadow": per_frame(prof.ms_shadow), "k_shade": per_frame(prof.ms_shade),
                     "k_resolve": per_frame(prof.ms_resolve),
                     "level1 k_hit": per_frame(prof.ms_l1_hit), "level1 k_shadow": per_frame(prof.ms_l1_shadow),
                     "level1 k_shade": per_frame(prof.ms_l1_shade), "levels>=2 (hit+shadow+shade)": per_frame(prof.ms_deep),
                     "k_tail": per_frame(prof.ms_tail)}
        level0 = {"k_hit": (prof.ms_hit, prof.n_hit), "k_shadow": (prof.ms_shadow, prof.n_shadow), "k_shade": (prof.ms_shade, prof.n_shade)}
        dom = max(level0, key=lambda k: level0[k][0])
        ms_dom_frame = per_frame(level0[dom][0])
        launches_dom = max(1, int(level0[dom][1]) // prof_frames)
        ms_all = sum(kernel_ms.values())
        # level-0 kernels of the bench scene (no CSG nodes) as ncu names them: template arguments <HAS_NODES, PRIMARY / LIGHT_ROWS>
        ncu_name = {"k_hit": "k_hit<0, 1>", "k_shadow": "k_shadow<0, 0>", "k_shade": "k_shade<1>"}[dom]
        fm = frame_metrics(a).get(ncu_name, {}) if world == 1 else {}
        # the dominant kernel is bound by the warp schedulers' issue slots (ncu: smsp__issue_active), not by HBM: its BVH is L1/L2
        # resident.  achieved = warp instructions the kernel executes per frame (ncu smsp__inst_executed.sum of the SAME workload,
        # committed under profiles/) / the kernel's time per frame measured live above; peak = issue rate measured live.
        sec = ms_dom_frame * 1e-3
        inst, lsu, dram = fm.get("inst"), fm.get("lsu_wavefronts"), fm.get("dram_bytes")
        roof = {"bound": "issue", "kernel": dom, "unit": "Gwarp-inst/s",
                "achieved": (inst / sec / 1e9) if inst and sec > 0 else None, "peak": peaks.issue_ginst,
                "frac": (inst / sec / 1e9 / peaks.issue_ginst) if inst and sec > 0 else None,
                "traffic": (dram / launches_dom) if dram else None,
                "peak_source": "measured in this run by rt_measure_peaks (csrc/rt_peaks.cuh: FFMA/IADD3/LOP3 mix, 16 warps per scheduler)",
                "count_source": "ncu smsp__inst_executed.sum / l1tex__data_pipe_lsu_wavefronts.sum / dram__bytes_* per frame of this "
                                "workload at N=1 (profiles/frame_metrics.json); time = CUDA events around each launch, this run",
                "kernel_ms_per_frame": ms_dom_frame, "launches_per_frame": launches_dom,
                "kernel_share_of_step": ms_dom_frame / max(1e-9, ms_all),
                "l1_lsu": {"achieved": (lsu / sec / 1e9) if lsu and sec > 0 else None, "peak": peaks.lsu_gwavefronts, "unit": "Gwavefronts/s (128 B)",
                           "frac": (lsu / sec / 1e9 / peaks.lsu_gwavefronts) if lsu and sec > 0 else None},
                "hbm": {"achieved": (dram / sec / 1e9) if dram and sec > 0 else None, "peak": hbm_peak, "unit": "GB/s", "peak_source": hbm_src,
                        "frac": (dram / sec / 1e9 / hbm_peak) if dram and sec > 0 else None}}
        if not fm:
            roof["note"] = "no ncu capture of this workload at N=1 under profiles/frame_metrics.json: counts unavailable, fractions null"
        # whole-frame FP32 figure from the device tallies x reference-formulation flop counts against the measured FMA peak
        n_lights = len(sc.lights)
        flops_frame = (FLOPS["raygen"] * tl[0] + FLOPS["node4"] * tl[5] + FLOPS["tri"] * tl[6] + FLOPS["prim"] * tl[10]
                       + FLOPS["light"] * n_lights * tl[9])
        step_s = ms_total / a.steps * 1e-3
        value = rays_frame * a.steps / (ms_total * 1e-3) / 1e6
        e2e_value = rays_frame * e2e_steps / (ms2_total * 1e-3) / 1e6
        knobs = {k: v for k, v in os.environ.items() if k.startswith("B200RT_")}
        out = {"metric": METRIC, "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
               "ms_per_step": ms_total / a.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
               "dtype": "f32", "data": "synthetic",
               "config": {**workload_config(a), "tiles": f"{sc.tile[0]}x{sc.tile[1]} interleaved over ranks", "fan_in": sharder.fan_in,
                          "l2_policy": "working set (BVH nodes+triangles+sample/queue buffers) exceeds the 126 MB L2; no flush between frames",
                          "wavefront": "batches of <= 64 Mi rays; level 0 of every batch on the caller's stream, the deeper levels + resolve of batch b on a "
                                       "high-priority side stream under level 0 of batch b + 1 (two work-buffer sets)",
                          "env_knobs": knobs},
               "rays_per_frame": dict(zip(RAY_KEYS, tl[:4])),
               "e2e": {"value": e2e_value, "unit": "Mrays/s", "ms_per_step": ms2_total / e2e_steps,
                       "h2d_bytes_per_step": C.sizeof(_abi.RtRenderJob), "d2h_bytes_per_step": W * H * 3 * 4,
                       "what": "frame rendered through the C ABI into pinned HOST memory every step as float32 (N = 1: rt_render_host, the frame leaves in chunks "
                               "as batches finish; N > 1: every rank copies its 1/N slice of the shared frame to one page-locked host frame over its own PCIe link; scene "
                               "resident: uploaded once like the reference's load_scene)"},
               "e2e_u8": {"value": rays_frame * e2e_steps / (ms3_total * 1e-3) / 1e6, "unit": "Mrays/s", "ms_per_step": ms3_total / e2e_steps,
                          "d2h_bytes_per_step": W * H * 3,
                          "what": "same, but main.py:31-33 (rot90, *255 truncation) done on the device and the uint8 PNG array copied"},
               "gpu_launches": launches_frame * a.steps,
               "image_sha256": image_sha, "image_equals_single_rank_render": equals_single,
               "roofline": roof,
               "fp32": {"achieved_tflops": flops_frame / step_s / 1e12, "peak_tflops": peaks.fp32_fma_tflops,
                        "frac": flops_frame / step_s / 1e12 / peaks.fp32_fma_tflops,
                        "flops_per_unit": FLOPS, "note": "whole frame; reference-formulation flop counts x device tallies"},
               "peaks_measured": peaks.as_dict(),
               "bytes_per_step_dram": sum(v.get("dram_bytes", 0) for v in frame_metrics(a).values()) or None,
               "tallies_rank0_per_frame": {k: {n: v for n, v in c.as_dict().items() if v and n not in ("kernel_launches", "error_flags")}
                                           for k, c in (("k_hit", prof.hit), ("k_shadow", prof.shadow), ("k_shade", prof.shade),
                                                        ("levels>=1", prof.secondary))},
               "kernel_ms_per_frame": kernel_ms,
               "kernel_ms_note": f"separate untimed pass of {prof_frames} frames with per-launch CUDA events, ONE launch at a time on one stream; `value` is timed with that "
                                 "instrumentation off and with the deeper levels overlapped (their sum here exceeds ms_per_step by what the overlap hides)",
               "setup_ms": {"scene_python": 1e3 * t_scene, **t_ingest, "upload": 1e3 * t_upload, "lbvh_build": 1e3 * t_bvh},
               "clocks": clocks}
        if not a.no_cpu_baseline and world == 1:
            out["cpu_baseline"] = cpu_sample(sc, a.cpu_seconds)
        print(json.dumps(out), flush=True)
    sharder.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)
