"""Turn gpurun_out/ ncu artefacts into the small text summaries committed under profiles/.
usage: python profiles/summarize.py <tag> [launches.csv] [report.ncu-rep]"""
import csv
import subprocess
import sys
from collections import defaultdict

KEYS = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct", "lts__t_bytes.sum", "l1tex__t_bytes.sum",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "smsp__sass_inst_executed_op_local_ld.sum",
        "smsp__sass_inst_executed_op_local_st.sum",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
        "gcc__cache_requests_type_instruction.sum.pct_of_peak_sustained_elapsed",
        "l1tex__data_pipe_lsu_wavefronts.sum.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum"]


def launches(path, out):
    rows = [r for r in csv.reader(open(path)) if len(r) > 10]
    h = rows[0]
    ki, vi, mi = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Name")
    data = [(r[ki].split("(")[0], float(r[vi].replace(",", ""))) for r in rows[1:] if r[mi] == "gpu__time_duration.sum"]
    agg, cnt = defaultdict(float), defaultdict(int)
    for k, v in data:
        agg[k] += v; cnt[k] += 1
    tot = sum(agg.values())
    out.write(f"# ncu --metrics gpu__time_duration.sum --clock-control none: {len(data)} launches, cold-cache serialised (compare SHARES)\n")
    out.write(f"{'kernel':48s} {'launches':>8s} {'total_ms':>10s} {'share':>7s} {'mean_us':>10s}\n")
    for k, v in sorted(agg.items(), key=lambda x: -x[1]):
        out.write(f"{k[:48]:48s} {cnt[k]:8d} {v/1e6:10.3f} {100*v/tot:6.1f}% {v/cnt[k]/1e3:10.1f}\n")
    out.write("\n# last launches in order (us)\n")
    for k, v in data[-26:]:
        out.write(f"{k[:48]:48s} {v/1e3:10.1f}\n")


def report(path, out):
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    h, u = rows[0], rows[1]
    for r in rows[2:]:
        out.write(f"\n# ncu --set full: {r[h.index('Kernel Name')][:80]}  (ID {r[0]})\n")
        for k in KEYS:
            if k in h:
                i = h.index(k)
                out.write(f"{k:90s} {r[i]:>16s} {u[i]}\n")


def traffic(workload, paths):
    """profiles/traffic.json: "<kernel>@<WxHxspp>" -> dram__bytes_read.sum + dram__bytes_write.sum of ONE launch
    (bench.py reports it as roofline.traffic for the same workload)."""
    import json
    import os
    unit_scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    dst = "profiles/traffic.json"
    table = json.load(open(dst)) if os.path.exists(dst) else {}
    for path in paths:
        raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        rows = list(csv.reader(raw.splitlines()))
        h, u, r = rows[0], rows[1], rows[2]
        tot = 0.0
        for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            i = h.index(k)
            tot += float(r[i].replace(",", "")) * unit_scale[u[i]]
        kern = r[h.index("Kernel Name")].split("(")[0].split("<")[0].replace("void ", "").strip()
        table[f"{kern}@{workload}"] = tot
    json.dump(table, open(dst, "w"), indent=1, sort_keys=True)
    print(json.dumps(table, indent=1))


def frame(workload, csv_path, tag):
    """profiles/frame_metrics.json["<WxHxspp>"][kernel] = per-FRAME sums over the launches of one frame (ncu --metrics ... of
    scripts/ncu_frame.sh): inst (smsp__inst_executed.sum), thread_inst, lsu_wavefronts, dram_bytes, launches, ncu_ms; plus a
    text table under profiles/<tag>.txt."""
    import json
    import os
    rows = [r for r in csv.reader(open(csv_path)) if len(r) > 10]
    h = rows[0]
    ki, vi, mi, ui, ii = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Name"), h.index("Metric Unit"), h.index("ID")
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-6, "us": 1e-3, "ms": 1.0, "nsecond": 1e-6, "usecond": 1e-3, "msecond": 1.0}
    per = defaultdict(lambda: defaultdict(float))
    pct = defaultdict(lambda: defaultdict(list))
    ids = defaultdict(set)
    for r in rows[1:]:
        k = r[ki].split("(")[0].replace("void ", "").strip()          # template arguments kept: k_shadow<0, 0> is the level-0 kernel
        v = float(r[vi].replace(",", "")) * scale.get(r[ui], 1.0)
        ids[k].add(r[ii])
        if "pct" in r[mi]:
            pct[k][r[mi]].append(v)
        else:
            per[k][r[mi]] += v
    table = {}
    for k, m in per.items():
        table[k] = {"inst": m.get("smsp__inst_executed.sum", 0.0), "thread_inst": m.get("smsp__thread_inst_executed.sum", 0.0),
                    "lsu_wavefronts": m.get("l1tex__data_pipe_lsu_wavefronts.sum", 0.0),
                    "dram_bytes": m.get("dram__bytes_read.sum", 0.0) + m.get("dram__bytes_write.sum", 0.0),
                    "launches": len(ids[k]), "ncu_ms": m.get("gpu__time_duration.sum", 0.0),
                    "issue_active_pct_mean": sum(pct[k].get("smsp__issue_active.avg.pct_of_peak_sustained_active", [0])) / max(1, len(ids[k])),
                    "lsu_pct_mean": sum(pct[k].get("l1tex__data_pipe_lsu_wavefronts.sum.pct_of_peak_sustained_elapsed", [0])) / max(1, len(ids[k])),
                    "l1_hit_pct_mean": sum(pct[k].get("l1tex__t_sector_hit_rate.pct", [0])) / max(1, len(ids[k])),
                    "l2_hit_pct_mean": sum(pct[k].get("lts__t_sector_hit_rate.pct", [0])) / max(1, len(ids[k]))}
    dst = "profiles/frame_metrics.json"
    allm = json.load(open(dst)) if os.path.exists(dst) else {}
    allm[workload] = table
    json.dump(allm, open(dst, "w"), indent=1, sort_keys=True)
    with open(f"profiles/{tag}.txt", "w") as out:
        out.write(f"# one frame of {workload}: per-kernel sums over its launches (ncu --metrics, --clock-control none; times are serialised, cold)\n")
        out.write(f"{'kernel':16s} {'launches':>8s} {'ncu_ms':>9s} {'warp inst':>14s} {'lanes/inst':>10s} {'lsu wavefronts':>15s} {'dram GB':>9s} {'issue%':>7s} {'lsu%':>6s} {'L1hit%':>7s} {'L2hit%':>7s}\n")
        for k, t in sorted(table.items(), key=lambda kv: -kv[1]["ncu_ms"]):
            out.write(f"{k[:16]:16s} {t['launches']:8d} {t['ncu_ms']:9.3f} {t['inst']:14.4g} {t['thread_inst'] / max(1.0, t['inst']):10.2f} "
                      f"{t['lsu_wavefronts']:15.4g} {t['dram_bytes'] / 1e9:9.3f} {t['issue_active_pct_mean']:7.1f} {t['lsu_pct_mean']:6.1f} "
                      f"{t['l1_hit_pct_mean']:7.1f} {t['l2_hit_pct_mean']:7.1f}\n")
    print(open(f"profiles/{tag}.txt").read())


if __name__ == "__main__":
    if sys.argv[1] == "--frame":     # python profiles/summarize.py --frame 3840x2160x64 gpurun_out/x_metrics.csv r02a_frame_4k64
        frame(sys.argv[2], sys.argv[3], sys.argv[4])
        sys.exit(0)
    if sys.argv[1] == "--traffic":   # python profiles/summarize.py --traffic 3840x2160x64 a.ncu-rep b.ncu-rep ...
        traffic(sys.argv[2], sys.argv[3:])
        sys.exit(0)
    tag = sys.argv[1]
    with open(f"profiles/{tag}.txt", "w") as out:
        for p in sys.argv[2:]:
            if p.endswith(".csv"):
                launches(p, out)
            else:
                report(p, out)
    print(open(f"profiles/{tag}.txt").read())
