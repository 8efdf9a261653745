#!/bin/bash
# usage: bench_variants.sh lib1.so lib2.so ...   (files under python-raytracer_b200/csrc); one bench line each
for v in "$@"; do
  B200RT_LIB=$PWD/python-raytracer_b200/csrc/$v python bench.py --steps 3 --warmup 3 --no-cpu-baseline 2>&1 | tail -1 > gpurun_out/var_$v.json
  python - "$v" <<'PY'
import json, sys
v = sys.argv[1]
try:
    d = json.load(open(f"gpurun_out/var_{v}.json"))
    print(v, round(d["ms_per_step"], 2), "ms", round(d["value"]), "Mrays/s", {k[:8]: round(x, 1) for k, x in d["kernel_ms_per_frame"].items()}, d["roofline"]["kernel"], round(d["roofline"]["frac"], 3))
except Exception as e:
    print(v, "FAILED", e, open(f"gpurun_out/var_{v}.json").read()[-500:])
PY
done
