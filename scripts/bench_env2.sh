#!/bin/bash
# usage: bench_env2.sh "A=1 B=2" "A=3 B=4" ... -- one short bench run per environment assignment list
i=0
for v in "$@"; do
  i=$((i+1)); echo "$v"
  env $v python bench.py --steps 3 --warmup 3 --no-cpu-baseline 2>&1 | tail -1 > gpurun_out/env_$i.json
  python - $i <<'PY'
import json, sys
i = sys.argv[1]
try:
    d = json.load(open(f"gpurun_out/env_{i}.json"))
    print(" |", round(d["ms_per_step"], 2), "ms", round(d["value"]), "Mrays/s", {k[:8]: round(x, 1) for k, x in d["kernel_ms_per_frame"].items()}, "e2e", round(d["e2e"]["ms_per_step"], 2))
except Exception as e:
    print("FAILED", e, open(f"gpurun_out/env_{i}.json").read()[-800:])
PY
done
