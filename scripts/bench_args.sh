#!/bin/bash
# usage: bench_args.sh "args1" "args2" ...  -- one short bench run per argument string
i=0
for v in "$@"; do
  i=$((i+1))
  python bench.py --steps 3 --warmup 3 --no-cpu-baseline $v 2>&1 | tail -1 > gpurun_out/arg_$i.json
  python - "$v" $i <<'PY'
import json, sys
v, i = sys.argv[1], sys.argv[2]
try:
    d = json.load(open(f"gpurun_out/arg_{i}.json"))
    print(v, "|", round(d["ms_per_step"], 2), "ms", round(d["value"]), "Mrays/s", {k[:8]: round(x, 1) for k, x in d["kernel_ms_per_frame"].items()}, "e2e", round(d["e2e"]["ms_per_step"], 2))
except Exception as e:
    print(v, "FAILED", e, open(f"gpurun_out/arg_{i}.json").read()[-800:])
PY
done
