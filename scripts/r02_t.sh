python -m pytest tests -m gpu -x -q 2>&1 | tail -12 > gpurun_out/r02t_pytest.log; tail -6 gpurun_out/r02t_pytest.log
python bench.py --steps 8 --warmup 3 --no-cpu-baseline > gpurun_out/r02t_bench.json 2> gpurun_out/r02t.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02t_bench.json').read().strip().splitlines()[-1])
print("ms", round(d["ms_per_step"],2), "e2e", round(d["e2e"]["ms_per_step"],2), {a: round(b,2) for a,b in d["kernel_ms_per_frame"].items()}, d["image_sha256"][:8], d["setup_ms"])
print({k: v for k, v in d["tallies_rank0_per_frame"]["k_shadow"].items()}, {k: v for k, v in d["tallies_rank0_per_frame"]["k_hit"].items()})
PY
tail -3 gpurun_out/r02t.err
