python -m pytest tests -m gpu -x -q 2>&1 | tail -4 > gpurun_out/r02q_pytest.log; tail -2 gpurun_out/r02q_pytest.log
python bench.py --steps 8 --warmup 3 --no-cpu-baseline > gpurun_out/r02q_bench.json 2> gpurun_out/r02q.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02q_bench.json').read().strip().splitlines()[-1])
print("ms", round(d["ms_per_step"],2), "e2e", round(d["e2e"]["ms_per_step"],2), {a: round(b,2) for a,b in d["kernel_ms_per_frame"].items()}, d["image_sha256"][:8])
PY
python scripts/bundled_scenes_timing.py Novel 2>&1 | cut -c1-200
