set -x
python -m pytest tests -m gpu -q 2>&1 | tail -30 > gpurun_out/r02f_pytest.log; tail -6 gpurun_out/r02f_pytest.log
python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r02f_bench.json 2> gpurun_out/r02f_bench.err; tail -3 gpurun_out/r02f_bench.err
python - gpurun_out/r02f_bench.json <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(round(d["ms_per_step"],2), round(d["e2e"]["ms_per_step"],2), round(d["e2e_u8"]["ms_per_step"],2), {k:round(v,2) for k,v in d["kernel_ms_per_frame"].items()}, d["image_sha256"][:12], d["roofline"]["frac"])
PY
