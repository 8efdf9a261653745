set -x
python -m pytest tests -m gpu -q 2>&1 | tail -30 > gpurun_out/r02g_pytest.log; tail -6 gpurun_out/r02g_pytest.log
python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r02g_bench.json 2> gpurun_out/r02g_bench.err; tail -3 gpurun_out/r02g_bench.err
python - gpurun_out/r02g_bench.json <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(round(d["ms_per_step"],2), round(d["e2e"]["ms_per_step"],2), round(d["e2e_u8"]["ms_per_step"],2), {k:round(v,2) for k,v in d["kernel_ms_per_frame"].items()}, d["image_sha256"][:12], d["roofline"]["frac"])
PY
python scripts/frame_once.py > gpurun_out/r02g_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:k_hit -s 3 -c 1 -f -o gpurun_out/r02g_hit python scripts/frame_once.py > gpurun_out/r02g_ncu1.log 2>&1
tail -n 3 gpurun_out/r02g_ncu1.log
