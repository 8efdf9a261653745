set -x
python -m pytest tests -m gpu -q 2>&1 | tail -40 > gpurun_out/r02b_pytest.log; tail -8 gpurun_out/r02b_pytest.log
L=$PWD/python-raytracer_b200/csrc
python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r02b_bench_packet8.json 2> gpurun_out/r02b_bench.err
for v in nopacket pk7 pk6; do B200RT_LIB=$L/libb200rt_$v.so python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r02b_bench_$v.json 2>> gpurun_out/r02b_bench.err; done
python scripts/bundled_scenes_timing.py Novel > gpurun_out/r02b_novel.jsonl 2>&1; cat gpurun_out/r02b_novel.jsonl
for f in gpurun_out/r02b_bench*.json; do python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], round(d["ms_per_step"],2), round(d["e2e"]["ms_per_step"],2), round(d["e2e_u8"]["ms_per_step"],2), {k:round(v,2) for k,v in d["kernel_ms_per_frame"].items()}, d["image_sha256"][:12], d["peaks_measured"])
except Exception as e: print(sys.argv[1], "ERR", e)
PY
done
tail -5 gpurun_out/r02b_bench.err
