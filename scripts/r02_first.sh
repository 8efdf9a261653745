# first GPU call of round 2: parity suite, bench (default + streaming-record variant + L2 persist), Novel scene timings, one-frame ncu counts
set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -40 > gpurun_out/r02a_pytest.log; tail -5 gpurun_out/r02a_pytest.log
python bench.py --steps 10 --warmup 3 > gpurun_out/r02a_bench.json 2> gpurun_out/r02a_bench.err; tail -c 1500 gpurun_out/r02a_bench.json
B200RT_LIB=$PWD/python-raytracer_b200/csrc/libb200rt_stream.so python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r02a_bench_stream.json 2> gpurun_out/r02a_bench_stream.err
B200RT_L2_PERSIST=64 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r02a_bench_l2.json 2> gpurun_out/r02a_bench_l2.err
B200RT_L2_PERSIST=64 B200RT_LIB=$PWD/python-raytracer_b200/csrc/libb200rt_stream.so python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r02a_bench_l2_stream.json 2> gpurun_out/r02a_bench_l2_stream.err
python scripts/bundled_scenes_timing.py Novel > gpurun_out/r02a_novel.jsonl 2>&1; cat gpurun_out/r02a_novel.jsonl
bash scripts/ncu_frame.sh r02a_frame
for f in gpurun_out/r02a_bench*.json; do python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], round(d["ms_per_step"],2), round(d["e2e"]["ms_per_step"],2), {k:round(v,2) for k,v in d["kernel_ms_per_frame"].items()})
except Exception as e: print(sys.argv[1], "ERR", e)
PY
done
