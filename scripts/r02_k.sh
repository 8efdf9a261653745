# stream modes: 1 serial, 2 alternate batches, 3 level-0 / deeper-levels pipeline; batch sizes
run() { tag=$1; shift; env "$@" > gpurun_out/r02k_$tag.json 2>> gpurun_out/r02k.err; python - gpurun_out/r02k_$tag.json $tag <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[2], "ms", round(d["ms_per_step"],2), "e2e", round(d["e2e"]["ms_per_step"],2), d["image_sha256"][:8])
except Exception as e: print(sys.argv[2], "ERR", e)
PY
}
B="python bench.py --steps 8 --warmup 3 --no-cpu-baseline"
run m3_64 $B
run m2_64 B200RT_STREAMS=2 $B
run m1_64 B200RT_STREAMS=1 $B
run m3_128 $B --batch-rays 134217728
run m3_32 $B --batch-rays 33554432
run m3_96 $B --batch-rays 100663296
python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/r02k_pytest.log; tail -3 gpurun_out/r02k_pytest.log
python scripts/bundled_scenes_timing.py Novel > gpurun_out/r02k_novel.jsonl 2>&1; cat gpurun_out/r02k_novel.jsonl
tail -5 gpurun_out/r02k.err
