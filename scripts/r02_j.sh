# FFMA2 slab tests + two-stream batch interleave: parity suite, bench with both streams / one stream, Novel scenes
run() { tag=$1; shift; env "$@" python bench.py --steps 8 --warmup 3 --no-cpu-baseline > gpurun_out/r02j_$tag.json 2>> gpurun_out/r02j.err; python - gpurun_out/r02j_$tag.json $tag <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    k=d["kernel_ms_per_frame"]
    print(sys.argv[2], "ms", round(d["ms_per_step"],2), "e2e", round(d["e2e"]["ms_per_step"],2), {a: round(b,2) for a,b in k.items()}, d["image_sha256"][:8])
except Exception as e: print(sys.argv[2], "ERR", e)
PY
}
python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/r02j_pytest.log; tail -3 gpurun_out/r02j_pytest.log
run two X=1
run one B200RT_STREAMS=1

python bench.py --steps 8 --warmup 3 --no-cpu-baseline --batch-rays 134217728 > gpurun_out/r02j_two128.json 2>> gpurun_out/r02j.err; python -c "
import json; d=json.loads(open('gpurun_out/r02j_two128.json').read().strip().splitlines()[-1]); print('two128', round(d['ms_per_step'],2))"
B200RT_STREAMS=1 python bench.py --steps 8 --warmup 3 --no-cpu-baseline --batch-rays 134217728 > gpurun_out/r02j_one128b.json 2>> gpurun_out/r02j.err; python -c "
import json; d=json.loads(open('gpurun_out/r02j_one128b.json').read().strip().splitlines()[-1]); print('one128', round(d['ms_per_step'],2))"
python scripts/bundled_scenes_timing.py Novel > gpurun_out/r02j_novel.jsonl 2>&1; cat gpurun_out/r02j_novel.jsonl
B200RT_STREAMS=1 python scripts/bundled_scenes_timing.py Novel > gpurun_out/r02j_novel_one.jsonl 2>&1; cat gpurun_out/r02j_novel_one.jsonl
tail -5 gpurun_out/r02j.err
