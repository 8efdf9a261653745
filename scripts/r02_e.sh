set -x
python -m pytest tests/test_multigpu.py -m gpu -q 2>&1 | tail -15
for f in shared gather; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 --fan-in $f > gpurun_out/r02e_bench_n2_$f.json 2> gpurun_out/r02e_bench_n2_$f.err
tail -3 gpurun_out/r02e_bench_n2_$f.err
python - gpurun_out/r02e_bench_n2_$f.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], "value ms", round(d["ms_per_step"],2), "e2e", round(d["e2e"]["ms_per_step"],2), "e2e_u8", round(d["e2e_u8"]["ms_per_step"],2), d["image_sha256"][:12], d["image_equals_single_rank_render"], d["config"]["fan_in"])
except Exception as e: print(sys.argv[1], "ERR", e)
PY
done
