# usage: bash scripts/scale_bench.sh <tag>   (under gpurun --gpus 8): bench at N = 1, 2, 4, 8 back to back + the bit-identity check
T=${1:-scale}
python bench.py --no-cpu-baseline 2>/dev/null | tail -1 > gpurun_out/${T}_n1.json
for N in 2 4 8; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + N)) bench.py --gpus $N --no-cpu-baseline 2>gpurun_out/${T}_n$N.err | tail -1 > gpurun_out/${T}_n$N.json
done
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 tests/run_multigpu_check.py 2>&1 | tail -3
for N in 1 2 4 8; do python -c "
import json,sys
d=json.load(open('gpurun_out/${T}_n$N.json')); print($N, round(d['ms_per_step'],2), 'ms', round(d['value']), d['unit'], 'e2e', round(d['e2e']['ms_per_step'],2))"; done
