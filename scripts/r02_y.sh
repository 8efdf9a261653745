python -m pytest tests/test_multigpu.py -m gpu -q 2>&1 | tail -15
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29502 bench.py --gpus 2 --no-cpu-baseline 2>gpurun_out/r02y_n2.err | tail -1 > gpurun_out/r02y_n2.json
python -c "
import json; d=json.load(open('gpurun_out/r02y_n2.json')); print(2, round(d['ms_per_step'],2), 'e2e', round(d['e2e']['ms_per_step'],2), 'u8', round(d['e2e_u8']['ms_per_step'],2), d['image_sha256'][:8], d['image_equals_single_rank_render'])"
tail -3 gpurun_out/r02y_n2.err
