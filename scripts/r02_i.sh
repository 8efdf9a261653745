# static-scratch entry point alone in a fresh process, then the whole GPU suite (it runs last there), then the bench
set -x
python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -15 > gpurun_out/r02i_noalloc.log; tail -5 gpurun_out/r02i_noalloc.log
python -m pytest tests -m gpu -x -q --durations=15 2>&1 | tail -40 > gpurun_out/r02i_pytest.log; tail -25 gpurun_out/r02i_pytest.log
python bench.py --steps 10 --warmup 3 > gpurun_out/r02i_bench.json 2> gpurun_out/r02i_bench.err; tail -c 3000 gpurun_out/r02i_bench.json
python scripts/bundled_scenes_timing.py Novel > gpurun_out/r02i_novel.jsonl 2>&1; cat gpurun_out/r02i_novel.jsonl
