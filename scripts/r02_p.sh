# round-2 profile collection: per-class CSG timings, per-frame ncu counts at 4K x 64, CSG kernel full captures, bench launch list
python -m pytest tests -m gpu -x -q 2>&1 | tail -4 > gpurun_out/r02p_pytest.log; tail -2 gpurun_out/r02p_pytest.log
python scripts/csg_profile.py > gpurun_out/r02p_csg_profile.jsonl 2> gpurun_out/r02p_csg_profile.err; cut -c1-700 gpurun_out/r02p_csg_profile.jsonl
bash scripts/ncu_frame.sh r02p_frame
bash scripts/ncu_csg.sh r02p_csg
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r02p_bench_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02p_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r02p_ncu_ll.log 2>&1
ls -la gpurun_out | tail -20
