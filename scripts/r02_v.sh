# final profile collection of round 2 (after PLOC, parameter-bank constants, scene in global memory for the general kernels)
python -m pytest tests -m gpu -x -q 2>&1 | tail -4 > gpurun_out/r02v_pytest.log; tail -2 gpurun_out/r02v_pytest.log
python scripts/csg_profile.py > gpurun_out/r02v_csg_profile.jsonl 2> gpurun_out/r02v_csg_profile.err; cut -c1-420 gpurun_out/r02v_csg_profile.jsonl
bash scripts/ncu_frame.sh r02v_frame
bash scripts/ncu_csg.sh r02v_csg
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r02v_bench_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02v_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r02v_ncu_ll.log 2>&1
# full captures of the level-0 kernels on a heavy batch of the 4K x 64 frame (frame_once: 2 warm frames, then one profiled frame of 8 batches; batch 4 = launches 3..)
for K in k_shadow k_hit k_shade; do
  ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:"$K<0, [01]>|$K<1>" -s 3 -c 1 -f -o gpurun_out/r02v_4k64_$K python scripts/frame_once.py > gpurun_out/r02v_ncu_$K.log 2>&1
done
python scripts/bundled_scenes_timing.py > gpurun_out/r02v_bundled_scenes.jsonl 2>&1; tail -4 gpurun_out/r02v_bundled_scenes.jsonl | cut -c1-200
ls gpurun_out | grep r02v
