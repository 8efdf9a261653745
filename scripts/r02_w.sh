python -m pytest tests -m gpu -x -q 2>&1 | tail -12 > gpurun_out/r02w_pytest.log; tail -3 gpurun_out/r02w_pytest.log
python bench.py --steps 8 --warmup 3 --no-cpu-baseline > gpurun_out/r02w_bench.json 2> gpurun_out/r02w.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02w_bench.json').read().strip().splitlines()[-1])
print("ms", round(d["ms_per_step"],2), "e2e", round(d["e2e"]["ms_per_step"],2), "u8", round(d["e2e_u8"]["ms_per_step"],2), d["image_sha256"][:8], d["roofline"]["frac"], d["roofline"]["l1_lsu"]["frac"])
PY
tail -3 gpurun_out/r02w.err
# launch list of the bench command: the LBVH build (PLOC rounds) is ~450 launches, the frames follow
ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 440 -c 500 --csv --log-file gpurun_out/r02w_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r02w_ncu_ll.log 2>&1
for K in k_shadow k_hit k_shade; do
  ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:$K -s 9 -c 1 -f -o gpurun_out/r02w_4k64_$K python scripts/frame_once.py > gpurun_out/r02w_ncu_$K.log 2>&1
done
ls gpurun_out | grep r02w
