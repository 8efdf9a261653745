set -e
CMD="python bench.py --width 1920 --height 1080 --spp 16 --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/r01c_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r01c_launches.csv $CMD > gpurun_out/r01c_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_primary -s 3 -c 1 -f -o gpurun_out/r01c_primary $CMD > gpurun_out/r01c_ncu2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_secondary -s 27 -c 2 -f -o gpurun_out/r01c_secondary $CMD > gpurun_out/r01c_ncu3.log 2>&1
cat gpurun_out/r01c_plain.log | tail -2
