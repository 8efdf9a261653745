CMD="python bench.py --width 1920 --height 1080 --spp 16 --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/ll_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/ll_launches.csv $CMD > gpurun_out/ll_ncu.log 2>&1
