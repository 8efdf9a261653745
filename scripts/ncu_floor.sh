set -e
CMD="python bench.py --width 1920 --height 1080 --spp 16 --steps 2 --warmup 3 --no-cpu-baseline --grid 4 4 --spheres 0"
$CMD > gpurun_out/ncu_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_shadow -s 10 -c 1 -f -o gpurun_out/floor_shadow $CMD > gpurun_out/ncu_full.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_hit -s 10 -c 1 -f -o gpurun_out/floor_hit $CMD > gpurun_out/ncu_full2.log 2>&1
