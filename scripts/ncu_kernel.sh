# usage: ncu_kernel.sh <tag> <kernel regex> [skip]   -- one --set full capture of one launch at 1920x1080x16
set -e
CMD="python bench.py --width 1920 --height 1080 --spp 16 --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/ncu_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:$2 -s ${3:-3} -c 1 -f -o gpurun_out/$1 $CMD > gpurun_out/ncu_full.log 2>&1
tail -1 gpurun_out/ncu_plain.log | cut -c1-200
