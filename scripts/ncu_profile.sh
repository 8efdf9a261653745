# usage: bash scripts/ncu_profile.sh <tag>   (under gpurun, one GPU)
# launch list of the bench command (gpu__time_duration per launch) + one --set full capture of each level-0 kernel
set -e
T=${1:-prof}
CMD="python bench.py --width 1920 --height 1080 --spp 16 --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/${T}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${T}_launches.csv $CMD > gpurun_out/${T}_ncu1.log 2>&1
# per frame: 10 launches of each of k_hit / k_shadow / k_shade (level 0 first); skip the first frame
for K in k_shadow k_hit k_shade; do
  ncu --set full --clock-control none --import-source on -k regex:$K -s 10 -c 1 -f -o gpurun_out/${T}_$K $CMD > gpurun_out/${T}_ncu_$K.log 2>&1
done
tail -1 gpurun_out/${T}_plain.log | cut -c1-400
