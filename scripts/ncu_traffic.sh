# usage: bash scripts/ncu_traffic.sh <tag>   (under gpurun, one GPU)
# one --set full capture of a level-0 launch of each kernel (and of one k_tail launch) at the FULL bench workload (3840x2160x64):
# the DRAM bytes per launch that bench.py reports as roofline.traffic (profiles/summarize.py --traffic -> profiles/traffic.json)
set -e
T=${1:-traffic}
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/${T}_plain.log 2>&1
# launches per 128 Mi-ray batch: k_hit / k_shadow / k_shade at levels 0, 1, 2, then one k_tail; 4 batches per frame.
# Skipping 15 launches of a base name lands on the level-0 launch of the second frame's SECOND batch (the batches are column
# bands of the frame: the outer two hold little geometry and cost a third of the inner two).
for K in k_shadow:15 k_hit:15 k_shade:15 k_tail:5; do
  ncu --set full --clock-control none --import-source on -k regex:${K%%:*} -s ${K##*:} -c 1 -f -o gpurun_out/${T}_${K%%:*} $CMD > gpurun_out/${T}_ncu_${K%%:*}.log 2>&1 || echo "capture $K failed"
done
tail -1 gpurun_out/${T}_plain.log | cut -c1-300
