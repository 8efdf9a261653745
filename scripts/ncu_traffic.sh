# usage: bash scripts/ncu_traffic.sh <tag>   (under gpurun, one GPU)
# one --set full capture of a level-0 launch of each kernel at the FULL bench workload (3840x2160x64): the DRAM bytes
# per launch that bench.py reports as roofline.traffic (profiles/summarize.py --traffic turns the reports into profiles/traffic.json)
set -e
T=${1:-traffic}
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/${T}_plain.log 2>&1
for K in k_shadow k_hit k_shade; do
  ncu --set full --clock-control none --import-source on -k regex:$K -s 10 -c 1 -f -o gpurun_out/${T}_$K $CMD > gpurun_out/${T}_ncu_$K.log 2>&1
done
tail -1 gpurun_out/${T}_plain.log | cut -c1-300
