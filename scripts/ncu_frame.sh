# usage: bash scripts/ncu_frame.sh <tag> [frame_once.py args]   (under gpurun, one GPU)
# Per-launch instruction / L1-wavefront / DRAM-byte counts of EVERY launch of ONE frame (after two warm-up frames): the per-frame
# sums per kernel are what bench.py's roofline divides by the live kernel time (profiles/summarize.py --frame -> profiles/frame_metrics.json).
T=${1:-frame}
shift
M=smsp__inst_executed.sum,smsp__thread_inst_executed.sum,l1tex__data_pipe_lsu_wavefronts.sum,dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,l1tex__data_pipe_lsu_wavefronts.sum.pct_of_peak_sustained_elapsed,lts__t_sector_hit_rate.pct,l1tex__t_sector_hit_rate.pct
python scripts/frame_once.py "$@" > gpurun_out/${T}_plain.log 2>&1 &&
ncu --metrics $M --clock-control none --profile-from-start off --csv --log-file gpurun_out/${T}_metrics.csv python scripts/frame_once.py "$@" > gpurun_out/${T}_ncu.log 2>&1
tail -2 gpurun_out/${T}_ncu.log
