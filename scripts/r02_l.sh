# source-level captures of the level-0 kernels at 1920x1080x16 (one launch each)
bash scripts/ncu_kernel.sh r02l_shadow "k_shadow" 3
bash scripts/ncu_kernel.sh r02l_hit "k_hit" 3
ls -la gpurun_out/*.ncu-rep
