# end-of-round check: smoke(), the default bench line (with the CPU baseline), the reference arm, CSG captures of the final kernels
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2 | cut -c1-400
python bench.py > gpurun_out/r02_final_bench.json 2> gpurun_out/r02_final_bench.err; tail -c 600 gpurun_out/r02_final_bench.json
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r02_final_reference.json 2>/dev/null; cut -c1-300 gpurun_out/r02_final_reference.json
bash scripts/ncu_csg.sh r02z_csg
python scripts/csg_profile.py > gpurun_out/r02z_csg_profile.jsonl 2>/dev/null; cut -c1-330 gpurun_out/r02z_csg_profile.jsonl
