set -x
python -m pytest tests/test_gpu_parity.py -m gpu -q -k "large_mesh or lbvh" 2>&1 | tail -15
L=$PWD/python-raytracer_b200/csrc
python scripts/frame_once.py > gpurun_out/r02c_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:k_shadow -s 3 -c 1 -f -o gpurun_out/r02c_shadow_packet python scripts/frame_once.py > gpurun_out/r02c_ncu1.log 2>&1
B200RT_LIB=$L/libb200rt_nopacket.so python scripts/frame_once.py > gpurun_out/r02c_plain2.log 2>&1 &&
B200RT_LIB=$L/libb200rt_nopacket.so ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:k_shadow -s 3 -c 1 -f -o gpurun_out/r02c_shadow_perlane python scripts/frame_once.py > gpurun_out/r02c_ncu2.log 2>&1
tail -3 gpurun_out/r02c_ncu1.log gpurun_out/r02c_ncu2.log
