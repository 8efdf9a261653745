for v in "" _xldg; do B200RT_LIB=$PWD/python-raytracer_b200/csrc/libb200rt$v.so python bench.py --steps 6 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); k=d['kernel_ms_per_frame']; print('$v', round(d['ms_per_step'],2), round(k['k_hit'],2), round(k['k_shadow'],2), d['image_sha256'][:8])"; done
