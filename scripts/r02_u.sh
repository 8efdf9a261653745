# variants must REPLACE libb200rt.so: the torch extension the bench renders through links to it by name
D=python-raytracer_b200/csrc
cp $D/libb200rt.so $D/libb200rt_keep.so
for v in _keep _xldg _s9 _h7; do cp $D/libb200rt$v.so $D/libb200rt.so; python bench.py --steps 6 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); k=d['kernel_ms_per_frame']; print('$v', round(d['ms_per_step'],2), round(k['k_hit'],2), round(k['k_shadow'],2), d['image_sha256'][:8])"; done
cp $D/libb200rt_keep.so $D/libb200rt.so
