for v in "" _nomesh; do echo "variant $v"; B200RT_LIB=$PWD/python-raytracer_b200/csrc/libb200rt$v.so python scripts/bundled_scenes_timing.py Novel 2>&1 | cut -c1-170; done
