# usage: bash scripts/ncu_csg.sh <tag>  (under gpurun) -- --set full captures of the CSG (HAS_NODES) kernels on NovelScene1 at 1024x512x8
set -e
T=${1:-csg}
cat > /tmp/csg_run.py <<'PY'
import sys, os
ROOT = os.environ.get("GRAFT_REPO_ROOT", "/root/repo")
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "oracle")]
from conftest import load_case
sc = load_case("NovelScene1_small", {"resolution": [1024, 512], "AA": {"jitter": True, "samples": 8}})
for _ in range(3):
    sc.render_u8()
print(sc.last_counters)
PY
python /tmp/csg_run.py > gpurun_out/${T}_plain.log 2>&1
for K in k_shadow k_hit; do
  ncu --set full --clock-control none --import-source on -k regex:$K -s 10 -c 1 -f -o gpurun_out/${T}_$K python /tmp/csg_run.py > gpurun_out/${T}_ncu_$K.log 2>&1
done
tail -1 gpurun_out/${T}_plain.log | cut -c1-300
