"""Dev container only: renders tests/fuzz_scenes.py cases with the UNMODIFIED reference (tier 0) and with the C oracle
(tier 1) and reports every difference.    python oracle/fuzz_check.py SEED [SEED ...]"""
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
for p in (ROOT, HERE, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)
import fuzz_scenes  # noqa: E402
import ref_runner  # noqa: E402


def check(seed, jitter_seed=0x5EED):
    import oracle
    import python_raytracer_b200 as rt
    rt.scene_parser.VERBOSE = False
    data = fuzz_scenes.make(seed)
    t0 = time.time()
    ref, cpu = ref_runner.render_reference("scenes/TwoSpheresPlane.json", overrides=data, seed=jitter_seed)
    t1 = time.time()
    sc = rt.scene_parser.load_scene_dict(data, os.path.join(ROOT, "tests", "golden", "scenes"))
    sc.jitter_seed = jitter_seed
    img, ctr = oracle.render(sc.flatten(), rt._abi)
    a, b = ref_runner.to_png_array(ref).astype(int), ref_runner.to_png_array(img).astype(int)
    d8 = np.abs(a - b).max(axis=2)
    exact = np.array_equal(ref.astype(np.float32), img.astype(np.float32))
    print(f"seed {seed}: ref {t1-t0:.0f}s wall ({cpu:.0f}s cpu)  float-exact={exact}  u8 differing pixels={int((d8 > 0).sum())} "
          f"max={int(d8.max())}  max|df|={np.abs(ref - img).max():.3g}  errors=0x{ctr['error_flags']:x}", flush=True)
    if not exact:
        bad = np.argwhere(np.abs(ref - img).max(axis=2) > 0)
        print("   first differing (col,row):", bad[:8].tolist())
    return exact


if __name__ == "__main__":
    ok = all([check(int(s)) for s in sys.argv[1:]])
    sys.exit(0 if ok else 1)
