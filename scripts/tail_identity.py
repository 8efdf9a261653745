"""Are the images independent of the level at which k_tail takes over (B200RT_TAIL_LEVEL)?  Renders a few scenes once per
setting in a subprocess and compares the float32 frames bit for bit."""
import os, subprocess, sys, json
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if len(sys.argv) > 1 and sys.argv[1] == "child":
    sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
    from conftest import load_case
    from python_raytracer_b200 import synthetic
    scenes = {"MirrorRefraction": load_case("MirrorRefraction"), "Fuzz12": load_case("Fuzz12"),
              "synthetic": synthetic.make_scene(width=480, height=270, aa_samples=4, grid=(128, 64), n_spheres=64)}
    for name, sc in scenes.items():
        np.save(os.path.join(ROOT, "gpurun_out", f"tailid_{name}_{sys.argv[2]}.npy"), sc.render_region(0, sc.vc.width, 0, sc.vc.height))
    sys.exit(0)
levels = ["0", "1", "2", "3", "5"]
for lv in levels:
    subprocess.run([sys.executable, __file__, "child", lv], env={**os.environ, "B200RT_TAIL_LEVEL": lv}, check=True)
for name in ("MirrorRefraction", "Fuzz12", "synthetic"):
    base = np.load(os.path.join(ROOT, "gpurun_out", f"tailid_{name}_0.npy"))
    for lv in levels[1:]:
        img = np.load(os.path.join(ROOT, "gpurun_out", f"tailid_{name}_{lv}.npy"))
        d = np.abs(img - base).max(axis=2)
        bad = np.argwhere(d > 0)
        print(json.dumps({"scene": name, "tail_level": lv, "pixels_differing": int(len(bad)), "max_abs": float(d.max()),
                          "first": bad[:5].tolist(), "vals": [[base[tuple(b)].tolist(), img[tuple(b)].tolist()] for b in bad[:2]]}))
for f in os.listdir(os.path.join(ROOT, "gpurun_out")):
    if f.startswith("tailid_"):
        os.remove(os.path.join(ROOT, "gpurun_out", f))
