"""How evenly are the 32 lanes of a k_primary warp loaded?  Uses tests/hostsim (host build of the per-ray device
functions) to count the LBVH nodes each (pixel, sample) ray visits in its first cast_ray level, groups rays into warps
exactly like k_primary does (ray i -> pixel i / N in tile order, slot i % N) and prints sum / (32 * max) per warp."""
import ctypes as C, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "tests", "hostsim"), os.path.join(ROOT, "oracle")]
import hostsim, python_raytracer_b200 as rt
from python_raytracer_b200 import synthetic
W, H, spp = (int(x) for x in (sys.argv[1:4] if len(sys.argv) > 3 else (480, 270, 16)))
rt.scene_parser.VERBOSE = False
sc = synthetic.make_scene(width=W, height=H, aa_samples=spp, grid=(1024, 512), n_spheres=64)
fs = sc.flatten()
job = sc.make_job(0, W, 0, H)
cols, rows, valid = hostsim.job_pixels(job)
cols, rows = cols[valid], rows[valid]
# a sample of tiles, keeping tile order
sel = np.arange(0, len(cols) - 64, 64 * 7)[:, None] + np.arange(64)[None, :]
cols, rows = np.ascontiguousarray(cols[sel.reshape(-1)]), np.ascontiguousarray(rows[sel.reshape(-1)])
out = np.zeros(len(cols) * spp, dtype=np.uint32)
L = hostsim.lib()
L.hostsim_node_counts.argtypes = [C.c_void_p] * 3 + [C.c_int64, C.c_void_p]
assert L.hostsim_node_counts(C.addressof(fs.desc), cols.ctypes.data, rows.ctypes.data, len(cols), out.ctypes.data) == 0
w = out.reshape(-1, 32).astype(np.float64)
mx = w.max(axis=1)
busy = mx > 0
print(f"{W}x{H}x{spp}: warps {len(w)}, with mesh work {busy.sum()}, nodes/ray {w.mean():.1f}")
print(f"lane utilisation sum/(32*max) over busy warps, weighted by max: {w[busy].sum() / (32 * mx[busy].sum()):.3f}")
print("unweighted mean:", (w[busy].sum(axis=1) / (32 * mx[busy])).mean())
