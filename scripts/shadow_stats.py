"""Where the level-0 shadow rays of the bench scene spend their BVH work (host twin of the device functions, tests/hostsim):
per hit-object kind x light x occluded: rays, mesh nodes, triangle tests, object-BVH node + primitive tests per ray."""
import ctypes as C, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "tests", "hostsim")]
import numpy as np
import hostsim
from python_raytracer_b200 import synthetic
spp = int(os.environ.get("SPP", "4"))
sc = synthetic.make_scene(aa_samples=spp)
flat = sc.flatten()
rng = np.random.default_rng(3)
npx = int(os.environ.get("NPX", "20000"))
cols = rng.integers(0, sc.vc.width, npx).astype(np.int32); rows = rng.integers(0, sc.vc.height, npx).astype(np.int32)
L = len(sc.lights)
out = np.zeros((npx, spp, L, 5), np.int32)
lib = hostsim.lib()
lib.hostsim_shadow_stats.argtypes = [C.c_void_p] * 3 + [C.c_int64, C.c_void_p]
assert lib.hostsim_shadow_stats(C.addressof(flat.desc), cols.ctypes.data, rows.ctypes.data, npx, out.ctypes.data) == 0
o = out.reshape(-1, L, 5)
kinds = {-1: "miss", 0: "sphere", 1: "plane", 3: "mesh"}
tot = o[..., 2:].sum(axis=(0, 1))
print("total per ray: nodes %.2f tris %.2f prims %.2f" % tuple(tot / (o[..., 0] >= 0).sum()))
for k, name in kinds.items():
    for li in range(L):
        for occ in (0, 1):
            m = (o[:, li, 0] == k) & (o[:, li, 1] == occ)
            if m.sum() == 0 or k < 0: continue
            r = o[m, li]
            print(f"{name:7s} light {li} occluded {occ}: rays {m.sum():8d} ({100 * m.sum() / (o[..., 0] >= 0).sum():5.1f} %)  nodes/ray {r[:, 2].mean():6.2f}  tris/ray {r[:, 3].mean():5.2f}  prims/ray {r[:, 4].mean():5.2f}"
                  f"  share of nodes {100 * r[:, 2].sum() / tot[0]:5.1f} %  of prims {100 * r[:, 4].sum() / tot[2]:5.1f} %  zero-node rays {100 * (r[:, 2] == 0).mean():5.1f} %")
