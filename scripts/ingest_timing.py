import sys, time, numpy as np, torch
sys.path.insert(0, '/root/repo')
import python_raytracer_b200 as rt
from python_raytracer_b200 import geometry, synthetic, _abi
torch.cuda.init(); torch.zeros(1).cuda()
vf = synthetic.displaced_torus(1024, 512, 1234)
for rep in range(3):
    for how in ("device", "host"):
        t0 = time.perf_counter()
        geometry.Mesh("m", "mesh", [], (0.0, 0.0, 0.0), 1.0, vf, False, None, ingest=how)
        print(rep, how, round(1e3 * (time.perf_counter() - t0), 1), "ms")
