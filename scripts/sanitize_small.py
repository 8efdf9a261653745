"""Small renders of every kernel variant for compute-sanitizer (memcheck / initcheck): lean scene with a mesh, CSG scene without a
mesh, CSG scene with meshes, several wavefront batches on two streams, the static-scratch entry point, the pinned host path."""
import ctypes as C, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import numpy as np, torch
from conftest import load_case
import python_raytracer_b200 as rt
abi = rt._abi
for case, over in (("MirrorRefraction", {"resolution": [96, 64]}), ("TorusMesh_smooth", {"resolution": [64, 64]}),
                   ("NovelScene1_small", {"resolution": [64, 32]}), ("MeshInNode", {"resolution": [64, 48]})):
    sc = load_case(case, over)
    W, H = sc.vc.width, sc.vc.height
    h = sc.upload()
    ref = sc.render_region(0, W, 0, H).copy()
    job = sc.make_job(0, W, 0, H, max_batch_rays=1500)          # several batches: both work sets, both streams
    host = torch.empty(W * H * 3, dtype=torch.float32).pin_memory()
    abi.check(sc._lib, sc._lib.rt_render_host(h, C.byref(job), host.data_ptr(), None, None), "rt_render_host")
    out = torch.zeros(W * H * 3, dtype=torch.float32, device="cuda")
    abi.check(sc._lib, sc._lib.rt_render_static(h, C.byref(job), out.data_ptr(), None, None), "rt_render_static")
    torch.cuda.synchronize()
    ok = np.array_equal(host.numpy().reshape(W, H, 3), ref) and np.array_equal(out.cpu().numpy().reshape(W, H, 3), ref)
    print(case, "identical" if ok else "DIFFERENT", flush=True)
    sc.close()
print("done")
