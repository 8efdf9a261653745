"""Tier-0 oracle: run the UNMODIFIED reference renderer through the glm/igl shims.

TEST INFRASTRUCTURE ONLY, and only usable in the dev container: it imports
`/root/reference/provided/*.py` from where they lie (nothing is copied) with
`oracle/shims` ahead of it on `sys.path`.  `/root/reference` does not exist on the GPU
box, so nothing under `tests/`, `bench.py` or `smoke()` calls this at run time; it is used
by `oracle/make_golden.py` to produce the committed fixtures in `tests/golden/`.

What is patched, and why (no reference file is edited):
  * `numpy.random.rand`  -> `oracle.philox.RandStream`, so jitter (`scene.py:63-65`) reads
    the counter-based table the CUDA kernels use.
  * `scene.tqdm`         -> a no-op progress object (one `update` per sample otherwise, `scene.py:71`).
  * cwd                  -> `/root/reference` so `./meshes/...` and `textures/...` resolve.
Strip fan-out stands in for `render.nu:10-13` (nushell/pueue are absent): one process per
strip over `Scene.render(subimage, tasks)`, concatenated on axis 0 as `glue.py:22` does.
"""
import contextlib
import io
import json
import multiprocessing as mp
import os
import sys
import tempfile
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REFERENCE = os.environ.get("RT_REFERENCE_ROOT", "/root/reference")


class _NoProgress:
    def __init__(self, *a, **k):
        pass

    def update(self, n=1):
        pass

    def close(self):
        pass


def _import_reference():
    shims = os.path.join(HERE, "shims")
    prov = os.path.join(REFERENCE, "provided")
    for p in (prov, shims):
        if p in sys.path:
            sys.path.remove(p)
        sys.path.insert(0, p)
    if HERE not in sys.path:
        sys.path.append(HERE)
    os.chdir(REFERENCE)
    import scene_parser  # noqa: the reference's own parser
    import scene
    scene.tqdm = _NoProgress
    return scene_parser, scene


def apply_overrides(data, overrides):
    """`overrides` is a dict merged into the scene JSON (top-level keys replaced; dict values merged)."""
    for k, v in (overrides or {}).items():
        if isinstance(v, dict) and isinstance(data.get(k), dict):
            data[k] = {**data[k], **v}
        else:
            data[k] = v
    return data


def load_reference_scene(scene_path, overrides=None, quiet=True):
    scene_parser, _ = _import_reference()
    path = scene_path if os.path.isabs(scene_path) else os.path.join(REFERENCE, scene_path)
    if overrides:
        with open(path) as fh:
            data = apply_overrides(json.load(fh), overrides)
        tmp = tempfile.NamedTemporaryFile("w", suffix=".json", delete=False)
        json.dump(data, tmp)
        tmp.close()
        path = tmp.name
    sink = io.StringIO()
    with contextlib.redirect_stdout(sink) if quiet else contextlib.nullcontext():
        sc = scene_parser.load_scene(path)
    if overrides:
        os.unlink(path)
    return sc


def _strip_job(args):
    scene_path, overrides, subimage, tasks, seed = args
    sc = load_reference_scene(scene_path, overrides)
    import philox
    W, H = sc.vc.width, sc.vc.height
    cols = np.array_split(np.arange(W), tasks)[subimage]
    if len(cols) == 0:
        return np.zeros((0, H, 3)), 0.0
    first = int(cols[0]) * H * sc.vc.dof_samples * sc.samples
    np.random.rand = philox.RandStream(first, seed)
    t0 = time.perf_counter()
    sink = io.StringIO()
    with contextlib.redirect_stdout(sink):  # Sphere.intersect prints on near-zero directions
        img = sc.render(subimage, tasks)
    return img, time.perf_counter() - t0


def render_reference(scene_path, overrides=None, tasks=None, procs=None, seed=0x5EED, strips=None):
    """Render with the reference. Returns (image (W_sel,H,3) float64, cpu_seconds_summed).

    tasks : number of column strips (default: one per core, capped at W)
    strips: optional subset of strip indices to render (for sampled timing/parity);
            the returned image is their concatenation in index order.
    """
    procs = procs or os.cpu_count()
    if tasks is None:
        tasks = procs * 4
    jobs = [(scene_path, overrides, n, tasks, seed) for n in (range(tasks) if strips is None else strips)]
    ctx = mp.get_context("fork")
    with ctx.Pool(min(procs, len(jobs))) as pool:
        parts = pool.map(_strip_job, jobs, chunksize=1)
    img = np.concatenate([p[0] for p in parts], axis=0)
    return img, sum(p[1] for p in parts)


def to_png_array(image):
    """`main.py:31-33`: rot90 then *255 with truncation."""
    return (np.rot90(image, k=1, axes=(0, 1)) * 255).astype(np.uint8)


if __name__ == "__main__":
    import argparse
    ap = argparse.ArgumentParser()
    ap.add_argument("scene")
    ap.add_argument("--out", default=None)
    ap.add_argument("--tasks", type=int, default=None)
    a = ap.parse_args()
    out = os.path.abspath(a.out) if a.out else None
    t0 = time.time()
    img, cpu = render_reference(a.scene, tasks=a.tasks)
    print(f"rendered {img.shape} in {time.time()-t0:.1f}s wall, {cpu:.1f}s cpu")
    if out:
        from PIL import Image
        Image.fromarray(to_png_array(img)).save(out)
