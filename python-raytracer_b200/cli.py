"""Command-line surface of the reference (`provided/main.py:14-35`, `provided/glue.py:11-27`), same flags:

    python -m python_raytracer_b200.cli main --infile S.json --outfile out.png [--subimage n --tasks N]
    python -m python_raytracer_b200.cli glue --dir temp --outfile out.png
    python -m python_raytracer_b200.cli render INFILE NUM_TASKS OUTFILE        (render.nu:1-23)

`main` without --subimage renders the frame and writes the PNG (rot90 + *255 truncation done on the device);
with --subimage it saves the float64 column strip as `<outfile>.npy` exactly like main.py:26-28, so strips made
here and strips made by the reference can be glued by either glue.  `im.show()` is left out (headless)."""
import argparse
import os
import sys

import numpy as np


def run_main(argv):
    ap = argparse.ArgumentParser(prog="main")
    ap.add_argument("-i", "--infile", type=str, default="scenes/BoxRGBLights.json")      # main.py:15-16
    ap.add_argument("-o", "--outfile", type=str, default="out.png")
    ap.add_argument("-s", "--subimage", type=int)
    ap.add_argument("-t", "--tasks", type=int)
    a = ap.parse_args(argv)
    from . import load_scene
    from PIL import Image
    sc = load_scene(a.infile)
    if a.subimage is not None and a.tasks is not None:                                    # main.py:26-28
        np.save(a.outfile, sc.render(a.subimage, a.tasks))
        return 0
    Image.fromarray(sc.render_u8()).save(a.outfile)                                       # main.py:30-34
    return 0


def glue_strips(directory):
    """glue.py:17-23: N.npy sorted by int(stem), concatenated along axis 0 (columns), rot90."""
    files = sorted((f for f in os.listdir(directory) if f.endswith(".npy")), key=lambda f: int(os.path.splitext(f)[0]))
    strips = [np.load(os.path.join(directory, f)) for f in files]
    image = np.concatenate(strips, axis=0)
    return np.rot90(image, k=1, axes=(0, 1))


def run_glue(argv):
    ap = argparse.ArgumentParser(prog="glue")
    ap.add_argument("-d", "--dir", type=str, default="temp")                              # glue.py:12-13
    ap.add_argument("-o", "--outfile", type=str, default="out.png")
    a = ap.parse_args(argv)
    from PIL import Image
    Image.fromarray((glue_strips(a.dir) * 255).astype(np.uint8)).save(a.outfile)          # glue.py:25-26
    return 0


def run_render(argv):
    """`render.nu infile num_tasks outfile` (render.nu:1-23): NUM_TASKS column strips into ./temp/<n>.npy, glue, remove
    temp.  The reference queues one `main.py --subimage n` process per strip on pueue; here the strips are rendered by
    this process (scene uploaded once), or, when launched under torchrun, strip n by rank n % WORLD_SIZE on that
    rank's GPU, with rank 0 gluing after a barrier."""
    ap = argparse.ArgumentParser(prog="render")
    ap.add_argument("infile", type=str)
    ap.add_argument("num_tasks", type=int)
    ap.add_argument("outfile", type=str)
    ap.add_argument("--temp", type=str, default="temp")
    a = ap.parse_args(argv)
    import shutil
    from . import load_scene
    from PIL import Image
    world, rank = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0"))
    if world > 1:
        import torch
        import torch.distributed as dist
        if torch.cuda.is_available():
            torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
        dist.init_process_group("gloo")          # control plane only: the strips travel through ./temp like the reference's
    os.makedirs(a.temp, exist_ok=True)                                                    # render.nu:5
    sc = load_scene(a.infile)
    for n in range(rank, a.num_tasks, world):                                             # render.nu:9-12
        np.save(os.path.join(a.temp, str(n)), sc.render(n, a.num_tasks))
    if world > 1:
        dist.barrier()
    if rank == 0:
        Image.fromarray((glue_strips(a.temp) * 255).astype(np.uint8)).save(a.outfile)     # render.nu:14, glue.py:25-26
        shutil.rmtree(a.temp)                                                             # render.nu:18
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def main(argv=None):
    argv = list(sys.argv[1:] if argv is None else argv)
    table = {"main": run_main, "glue": run_glue, "render": run_render}
    if not argv or argv[0] not in table:
        print(__doc__)
        return 2
    return table[argv[0]](argv[1:])


if __name__ == "__main__":
    sys.exit(main())
