"""Command-line surface of the reference (`provided/main.py:14-35`, `provided/glue.py:11-27`), same flags:

    python -m python_raytracer_b200.cli main --infile S.json --outfile out.png [--subimage n --tasks N]
    python -m python_raytracer_b200.cli glue --dir temp --outfile out.png

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


def main(argv=None):
    argv = list(sys.argv[1:] if argv is None else argv)
    if not argv or argv[0] not in ("main", "glue"):
        print(__doc__)
        return 2
    return run_main(argv[1:]) if argv[0] == "main" else run_glue(argv[1:])


if __name__ == "__main__":
    sys.exit(main())
