"""Fixtures for the statistical cross-check against the AUTHOR's own jittered renders (the stored PNGs of the
scenes that use unseeded jitter cannot be reproduced bit for bit; their block means can).  Reads /root/reference/renders/<name>.png
(dev container only) and writes tests/golden/author_stats.npz: per scene the mean uint8 value of every BLOCK x BLOCK pixel block,
(H/B, W/B, 3) float32.  TEST INFRASTRUCTURE.  usage: python oracle/make_author_stats.py"""
import os
import numpy as np
from PIL import Image

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BLOCK = 8
SCENES = ["DepthOfField", "NovelScene1", "NovelScene2"]


def block_means(img_u8, b=BLOCK):
    """(H, W, 3) uint8 -> (H/b, W/b, 3) float32 block means (H, W multiples of b)."""
    a = np.asarray(img_u8, dtype=np.float32)
    H, W, _ = a.shape
    return a.reshape(H // b, b, W // b, b, 3).mean(axis=(1, 3))


if __name__ == "__main__":
    out = {}
    for name in SCENES:
        im = np.array(Image.open(f"/root/reference/renders/{name}.png").convert("RGB"))
        out[name] = block_means(im)
        print(name, im.shape, "->", out[name].shape)
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "author_stats.npz"), **out)
