"""Build csrc/libb200rt.so for sm_100a with nvcc (cross-compiles without a GPU)."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "libb200rt.so")
SRCS = ["b200rt.cu"]
DEPS = sorted(f for f in os.listdir(HERE) if f.endswith((".cu", ".cuh", ".hpp", ".h"))) + ["../../include/b200rt.h"]      # every source the library includes
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17", "-shared",
         "-Xcompiler", "-fPIC", "-Xptxas", "-v", "--expt-relaxed-constexpr"]


def build(force=False, verbose=False):
    global OUT
    extra = os.environ.get("B200RT_EXTRA_NVCC", "").split()       # developer knob for tuning experiments (-DRT_MIN_BLOCKS=...)
    if os.environ.get("B200RT_OUT"):
        OUT = os.path.join(HERE, os.environ["B200RT_OUT"]); force = True
    newest = max(os.path.getmtime(os.path.join(HERE, d)) for d in DEPS)
    if not force and os.path.exists(OUT) and os.path.getmtime(OUT) >= newest:
        return OUT
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    cmd = [nvcc] + FLAGS + extra + ["-o", OUT] + [os.path.join(HERE, s) for s in SRCS]
    r = subprocess.run(cmd, capture_output=True, text=True)
    with open(os.path.join(HERE, "build.log"), "w") as fh:
        fh.write(" ".join(cmd) + "\n" + r.stdout + r.stderr)
    if verbose or r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed, see csrc/build.log")
    return OUT


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose=True)
