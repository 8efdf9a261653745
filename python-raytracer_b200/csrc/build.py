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


TORCH_OUT = os.path.join(HERE, "libb200rt_torch.so")


def build_torch_ops(force=False, verbose=False):
    """The thin PyTorch extension (torch_ops.cpp -> torch.ops.b200rt.*): plain g++ against the torch headers, linked to
    libb200rt.so in the same directory (rpath $ORIGIN).  In-tree, so it travels to the GPU box with the snapshot."""
    src = os.path.join(HERE, "torch_ops.cpp")
    hdr = os.path.join(HERE, "..", "..", "include", "b200rt.h")
    if not force and os.path.exists(TORCH_OUT) and os.path.getmtime(TORCH_OUT) >= max(os.path.getmtime(src), os.path.getmtime(hdr)):
        return TORCH_OUT
    import torch
    from torch.utils import cpp_extension as ce
    inc = [f"-I{p}" for p in ce.include_paths()] + ["-I/usr/local/cuda/include"]
    libdir = os.path.join(os.path.dirname(torch.__file__), "lib")
    cmd = ["g++", "-O2", "-std=c++17", "-shared", "-fPIC", f"-D_GLIBCXX_USE_CXX11_ABI={int(torch._C._GLIBCXX_USE_CXX11_ABI)}",
           *inc, "-o", TORCH_OUT, src, f"-L{HERE}", "-lb200rt", f"-L{libdir}", "-ltorch", "-ltorch_cpu", "-lc10", "-lc10_cuda",
           "-L/usr/local/cuda/lib64", "-lcudart", "-Wl,-rpath,$ORIGIN", f"-Wl,-rpath,{libdir}"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if verbose or r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
    if r.returncode != 0:
        raise RuntimeError("building the torch extension failed")
    return TORCH_OUT


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose=True)
    build_torch_ops(force="--force" in sys.argv, verbose=True)
