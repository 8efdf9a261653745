"""SURVEY 8e under pytest: the N-rank tile-sharded render (one process per GPU, torchrun) equals the single-rank render bit for
bit, for both fan-ins (NVLink peer stores into rank 0's frame; NCCL gather + scatter kernel).  Needs >= 2 visible GPUs: skipped
on a one-GPU box (there the emulated-rank test of tests/test_gpu_parity.py and the gloo test of tests/test_distributed_cpu.py
cover the tile logic); `gpurun --gpus N -- python -m pytest tests/test_multigpu.py -m gpu` runs it for real."""
import json
import os
import socket
import subprocess
import sys

import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu


def _n_gpus():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


@pytest.mark.skipif(_n_gpus() < 2, reason="needs at least two GPUs on one box")
def test_n_rank_render_is_bit_identical_to_single_rank():
    n = min(8, _n_gpus())
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(ROOT, "tests", "run_multigpu_check.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    line = [l for l in r.stdout.splitlines() if l.startswith("{")][-1]
    out = json.loads(line)
    assert out["world"] == n and len(out["results"]) == 6
    assert all(v["bit_identical"] for v in out["results"].values()), out
