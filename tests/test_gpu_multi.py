"""Multi-GPU tests of the peer-memory communicator (need >= 2 GPUs on one node; skipped otherwise).

The checks themselves live in tools/multi_gpu_check.py (one process per GPU under torchrun): all-reduce kernels vs NCCL
and vs a fixed-order reduction, the fused all-reduce+optimizer kernel vs NCCL + single-GPU update, parameter broadcast,
and real training steps whose weights must agree bit for bit across ranks.
"""
import json
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs at least 2 GPUs")
def test_peer_comm_two_ranks():
    n = 2
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr", "127.0.0.1",
           "--master-port", "29517", os.path.join(ROOT, "tools", "multi_gpu_check.py")]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=600)
    line = [l for l in r.stdout.splitlines() if l.startswith("MULTI_GPU_RESULT ")]
    assert line, (r.stdout[-3000:], r.stderr[-3000:])
    res = json.loads(line[-1][len("MULTI_GPU_RESULT "):])
    assert res["ok"], res["failed"]
    assert r.returncode == 0
