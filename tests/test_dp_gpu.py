"""GPU, needs >= 2 devices (skipped on a one-GPU box): the data-parallel paths of PretrainStep on real NVLink —
the fused peer-memory optimizer (k3m_adamw_dp) against NCCL bf16 / fp32 all-reduce, via tools/check_dp.py under
torchrun."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_fused_peer_memory_optimizer_matches_nccl_paths():
    n = min(torch.cuda.device_count(), 8)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(n), "--master-addr", "127.0.0.1",
           "--master-port", str(29600 + os.getpid() % 300), os.path.join(ROOT, "tools", "check_dp.py")]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    print(res.stdout[-3000:])
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-2000:]
    assert "ALL OK" in res.stdout
