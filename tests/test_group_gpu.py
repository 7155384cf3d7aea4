"""T4: the fused NVLink averaging kernel against the NCCL + single-GPU path (needs >= 2 GPUs)."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("payload", ["fp32", "bf16", "u8"])
def test_fused_group_step_matches_nccl_path(payload):
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs at least 2 GPUs")
    world = 2 if n < 4 else (4 if n < 8 else 8)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", "29533", os.path.join(ROOT, "tests/multigpu/check_group.py"),
           payload]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    assert "GROUP_OK" in res.stdout, res.stdout[-3000:] + res.stderr[-3000:]
