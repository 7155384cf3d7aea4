"""T4: the fused NVLink averaging kernel against the NCCL + single-GPU path (needs >= 2 GPUs)."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_fused_group_kernel_runs_as_a_self_peer():
    """world = 1: the fused reduce-scatter / LAMB / all-gather kernel on a single GPU (its peers are itself), every
    payload codec, against the single-GPU LAMB kernels - so a one-GPU box executes group_lamb_kernel at all."""
    env = dict(os.environ, WORLD_SIZE="1")
    for payload in ("fp32", "adaptive"):
        res = subprocess.run([sys.executable, os.path.join(ROOT, "tests/multigpu/check_group.py"), payload],
                             capture_output=True, text=True, timeout=300, env=env)
        assert "GROUP_OK" in res.stdout, res.stdout[-3000:] + res.stderr[-3000:]


def test_hivemind_codecs_of_the_fused_kernel_match_the_oracle():
    """reduce_only at world = 1 returns decode(encode(g)): the kernel's fp16 / uniform-8-bit codecs against
    collab/compression.py (chunk by chunk, 131072 elements, Appendix D.2)."""
    res = subprocess.run([sys.executable, os.path.join(ROOT, "tests/multigpu/check_codecs.py")], capture_output=True,
                         text=True, timeout=300, env=dict(os.environ, WORLD_SIZE="1"))
    assert "CODECS_OK" in res.stdout, res.stdout[-3000:] + res.stderr[-3000:]


@pytest.mark.parametrize("payload,layout", [("fp32", "small"), ("bf16", "small"), ("u8", "small"), ("fp16", "small"),
                                            ("u8h", "small"), ("adaptive", "small"), ("fp32", "ttt"),
                                            ("adaptive", "ttt")])
def test_fused_group_step_matches_nccl_path(payload, layout):
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs at least 2 GPUs")
    world = 2 if n < 4 else (4 if n < 8 else 8)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", "29533", os.path.join(ROOT, "tests/multigpu/check_group.py"),
           payload, layout]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    assert "GROUP_OK" in res.stdout, res.stdout[-3000:] + res.stderr[-3000:]
