"""A multi-GPU swarm peer on real GPUs: run_trainer.py under torchrun (2 ranks = ONE peer, fused NVLink averaging and
sharded LAMB-8bit inside, leader on the TCP swarm) + an aux peer; the group's backup must be the complete single-peer
layout and load into the single-process inference path."""
import glob
import os
import re
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.timeout(900)
def test_two_gpu_group_peer_trains_and_checkpoints(tmp_path):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs at least 2 GPUs")
    work = str(tmp_path / "swarm")
    env = dict(os.environ, DALLE_B200_SWARM_STATE_DIR=str(tmp_path))
    cmd = [sys.executable, os.path.join(ROOT, "launch_swarm.py"), "create", "--peers", "1", "--ranks_per_peer", "2",
           "--device", "cuda", "--gpus", "2", "--model_preset", "small", "--batch", "2", "--target_batch_size", "16",
           "--matchmaking_time", "2", "--averaging_timeout", "20", "--allreduce_timeout", "10", "--refresh_period", "1",
           "--max_steps", "24", "--workdir", work, "--prefix", "gpugrp", "--linger", "4", "--backup_every_steps", "1",
           "--base_port", str(29800 + os.getpid() % 100)]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=840, cwd=str(tmp_path), env=env)
    logs = {os.path.basename(p): open(p, errors="ignore").read() for p in glob.glob(os.path.join(work, "*.log"))}
    assert res.returncode == 0, (res.stdout, res.stderr, logs)
    g = logs["peer0.log"]
    assert "Traceback" not in g, g[-4000:]
    epochs = [int(x) for x in re.findall(r"Current epoch: (\d+)", g)]
    assert epochs and max(epochs) >= 3, g[-4000:]
    assert len(set(re.findall(r"GROUP_PARAM_CHECKSUM (\S+)", g))) == 1, g[-2000:]
    # loss goes down over the run (same synthetic stream, lr warm-up is short in the small preset run)
    state = torch.load(os.path.join(work, "peer0_state.zip"), weights_only=False)
    assert set(state) == {"model", "training", "scheduler", "local_epoch"}
    n_params = sum(len(grp["params"]) for grp in state["training"]["param_groups"])
    st = {k: v for k, v in state["training"]["state"].items() if k != "local_epoch"}
    assert len(st) == n_params
    # the gathered optimizer state is complete: every quantised tensor has non-zero absmax in all of its blocks
    big = [v for v in st.values() if "absmax1" in v]
    assert big and all(bool((v["absmax2"] > 0).all()) for v in big), "optimizer state of some shard is missing"
    # the checkpoint loads into a plain single-process model
    from dalle_b200.config import get_config
    from dalle_b200.models.dalle import DALLE, ModelWrapper

    m = ModelWrapper(DALLE(get_config("small")))
    m.load_state_dict(state["model"])
