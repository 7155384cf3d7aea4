"""T2 end to end (BASELINE config 1): tiny DALL-E, 2 collaborative trainer peers + 1 aux peer on CPU over the TCP
DHT, synthetic VQGAN codes; epochs advance, the aux peer sees signed metrics and writes a checkpoint."""
import glob
import os
import re
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.timeout(600)
def test_two_peers_and_aux_on_cpu(tmp_path):
    work = str(tmp_path / "swarm")
    cmd = [sys.executable, os.path.join(ROOT, "launch_swarm.py"), "create", "--peers", "2", "--device", "cpu",
           "--model_preset", "tiny", "--batch", "2", "--target_batch_size", "8", "--matchmaking_time", "2",
           "--averaging_timeout", "20", "--allreduce_timeout", "10", "--refresh_period", "1", "--max_steps", "14",
           "--workdir", work, "--prefix", "e2e", "--linger", "4", "--backup_every_steps", "1"]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=540, cwd=str(tmp_path))
    logs = {os.path.basename(p): open(p, errors="ignore").read() for p in glob.glob(os.path.join(work, "*.log"))}
    assert res.returncode == 0, (res.stdout, res.stderr, logs)
    for name in ("peer0.log", "peer1.log"):
        epochs = [int(x) for x in re.findall(r"Current epoch: (\d+)", logs[name])]
        assert epochs and max(epochs) >= 2, logs[name][-3000:]
        assert "Traceback" not in logs[name], logs[name][-3000:]
    assert re.search(r"Epoch #\d+\s+loss = \d", logs["aux.log"]), logs["aux.log"][-3000:]
    # the local rolling backup has the reference layout (Appendix A.3)
    state = torch.load(os.path.join(work, "peer0_state.zip"), weights_only=False)
    assert set(state) == {"model", "training", "scheduler", "local_epoch"}
    assert "model.to_logits.1.weight" in state["model"] and "local_epoch" in state["training"]["state"]
