"""A multi-process swarm peer on CPU (gloo): the ranks of one torchrun launch form ONE peer.  Two such
groups of two ranks + an aux peer train the tiny model over the TCP DHT; afterwards the group's backup is checked to
be the complete single-peer layout (Appendix A.3 / A.2)."""
import glob
import os
import re
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.timeout(900)
def test_two_groups_of_two_ranks_and_aux_on_cpu(tmp_path):
    work = str(tmp_path / "swarm")
    env = dict(os.environ, DALLE_B200_SWARM_STATE_DIR=str(tmp_path), DALLE_B200_TARGET_GROUP_SIZE="2")
    cmd = [sys.executable, os.path.join(ROOT, "launch_swarm.py"), "create", "--peers", "2", "--ranks_per_peer", "2",
           "--device", "cpu", "--model_preset", "tiny", "--batch", "1", "--target_batch_size", "8",
           "--matchmaking_time", "2", "--averaging_timeout", "20", "--allreduce_timeout", "10", "--refresh_period", "1",
           "--max_steps", "14", "--workdir", work, "--prefix", "grp", "--linger", "4", "--backup_every_steps", "1",
           "--base_port", str(29700 + os.getpid() % 200)]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=840, cwd=str(tmp_path), env=env)
    logs = {os.path.basename(p): open(p, errors="ignore").read() for p in glob.glob(os.path.join(work, "*.log"))}
    assert res.returncode == 0, (res.stdout, res.stderr, logs)
    for name in ("peer0.log", "peer1.log"):
        assert "Traceback" not in logs[name], logs[name][-4000:]
        epochs = [int(x) for x in re.findall(r"Current epoch: (\d+)", logs[name])]
        assert epochs and max(epochs) >= 2, logs[name][-4000:]
        # both ranks of the group end with identical parameters
        sums = set(re.findall(r"GROUP_PARAM_CHECKSUM (\S+)", logs[name]))
        assert len(sums) == 1, (sums, logs[name][-2000:])
    assert "swarm-averaged" in logs["peer0.log"] + logs["peer1.log"]
    assert re.search(r"Epoch #\d+\s+loss = \d", logs["aux.log"]), logs["aux.log"][-3000:]
    # the group's rolling backup is the complete reference layout and loads into a single-process optimizer
    state = torch.load(os.path.join(work, "peer0_state.zip"), weights_only=False)
    assert set(state) == {"model", "training", "scheduler", "local_epoch"}
    assert "model.to_logits.1.weight" in state["model"] and "local_epoch" in state["training"]["state"]
    n_params = sum(len(g["params"]) for g in state["training"]["param_groups"])
    assert len([k for k in state["training"]["state"] if k != "local_epoch"]) == n_params

    # (1) the checkpoint a GROUP wrote drives the public inference entry point
    sys.path.insert(0, os.path.join(ROOT, "inference"))
    import pickle

    import run_inference

    ckpt = os.path.join(work, "group_model.pt")
    torch.save(state["model"], ckpt)
    q = tmp_path / "queries.txt"
    q.write_text("a cat\n")
    run_inference.main(["--queries", str(q), "--temperature", "1.0", "--top-k", "4", "--model", ckpt, "--output-dir",
                        str(tmp_path / "inf"), "--model-preset", "tiny", "--batch-size", "1", "--n-iters", "1"])
    rec = pickle.load(open(tmp_path / "inf" / "a cat.pickle", "rb"))
    assert rec["codes"][0].shape == (256,)

    # (2) the group is restarted from its state.zip (alone, no swarm): it resumes at the saved epoch, not at 0
    saved_epoch = int(state["local_epoch"])
    assert saved_epoch >= 1
    log2 = os.path.join(work, "restart.log")
    cmd2 = _trainer_cmd(work, "peer0", "", "grp_restart", 2, 29950 + os.getpid() % 40, 4)
    i = cmd2.index("--initial_peers")
    del cmd2[i:i + 2]
    p2 = _spawn(cmd2, log2, dict(env, PYTHONPATH=ROOT + os.pathsep + env.get("PYTHONPATH", ""), PYTHONUNBUFFERED="1"), work)
    assert p2.wait(timeout=400) == 0, open(log2, errors="ignore").read()[-4000:]
    txt = open(log2, errors="ignore").read()
    assert "Loaded state" in txt and "Traceback" not in txt, txt[-3000:]
    epochs = [int(x) for x in re.findall(r"Current epoch: (\d+)", txt)]
    assert epochs and min(epochs) >= saved_epoch, (saved_epoch, epochs, txt[-3000:])
    assert len(set(re.findall(r"GROUP_PARAM_CHECKSUM (\S+)", txt))) == 1


def _spawn(cmd, log, env, cwd):
    return subprocess.Popen(cmd, stdout=open(log, "w"), stderr=subprocess.STDOUT, env=env, cwd=cwd)


def _trainer_cmd(work, name, maddr, prefix, ranks, port, steps, extra=()):
    launcher = [sys.executable] if ranks == 1 else [
        sys.executable, "-m", "torch.distributed.run", "--nnodes", "1", "--nproc-per-node", str(ranks),
        "--master-addr", "127.0.0.1", "--master-port", str(port)]
    return [*launcher, os.path.join(ROOT, "run_trainer.py"), "--experiment_prefix", prefix, "--authorize", "false",
            "--model_preset", "tiny", "--no_cuda", "true", "--target_batch_size", "8", "--matchmaking_time", "2",
            "--averaging_timeout", "20", "--allreduce_timeout", "10", "--initial_peers", maddr,
            "--host_maddrs", "/ip4/127.0.0.1/tcp/0", "--synthetic_data", "true", "--per_device_train_batch_size", "1",
            "--gradient_accumulation_steps", "1", "--output_dir", os.path.join(work, f"{name}_out"),
            "--state_path", os.path.join(work, f"{name}_state.zip"), "--logging_steps", "10",
            "--backup_every_steps", "1", "--max_steps", str(steps), *extra]


@pytest.mark.timeout(1200)
def test_group_nan_rollback_mixed_swarm_and_late_joiner(tmp_path):
    """One swarm with a two-rank group peer and a single-process peer (same averager schema: they average together);
    a NaN is injected into ONE rank of the group, which must roll back as a whole; a second group joins late and
    catches up by downloading the state."""
    import time

    sys.path.insert(0, ROOT)
    from launch_swarm import _wait_for_maddr

    work = str(tmp_path / "swarm")
    os.makedirs(work)
    env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""), PYTHONUNBUFFERED="1")
    port = 29900 + os.getpid() % 90
    procs = []
    try:
        aux_log = os.path.join(work, "aux.log")
        procs.append(_spawn([sys.executable, os.path.join(ROOT, "run_aux_peer.py"), "--experiment_prefix", "mix",
                             "--authorize", "false", "--model_preset", "tiny", "--no_cuda", "true",
                             "--target_batch_size", "8", "--matchmaking_time", "2", "--averaging_timeout", "20",
                             "--allreduce_timeout", "10", "--host_maddrs", "/ip4/127.0.0.1/tcp/0", "--refresh_period", "1",
                             "--local_path", os.path.join(work, "aux_repo"), "--output_dir", os.path.join(work, "aux_out")],
                            aux_log, env, work))
        maddr = _wait_for_maddr(aux_log)
        genv = dict(env, DALLE_B200_INJECT_NAN="1:9")     # rank 1 of the group, after its 9th micro-step
        procs.append(_spawn(_trainer_cmd(work, "group", maddr, "mix", 2, port, 36), os.path.join(work, "group.log"),
                            genv, work))
        procs.append(_spawn(_trainer_cmd(work, "single", maddr, "mix", 1, 0, 80), os.path.join(work, "single.log"),
                            env, work))
        # the late joiner starts once the swarm is a few epochs in
        t0 = time.time()
        while time.time() - t0 < 300:
            txt = open(os.path.join(work, "group.log"), errors="ignore").read()
            if len(re.findall(r"Current epoch: [3-9]", txt)) > 0:
                break
            time.sleep(1.0)
        procs.append(_spawn(_trainer_cmd(work, "late", maddr, "mix", 2, port + 1, 12), os.path.join(work, "late.log"),
                            env, work))
        for p in procs[1:]:
            p.wait(timeout=600)
    finally:
        for p in procs:
            if p.poll() is None:
                p.terminate()
    logs = {os.path.basename(p): open(p, errors="ignore").read() for p in glob.glob(os.path.join(work, "*.log"))}
    for name in ("group.log", "single.log", "late.log"):
        assert "Traceback" not in logs[name], logs[name][-4000:]
    g = logs["group.log"]
    assert "failpoint: injected NaN" in g
    assert len(re.findall(r"Parameters are invalid, reloading model from earlier state", g)) == 2, g[-4000:]  # both ranks
    assert len(set(re.findall(r"GROUP_PARAM_CHECKSUM (\S+)", g))) == 1, g[-2000:]
    assert "swarm-averaged" in g, g[-4000:]                      # the group and the single peer did average together
    late = logs["late.log"]
    first = int(re.findall(r"Current epoch: (\d+)", late)[0])
    assert first >= 3, late[-4000:]                             # started from the downloaded state, not from epoch 0
    assert len(set(re.findall(r"GROUP_PARAM_CHECKSUM (\S+)", late))) == 1
