#!/usr/bin/env python3
"""Local swarm launcher - the credential-free counterpart of reference manage_scaleset.py (which provisions an Azure
spot scale set whose cloud-init starts ``run_trainer.py``).  Starts, on this machine:

    1 auxiliary peer (the bootstrap / monitor / checkpointer)  +  N trainer peers (one per GPU, or CPU peers)

    python launch_swarm.py create --peers 2 --device cpu --model_preset tiny --target_batch_size 16
    python launch_swarm.py delete            # stop a swarm started with --detach

Peers are plain subprocesses of this repository's entry points; nothing here needs cloud credentials.
"""
import argparse
import json
import os
import re
import signal
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
STATE = os.path.join(os.environ.get("DALLE_B200_SWARM_STATE_DIR", os.path.expanduser("~")), ".dalle_b200_swarm.json")


def _wait_for_maddr(log_path, timeout=120.0):
    t0 = time.time()
    while time.time() - t0 < timeout:
        if os.path.exists(log_path):
            m = re.search(r"--initial_peers (\S+)", open(log_path, errors="ignore").read())
            if m:
                return re.sub(r"\x1b\[[0-9;]*m", "", m.group(1))
        time.sleep(0.2)
    raise TimeoutError(f"aux peer did not announce its address, see {log_path}")


def create(a, extra):
    os.makedirs(a.workdir, exist_ok=True)
    env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""), PYTHONUNBUFFERED="1")
    common = ["--experiment_prefix", a.prefix, "--authorize", "false", "--model_preset", a.model_preset,
              "--target_batch_size", str(a.target_batch_size), "--matchmaking_time", str(a.matchmaking_time),
              "--averaging_timeout", str(a.averaging_timeout), "--allreduce_timeout", str(a.allreduce_timeout)]
    if a.device == "cpu":
        common += ["--no_cuda", "true"]
    procs = []
    aux_log = os.path.join(a.workdir, "aux.log")
    aux_cmd = [sys.executable, os.path.join(ROOT, "run_aux_peer.py"), *common, "--host_maddrs", "/ip4/127.0.0.1/tcp/0",
               "--refresh_period", str(a.refresh_period), "--local_path", os.path.join(a.workdir, "aux_repo"),
               "--output_dir", os.path.join(a.workdir, "aux_out"),
               "--assist_in_averaging", "true" if a.aux_assists else "false"]
    procs.append(subprocess.Popen(aux_cmd, stdout=open(aux_log, "w"), stderr=subprocess.STDOUT, env=env, cwd=a.workdir))
    maddr = _wait_for_maddr(aux_log)
    R = max(1, a.ranks_per_peer)
    for i in range(a.peers):
        penv = dict(env)
        if a.device == "cuda":
            first = (i * R) % max(1, a.gpus)
            penv["CUDA_VISIBLE_DEVICES"] = ",".join(str((first + r) % max(1, a.gpus)) for r in range(R))
        # a peer of several ranks is one torchrun launch: the ranks form ONE swarm peer (NVLink-fused averaging inside)
        launcher = [sys.executable] if R == 1 else [
            sys.executable, "-m", "torch.distributed.run", "--nnodes", "1", "--nproc-per-node", str(R),
            "--master-addr", "127.0.0.1", "--master-port", str(a.base_port + i)]
        cmd = [*launcher, os.path.join(ROOT, "run_trainer.py"), *common, "--initial_peers", maddr,
               "--host_maddrs", "/ip4/127.0.0.1/tcp/0", "--synthetic_data", "true",
               "--per_device_train_batch_size", str(a.batch), "--gradient_accumulation_steps", "1",
               "--output_dir", os.path.join(a.workdir, f"peer{i}_out"),
               "--state_path", os.path.join(a.workdir, f"peer{i}_state.zip"), "--logging_steps", "10",
               "--max_steps", str(a.max_steps), *extra]
        procs.append(subprocess.Popen(cmd, stdout=open(os.path.join(a.workdir, f"peer{i}.log"), "w"),
                                      stderr=subprocess.STDOUT, env=penv, cwd=a.workdir))
    json.dump({"pids": [p.pid for p in procs], "workdir": a.workdir}, open(STATE, "w"))
    print(f"swarm '{a.prefix}': aux + {a.peers} trainer peer(s); logs in {a.workdir}; bootstrap {maddr}")
    if a.detach:
        return 0
    try:
        for p in procs[1:]:
            p.wait()
    except KeyboardInterrupt:
        pass
    finally:
        time.sleep(a.linger)
        for p in procs:
            if p.poll() is None:
                p.send_signal(signal.SIGTERM)
    return 0


def delete(a, extra):
    if not os.path.exists(STATE):
        print("no swarm recorded")
        return 0
    for pid in json.load(open(STATE))["pids"]:
        try:
            os.kill(pid, signal.SIGTERM)  # exact PIDs this launcher started
        except ProcessLookupError:
            pass
    os.remove(STATE)
    return 0


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("action", choices=["create", "delete"])
    ap.add_argument("--peers", type=int, default=4, help="trainer peers (reference SWARM_SIZE = 4)")
    ap.add_argument("--device", choices=["cpu", "cuda"], default="cuda")
    ap.add_argument("--gpus", type=int, default=8)
    ap.add_argument("--ranks_per_peer", type=int, default=1,
                    help="GPUs (processes) per trainer peer; > 1 launches each peer under torchrun as one group")
    ap.add_argument("--base_port", type=int, default=29700, help="torchrun rendezvous port of peer i = base_port + i")
    ap.add_argument("--prefix", default="my-model")
    ap.add_argument("--model_preset", default="ttt")
    ap.add_argument("--batch", type=int, default=1, help="per-device micro-batch (reference cloud workers use 1)")
    ap.add_argument("--target_batch_size", type=int, default=4096)
    ap.add_argument("--matchmaking_time", type=float, default=15.0)
    ap.add_argument("--averaging_timeout", type=float, default=180.0)
    ap.add_argument("--allreduce_timeout", type=float, default=60.0)
    ap.add_argument("--refresh_period", type=float, default=10.0)
    ap.add_argument("--max_steps", type=int, default=10 ** 20)
    ap.add_argument("--workdir", default=os.path.join(ROOT, "swarm_run"))
    ap.add_argument("--aux_assists", action="store_true")
    ap.add_argument("--detach", action="store_true")
    ap.add_argument("--linger", type=float, default=0.0, help="keep the aux peer alive this long after trainers exit")
    a, extra = ap.parse_known_args(argv)
    return {"create": create, "delete": delete}[a.action](a, extra)


if __name__ == "__main__":
    sys.exit(main())
