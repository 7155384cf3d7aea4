"""BASELINE config 5 on one box: two collaborative peers, each a group of G GPUs (NVLink-fused averaging inside),
plus run_aux_peer.py, talking over the TCP DHT.  Usage: python tools/run_config5.py [gpus_per_group] [preset]."""
import os, re, subprocess, sys, time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G = int(sys.argv[1]) if len(sys.argv) > 1 else 2
preset = sys.argv[2] if len(sys.argv) > 2 else "small"
work = os.path.join(ROOT, "gpurun_out", "config5")
os.makedirs(work, exist_ok=True)
env = dict(os.environ, PYTHONPATH=ROOT, PYTHONUNBUFFERED="1")
common = ["--experiment_prefix", "cfg5", "--authorize", "false", "--model_preset", preset, "--target_batch_size", "64",
          "--matchmaking_time", "20", "--averaging_timeout", "60", "--allreduce_timeout", "30"]
aux_log = os.path.join(work, "aux.log")
aux = subprocess.Popen([sys.executable, os.path.join(ROOT, "run_aux_peer.py"), *common, "--no_cuda", "true",
                        "--host_maddrs", "/ip4/127.0.0.1/tcp/0", "--refresh_period", "1", "--store_checkpoints", "false",
                        "--output_dir", os.path.join(work, "aux_out")],
                       stdout=open(aux_log, "w"), stderr=subprocess.STDOUT, env=dict(env, CUDA_VISIBLE_DEVICES=""), cwd=work)
maddr = None
for _ in range(600):
    m = re.search(r"--initial_peers (\S+)", open(aux_log, errors="ignore").read()) if os.path.exists(aux_log) else None
    if m:
        maddr = re.sub(r"\x1b\[[0-9;]*m", "", m.group(1)); break
    time.sleep(0.2)
assert maddr, open(aux_log).read()[-2000:]
groups = []
for gi in range(2):
    devs = ",".join(str(gi * G + i) for i in range(G))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={G}", "--master-addr", "127.0.0.1",
           "--master-port", str(29600 + gi), os.path.join(ROOT, "run_trainer.py"), *common, "--initial_peers", maddr,
           "--host_maddrs", "/ip4/127.0.0.1/tcp/0", "--per_device_train_batch_size", "4", "--max_steps", "24",
           "--seed", "0", "--learning_rate", "0.01", "--warmup_steps", "1"]
    groups.append(subprocess.Popen(cmd, stdout=open(os.path.join(work, f"group{gi}.log"), "w"), stderr=subprocess.STDOUT,
                                   env=dict(env, CUDA_VISIBLE_DEVICES=devs, DALLE_B200_TARGET_GROUP_SIZE="2", DALLE_B200_DATA_SEED=str(gi + 1)), cwd=work))
rcs = [g.wait(timeout=500) for g in groups]
time.sleep(3)
aux.terminate()
logs = [open(os.path.join(work, f"group{gi}.log"), errors="ignore").read() for gi in range(2)]
auxl = open(aux_log, errors="ignore").read()
done = []
for l in logs:   # run_trainer.py logs "GROUP_PARAM_CHECKSUM <sum> epoch <n>" on every rank at shutdown
    m = re.findall(r"GROUP_PARAM_CHECKSUM (\S+) epoch (\d+)", l)
    done.append(re.match(r"(\d+) (\S+)", f"{m[-1][1]} {m[-1][0]}") if m and len(set(m)) == 1 else None)
print("rcs", rcs, "done", [d.groups() if d else None for d in done])
rounds = [[float(t) for t in re.findall(r"swarm-averaged gradients, ([0-9.]+)s", l)] for l in logs]
print("global-step times of swarm-averaged epochs (reduce inside the group + swarm round + group LAMB step), s:", rounds)
print("swarm-averaged rounds:", [l.count("swarm-averaged") for l in logs], "group-local:", [l.count("group-local") for l in logs])
print("aux epochs:", re.findall(r"Epoch #(\d+)", auxl)[-5:])
ok = all(r == 0 for r in rcs) and all(done) and all(int(d.group(1)) >= 2 for d in done) and \
    all(l.count("swarm-averaged") >= 1 for l in logs) and re.search(r"Epoch #\d+", auxl)
if ok:
    a, b = float(done[0].group(2)), float(done[1].group(2))
    # gradients travel 8-bit compressed and every peer applies x += delta (hivemind semantics), so the two groups
    # agree up to the compression error, not bit for bit
    rel = abs(a - b) / max(abs(a), 1e-9)
    print(f"CONFIG5_OK param checksums {a} vs {b} (rel diff {rel:.2e})" if rel < 1e-4 else f"CONFIG5_DRIFT {a} {b}")
else:
    print("CONFIG5_FAIL")
    for l in logs: print(l[-2500:])
    print(auxl[-1500:])
