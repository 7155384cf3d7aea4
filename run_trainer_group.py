#!/usr/bin/env python3
"""Multi-GPU collaborative trainer: ONE swarm peer = the N GPUs of this launch (torchrun, one rank per GPU).

    python -m torch.distributed.run --nproc-per-node 4 --master-addr 127.0.0.1 run_trainer_group.py \
        --initial_peers /ip4/.../tcp/.../p2p/... --authorize false [same flags as run_trainer.py]

Kept as a name for launch scripts: ``run_trainer.py`` itself is group-aware (under torchrun its ranks form one peer:
same TrainingTask, dataset, callback, backups, state download and authorizer as a single-GPU peer; fused NVLink
averaging + sharded LAMB-8bit inside the group, the leader rank on the swarm's TCP plane), so this file only forwards
to it.  Covers BASELINE config 5 and the "one host feeds N accelerator workers" capability of the reference's
run_trainer_tpu.py + lib/training/tpu.py."""
import run_trainer

if __name__ == "__main__":
    _task, _ = run_trainer.main()
    run_trainer.shutdown(_task)
