#!/usr/bin/env python3
"""Multi-GPU collaborative trainer: ONE swarm peer = the N GPUs of this launch (torchrun, one rank per GPU).

    python -m torch.distributed.run --nproc-per-node 4 --master-addr 127.0.0.1 run_trainer_group.py \
        --initial_peers /ip4/.../tcp/.../p2p/... --authorize false [same flags as run_trainer.py]

Inside the group: fused NVLink averaging + sharded LAMB-8bit (no NCCL on that path); between groups: the leader rank
joins the TCP swarm (DHT progress, matchmaking, compressed butterfly all-reduce) on behalf of the whole group.
Covers BASELINE config 5 and the "one host feeds N accelerator workers" capability of the reference's
run_trainer_tpu.py."""
import logging
import os
import time

import torch

import utils
from arguments import (CollaborativeArguments, HFTrainerArguments, TrainingPeerArguments, parse_dataclasses)
from dalle_b200.collab import Float16Compression, SizeAdaptiveCompression, Uniform8BitQuantization
from dalle_b200.collab.group_optimizer import GroupCollaborativeOptimizer
from dalle_b200.config import get_config
from dalle_b200.dht import DHT, get_dht_time
from dalle_b200.trainer import GroupTrainer, init_distributed_from_env

utils.use_log_handler()
logger = logging.getLogger(__name__)


def main(argv=None):
    peer_args, trainer_args, collab_args = parse_dataclasses(
        (TrainingPeerArguments, HFTrainerArguments, CollaborativeArguments), argv)
    rank, local, world = init_distributed_from_env()
    device = torch.device("cuda", local)
    cfg = get_config(trainer_args.model_preset)
    a = trainer_args
    tr = GroupTrainer(cfg, device, lr=a.learning_rate, betas=(a.adam_beta1, a.adam_beta2), eps=a.adam_epsilon,
                      weight_decay=a.weight_decay, max_grad_norm=a.max_grad_norm, clamp_value=a.clamp_value,
                      warmup_steps=a.warmup_steps, total_steps=a.total_steps, seed=a.seed,
                      averaging="fused" if world > 1 else "nccl")
    dht = validators = pubkey = None
    if rank == 0:
        validators, pubkey = utils.make_validators(peer_args.experiment_prefix)
        dht = DHT(start=True, initial_peers=peer_args.initial_peers, client_mode=peer_args.client_mode,
                  host_maddrs=peer_args.host_maddrs, announce_maddrs=peer_args.announce_maddrs,
                  record_validators=validators, identity_path=peer_args.identity_path)
        utils.log_visible_maddrs(dht.get_visible_maddrs(), only_p2p=False)
    group_batch = a.per_device_train_batch_size * a.gradient_accumulation_steps * world
    comp = SizeAdaptiveCompression(2 ** 16 + 1, Float16Compression(), Uniform8BitQuantization())
    opt = GroupCollaborativeOptimizer(tr, dht=dht, run_id=peer_args.experiment_prefix,
                                      target_batch_size=collab_args.target_batch_size, batch_size_per_step=group_batch,
                                      matchmaking_time=collab_args.matchmaking_time,
                                      averaging_timeout=collab_args.averaging_timeout,
                                      allreduce_timeout=collab_args.allreduce_timeout, grad_compression=comp,
                                      target_group_size=int(os.environ.get("DALLE_B200_TARGET_GROUP_SIZE", "0")) or None,
                                      tracker_opts=dict(min_refresh_period=0.5, default_refresh_period=1.0,
                                                        max_refresh_period=3.0))
    gen = torch.Generator().manual_seed(1000 * a.seed + 17 * rank + int(os.environ.get("DALLE_B200_DATA_SEED", "0")))
    B = a.per_device_train_batch_size
    loss_sum, steps, reported = 0.0, 0, -1
    for it in range(int(min(a.max_steps, 10 ** 9))):
        text = torch.randint(2, cfg.num_text_tokens, (B, cfg.text_seq_len), generator=gen).to(device)
        image = torch.randint(0, cfg.num_image_tokens, (B, cfg.image_seq_len), generator=gen).to(device)
        loss = tr.accumulate(text, image)
        loss_sum += float(loss)
        steps += 1
        samples_before = opt.samples + group_batch
        if opt.step() and rank == 0 and opt.local_epoch != reported:
            reported = opt.local_epoch
            rec = utils.LocalMetrics(step=opt.local_epoch, samples_accumulated=samples_before, loss=loss_sum,
                                     samples_per_second=float(opt.tracker.performance_ema.samples_per_second),
                                     mini_steps=steps)
            logger.info(f"Current epoch: {opt.local_epoch}  local loss {loss_sum / steps:.4f}")
            dht.store(peer_args.experiment_prefix + "_metrics", rec.dict(),
                      get_dht_time() + peer_args.statistics_expiration, subkey=pubkey, return_future=True)
            loss_sum, steps = 0.0, 0
    torch.cuda.synchronize()
    # all ranks of a group must agree bit for bit on the parameters
    flat = opt.layout.flat
    ref = flat.clone()
    torch.distributed.broadcast(ref, src=0) if world > 1 else None
    assert torch.equal(ref, flat), "ranks of the group diverged"
    if rank == 0:
        logger.info(f"GROUP_DONE epoch {opt.local_epoch} param_checksum {float(flat.double().sum()):.6f}")
        time.sleep(1.0)
        opt.shutdown()
        dht.shutdown()
    if world > 1:
        torch.distributed.barrier()
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
