#!/usr/bin/env python3
"""Collaborative trainer entry point (same flags as reference run_trainer.py:26-60)."""
import hashlib
import logging
import os
from pathlib import Path

import torch

import callback
import utils
from arguments import (CollaborativeArguments, HFTrainerArguments, TrainingPeerArguments, parse_dataclasses)
from dalle_b200.training.loop import CollaborativeTrainer
from task import TrainingTask

utils.use_log_handler()
logger = logging.getLogger(__name__)


def main(argv=None, max_micro_steps=None):
    peer_args, trainer_args, collab_args = parse_dataclasses(
        (TrainingPeerArguments, HFTrainerArguments, CollaborativeArguments), argv)
    # under torchrun the ranks of the launch form ONE swarm peer (one process per GPU): rank 0 joins the swarm, all
    # ranks train, average inside the group over NVLink and checkpoint together
    from dalle_b200.trainer import init_distributed_from_env

    on_cpu = trainer_args.no_cuda or not torch.cuda.is_available()
    rank, local, world = init_distributed_from_env("gloo" if on_cpu else "nccl")
    if world > 1:
        trainer_args.local_rank = local   # -> trainer_args.device = cuda:<local rank>
    if rank == 0:
        logger.info(f"Trying {len(peer_args.initial_peers)} initial peers: {peer_args.initial_peers}")
    utils.log_process_rank(trainer_args)
    task = TrainingTask(peer_args, trainer_args, collab_args)
    model = task.model.to(trainer_args.device)
    assert trainer_args.do_train and not trainer_args.do_eval
    cb = callback.CollaborativeCallback(task, peer_args)
    seed = int.from_bytes(hashlib.sha256(task.local_public_key).digest()[:8], "little")
    trainer = CollaborativeTrainer(model=model, args=trainer_args, data_collator=task.data_collator,
                                   train_dataset=task.training_dataset, data_seed=seed,
                                   collaborative_optimizer=task.collaborative_optimizer, callbacks=[cb],
                                   max_micro_steps=max_micro_steps)
    latest = max(Path(trainer_args.output_dir).glob("checkpoint*"), key=os.path.getctime, default=None)
    steps = trainer.train(model_path=latest)
    return task, steps


def shutdown(task):
    task.collaborative_optimizer.shutdown()   # waits for a state-averaging round still in flight
    if task.group_world > 1:   # every rank of a multi-GPU peer must end with the same parameters, bit for bit
        flat = torch.cat([p.detach().double().reshape(-1) for p in task.model.parameters()])
        logger.info(f"GROUP_PARAM_CHECKSUM {float(flat.sum()):.10e} epoch {task.collaborative_optimizer.local_epoch}")
    if task.dht is not None:
        task.dht.shutdown()
    if torch.distributed.is_available() and torch.distributed.is_initialized():
        torch.distributed.barrier()
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    _task, _ = main()
    shutdown(_task)
