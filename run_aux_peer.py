#!/usr/bin/env python3
"""Auxiliary peer (same flags as reference run_aux_peer.py:85-152): welcomes peers into the DHT, aggregates the
signed per-peer metrics published under ``{prefix}_metrics``, periodically pulls the training state from peers and
writes ``model_state.pt`` / ``optimizer_state.pt`` checkpoints, optionally assists averaging with weight 0
(``--assist_in_averaging``, which the reference advertises but leaves unimplemented)."""
import logging
import os
import threading
import time

import torch

import utils
from arguments import (AuxiliaryPeerArguments, CollaborativeArguments, HFTrainerArguments, parse_dataclasses)
from task import TrainingTask

utils.use_log_handler()
logger = logging.getLogger(__name__)


class CheckpointHandler:
    def __init__(self, task: TrainingTask, peer_args: AuxiliaryPeerArguments):
        self.task, self.peer_args = task, peer_args
        self.save_checkpoint_step_interval = peer_args.save_checkpoint_step_interval
        self.local_path, self.upload_interval = peer_args.local_path, peer_args.upload_interval
        self.repo, self.last_upload_time, self.previous_step = None, None, -1
        os.makedirs(self.local_path, exist_ok=True)
        if self.upload_interval is not None and peer_args.repo_url:
            assert task.authorizer is not None, "Model uploading needs Hugging Face auth to be enabled"
            try:
                from huggingface_hub import Repository

                self.repo = Repository(local_dir=self.local_path, clone_from=peer_args.repo_url,
                                       use_auth_token=task.authorizer.hf_user_access_token)
            except Exception as e:
                logger.warning(f"Hub repository unavailable ({e!r}); checkpoints stay local")

    def should_save_state(self, cur_step: int) -> bool:
        return (self.save_checkpoint_step_interval is not None
                and cur_step - self.previous_step >= self.save_checkpoint_step_interval)

    def save_state(self, cur_step: int):
        logger.info("Saving state from peers")
        self.task.collaborative_optimizer.load_state_from_peers()
        self.previous_step = cur_step

    def is_time_to_upload(self) -> bool:
        if self.upload_interval is None:
            return False
        return self.last_upload_time is None or time.time() - self.last_upload_time >= self.upload_interval

    def upload_checkpoint(self, current_loss: float):
        self.last_upload_time = time.time()
        opt = self.task.collaborative_optimizer
        logger.info("Saving model")
        torch.save(self.task.model.state_dict(), f"{self.local_path}/model_state.pt")
        logger.info("Saving optimizer")
        torch.save(opt.state_dict(), f"{self.local_path}/optimizer_state.pt")
        if self.repo is None:
            return
        try:
            self.repo.git_pull()
            self.repo.push_to_hub(commit_message=f"Epoch {opt.local_epoch}, loss {current_loss:.3f}")
            logger.info("Finished uploading to Model Hub")
        except Exception:
            logger.exception("Uploading the checkpoint to HF Model Hub failed:")


def assist_averaging_in_background(task: TrainingTask, peer_args: AuxiliaryPeerArguments, stop: threading.Event):
    while not stop.is_set():
        time.sleep(peer_args.assist_refresh)
        try:
            task.collaborative_optimizer.step()
        except Exception:
            logger.exception("assisting averaging failed")


def aggregate(metrics):
    """(loss, alive peers, samples, samples/s, latest step) from LocalMetrics records; guards the empty epoch."""
    mini = sum(m.mini_steps for m in metrics)
    return (sum(m.loss for m in metrics) / mini if mini else float("nan"), len(metrics),
            sum(m.samples_accumulated for m in metrics), sum(m.samples_per_second for m in metrics),
            max(m.step for m in metrics))


def main(argv=None, max_iterations=None):
    peer_args, trainer_args, collab_args = parse_dataclasses(
        (AuxiliaryPeerArguments, HFTrainerArguments, CollaborativeArguments), argv)
    task = TrainingTask(peer_args, trainer_args, collab_args)
    task.auxiliary = bool(peer_args.assist_in_averaging)
    dht, opt = task.dht, task.collaborative_optimizer
    wandb = None
    if peer_args.wandb_project is not None:
        try:
            import wandb

            wandb.init(project=peer_args.wandb_project)
        except Exception as e:
            logger.warning(f"wandb unavailable: {e!r}")
            wandb = None
    handler = CheckpointHandler(task, peer_args) if peer_args.store_checkpoints else None
    stop = threading.Event()
    if peer_args.assist_in_averaging:
        assert not peer_args.client_mode, "client-mode peers cannot assist in averaging"
        threading.Thread(target=assist_averaging_in_background, args=(task, peer_args, stop), daemon=True,
                         name="AveragingAuxThread").start()
    current_step, history, it = 0, [], 0
    try:
        while max_iterations is None or it < max_iterations:
            it += 1
            entry = dht.get(peer_args.experiment_prefix + "_metrics", latest=True)
            if entry is not None and len(entry.value) > 0:
                metrics = [utils.LocalMetrics.parse_obj(v.value) for v in entry.value.values()]
                loss, alive, samples, perf, latest = aggregate(metrics)
                if latest != current_step:
                    current_step = latest
                    logger.info(f"Epoch #{current_step}\tloss = {loss:.5f}\tpeers = {alive}\tsamples/s = {perf:.1f}")
                    history.append({"loss": loss, "alive peers": alive, "samples": samples, "performance": perf,
                                    "step": latest})
                    if wandb is not None:
                        wandb.log(history[-1])
                    if handler is not None and handler.should_save_state(current_step):
                        handler.save_state(current_step)
                        if handler.is_time_to_upload():
                            handler.upload_checkpoint(loss)
            logger.debug("Peer is still alive...")
            time.sleep(peer_args.refresh_period)
    finally:
        stop.set()
    return task, history


if __name__ == "__main__":
    main()
