"""``CollaborativeCallback``: swarm metrics, NaN guard with rollback, local backups (reference callback.py:14-127).

Driven by this repository's own loop (``dalle_b200.training.loop``): ``on_step_end(loss=...)`` receives the loss
directly instead of digging it out of HF ``TrainerState.log_history``.
"""
from __future__ import annotations

import logging
import os
from typing import Any, Optional

import torch

from arguments import TrainingPeerArguments
from dalle_b200.dht import get_dht_time
from utils import LocalMetrics

logger = logging.getLogger(__name__)


class CollaborativeCallback:
    def __init__(self, task, args: TrainingPeerArguments):
        self.task = task
        self.dht, self.collaborative_optimizer = task.dht, task.collaborative_optimizer
        self.statistics_expiration = args.statistics_expiration
        self.backup_every_steps, self.state_path = args.backup_every_steps, args.state_path
        self.last_reported_collaboration_step = -1
        self.samples = self.steps = self.total_samples_processed = 0
        self.loss = 0.0

    # ---- hooks --------------------------------------------------------------------------------------
    def on_train_begin(self, *unused, **kwargs):
        have_backup = os.path.isfile(self.state_path)
        if have_backup:
            self.restore_from_backup(self.state_path)
            logger.info("Loaded state")
        logger.info("Loading state from peers")
        self.collaborative_optimizer.load_state_from_peers()
        if have_backup:
            self.restore_from_backup(self.state_path, check_step=True)

    def on_step_end(self, *unused, loss: Optional[float] = None, **kwargs):
        opt = self.collaborative_optimizer
        ok = self.params_are_finite()
        if hasattr(opt, "all_ranks_agree"):   # a multi-GPU peer rolls back as one: every rank or none
            ok = opt.all_ranks_agree(ok)
        if not ok:
            if not os.path.exists(self.state_path):
                raise RuntimeError("Encountered broken parameters, but there is no backup to fall back to.")
            logger.warning("Parameters are invalid, reloading model from earlier state")
            self.restore_from_backup(self.state_path)
            return
        if loss is not None:
            self.loss += float(loss)
            self.steps += 1
            if opt.local_epoch != self.last_reported_collaboration_step:
                self.last_reported_collaboration_step = opt.local_epoch
                self.total_samples_processed += self.samples
                sps = float(opt.tracker.performance_ema.samples_per_second)
                record = LocalMetrics(step=opt.local_epoch, samples_per_second=sps, samples_accumulated=self.samples,
                                      loss=self.loss, mini_steps=self.steps)
                logger.info(f"Current epoch: {opt.local_epoch}")
                logger.info(f"Your current contribution: {self.total_samples_processed} samples")
                logger.info(f"Performance: {sps} samples/sec")
                if self.steps:
                    logger.info(f"Local loss: {self.loss / self.steps}")
                self.loss, self.steps = 0.0, 0
                if opt.local_epoch == opt.tracker.global_epoch and self.dht is not None:
                    self.dht.store(key=opt.run_id + "_metrics", subkey=self.task.local_public_key,
                                   value=record.dict(), expiration_time=get_dht_time() + self.statistics_expiration,
                                   return_future=True)
                if self.backup_every_steps is not None and opt.local_epoch % self.backup_every_steps == 0:
                    self.backup_state()
        # contribution of the epoch that just finished is reported one call later (reference semantics)
        self.samples = opt.grad_averager.local_samples_accumulated

    # ---- guard & backups ----------------------------------------------------------------------------
    @torch.no_grad()
    def params_are_finite(self) -> bool:
        model = self.task.model
        flat = getattr(next(model.parameters()), "_flat_layout", None)
        if flat is not None:  # one reduction over the flat buffer instead of ~650 per-tensor scans + syncs
            return bool(torch.isfinite(flat.flat).all())
        return all(bool(torch.isfinite(p).all()) for p in model.parameters())

    @torch.no_grad()
    def backup_state(self) -> Any:
        opt = self.collaborative_optimizer
        # (in a multi-GPU peer state_dict() is collective - it gathers the sharded optimizer state - and the leader
        # rank writes the file)
        state = {"model": self.task.model.state_dict(), "training": opt.state_dict(),
                 "scheduler": opt.state_averager.scheduler.state_dict(), "local_epoch": opt.local_epoch}
        if not getattr(opt, "is_leader", True):
            return None
        logger.info("Saving backup")
        tmp = f"{self.state_path}.tmp{os.getpid()}"
        torch.save(state, tmp)
        os.replace(tmp, self.state_path)   # readers (other ranks restoring after a NaN) never see a partial file
        return None

    @torch.no_grad()
    def restore_from_backup(self, path, check_step: bool = False):
        state = torch.load(path, map_location="cpu", weights_only=False)
        opt = self.collaborative_optimizer
        if check_step and state["local_epoch"] < opt.local_epoch:
            logger.info("Bypassed restoring state from local backup: backup state is too old.")
            return
        self.task.model.load_state_dict(state["model"], strict=False)
        opt.load_state_dict(state["training"])
        opt.state_averager.scheduler.load_state_dict(state["scheduler"])
        opt.state_averager.local_epoch = state["local_epoch"]
        logger.info("Restored from a backup")
