"""Two-level collaborative step for a multi-GPU peer (level 2 over level 1; BASELINE config 5).

A "peer" of the swarm is a whole group of GPUs in one box.  Inside the group gradients are reduced and the optimizer
runs through the fused NVLink kernel (parallel/group.py); between groups the group LEADER (rank 0) speaks the TCP
swarm protocol - progress records in the DHT, matchmaking, butterfly all-reduce with fp16 / 8-bit compression -
with ``batch_size_per_step`` = the whole group's samples and weight = the group's accumulated samples.

Global step (when the swarm reached ``target_batch_size``):
  1. ``group.reduce_gradients``  : every rank ends with the group-mean gradient (NVLink reduce-scatter + all-gather)
  2. leader: D2H -> swarm all-reduce (weighted by samples, compressed) -> H2D; on failure keeps the group mean
  3. non-leaders zero their gradient buffer; ``group.step`` (reduce + LAMB-8bit shard + parameter all-gather) then
     reduces "leader's averaged gradient + zeros" - one code path whether or not other peers exist.
``torch.distributed`` (NCCL) is used only for the two tiny control broadcasts (decision flags).
This also covers the capability of the reference's TPU entry point (run_trainer_tpu.py + lib/training/tpu.py: one
peer = one host feeding N accelerator workers with host-side aggregation).
"""
from __future__ import annotations

import logging
import time
from typing import Optional

import torch
import torch.distributed as dist

from ..dht import DHT
from . import compression as C
from .averager import DecentralizedAverager
from .progress_tracker import ProgressTracker

logger = logging.getLogger(__name__)


class _Samples:
    local_samples_accumulated = 0
    reuse_grad_buffers = True


class GroupCollaborativeOptimizer:
    def __init__(self, trainer, *, dht: Optional[DHT], run_id: str, target_batch_size: int, batch_size_per_step: int,
                 matchmaking_time: float = 15.0, averaging_timeout: float = 180.0, allreduce_timeout: float = 60.0,
                 grad_compression: C.CompressionBase = C.NoCompression(), target_group_size: Optional[int] = None,
                 tracker_opts: Optional[dict] = None, **unused):
        self.trainer, self.run_id = trainer, run_id
        self.rank, self.world = trainer.rank, trainer.world
        self.leader = self.rank == 0
        self.batch_size_per_step, self.target_batch_size = batch_size_per_step, target_batch_size
        self.averaging_timeout, self._target_group_size = averaging_timeout, target_group_size
        self.local_epoch, self.samples, self.times = 0, 0, 0
        self.grad_averager = _Samples()
        self.layout = trainer.opt._fused.layout
        self.flags = torch.zeros(2, dtype=torch.int64, device=trainer.device)
        self.tracker = self.averager = None
        if self.leader:
            assert dht is not None
            self.dht = dht
            sizes = [(self.layout.offsets[i + 1] if i + 1 < len(self.layout.offsets) else self.layout.n_padded)
                     - self.layout.offsets[i] for i in range(len(self.layout.offsets))]
            self.averager = DecentralizedAverager([(s,) for s in sizes], dht, prefix=f"{run_id}_grad_averager",
                                                  min_matchmaking_time=matchmaking_time,
                                                  allreduce_timeout=allreduce_timeout, compression=grad_compression,
                                                  allow_state_sharing=False, pin_memory=True)
            self.tracker = ProgressTracker(dht, run_id, target_batch_size, **(tracker_opts or {}))
        self.last_round_averaged = False

    @property
    def param_groups(self):
        return self.trainer.opt.param_groups

    def step(self):
        """Call on EVERY rank after backward."""
        self.samples += self.batch_size_per_step
        self.times += 1
        self.grad_averager.local_samples_accumulated = self.samples
        if self.leader:
            self.tracker.report_local_progress(self.local_epoch, self.samples)
            self.flags[0] = int(self.tracker.ready_to_update_epoch)
            self.flags[1] = int(self.tracker.global_progress.num_peers > 1 or bool(self._target_group_size))
        if self.world > 1:
            dist.broadcast(self.flags, src=0)
        do_global, multi_peer = bool(self.flags[0].item()), bool(self.flags[1].item())
        if do_global:
            self._global_step(multi_peer)
        return do_global

    def _global_step(self, multi_peer: bool):
        tr, L = self.trainer, self.layout
        scale = 1.0 / (self.world * max(self.times, 1))
        t0 = time.perf_counter()
        ctx = self.tracker.pause_updates() if self.leader else _null()
        with ctx:
            if multi_peer and tr.group is not None:
                tr.group.reduce_gradients(grad_scale=scale)
                ok = 0
                if self.leader:
                    torch.cuda.synchronize()
                    self.averager._flat.copy_(L.flat_grad, non_blocking=False)
                    hint = self._target_group_size or (self.tracker.global_progress.num_peers or None)
                    res = self.averager.step(weight=float(self.samples), timeout=self.averaging_timeout,
                                             target_group_size=hint)
                    ok = int(res is not None)
                    if ok:
                        L.flat_grad.copy_(self.averager._flat, non_blocking=False)
                    self.flags[0] = ok
                dist.broadcast(self.flags, src=0)      # also orders "leader's H2D is done" before the fused step
                self.last_round_averaged = bool(self.flags[0].item())
                if not self.leader:
                    L.flat_grad.zero_()
                tr.group.step(grad_scale=1.0)
            elif multi_peer:   # single-GPU peer
                if self.leader:
                    torch.cuda.synchronize()
                    self.averager._flat.copy_(L.flat_grad)
                    self.averager._flat.mul_(scale)
                    res = self.averager.step(weight=float(self.samples), timeout=self.averaging_timeout,
                                             target_group_size=self._target_group_size)
                    self.last_round_averaged = res is not None
                    L.flat_grad.copy_(self.averager._flat)
                tr.opt.step(grad_scale=1.0, zero_grad=True)
            else:
                if tr.group is not None:
                    tr.group.step(grad_scale=scale)
                else:
                    tr.opt.step(grad_scale=scale, zero_grad=True)
            tr.sched.step()
            self.local_epoch += 1
            self.samples = self.times = 0
            if self.leader:
                self.local_epoch = max(self.local_epoch, self.tracker.global_epoch)
                self.tracker.update_epoch(self.local_epoch)
        if self.leader:
            logger.info(f"global step -> epoch {self.local_epoch} ({'swarm-averaged' if self.last_round_averaged else 'group-local'} "
                        f"gradients, {time.perf_counter() - t0:.2f}s)")

    def shutdown(self):
        if self.leader:
            self.tracker.shutdown()
            self.averager.shutdown()


class _null:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False
