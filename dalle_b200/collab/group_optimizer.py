"""``GroupOptimizer``: the collaborative optimizer of a multi-GPU peer (level 2 over level 1; BASELINE
config 5) with the SAME surface as :class:`dalle_b200.collab.Optimizer`, i.e. everything the reference's
``callback.py`` / ``run_aux_peer.py`` / ``hf_trainer.py`` poke at on ``hivemind.Optimizer``:
``step``, ``param_groups``, ``scheduler``, ``local_epoch``, ``run_id``, ``tracker.global_epoch``,
``tracker.performance_ema.samples_per_second``, ``grad_averager.local_samples_accumulated``,
``state_averager.scheduler`` / ``.local_epoch``, ``state_dict`` / ``load_state_dict`` / ``load_state_from_peers``,
``callbacks``.

One swarm "peer" is the whole group of ranks of a ``torchrun`` launch (one process per GPU).  Inside the group the
gradients are reduced and the sharded LAMB-8bit update runs through the fused NVLink kernel (parallel/group.py; on CPU
ranks - the gloo plumbing tests - through ``all_reduce`` + the reference LAMB); between groups the LEADER (rank 0)
speaks the TCP swarm protocol with ``batch_size_per_step`` = the whole group's samples:

* progress records, matchmaking and the compressed butterfly all-reduce of gradients - the leader's averager uses the
  real parameter shapes and the same DHT prefixes as a single-GPU ``Optimizer``, so groups, single-GPU peers and
  assisting aux peers can share one swarm;
* parameter (state) averaging after the update (``average_state_every``), delayed onto a background thread when
  ``delay_optimizer_step`` is set and applied - to every rank, by a broadcast - at a later ``step()``;
* ``load_state_from_peers``: the leader downloads, every rank receives parameters + optimizer state + epoch +
  scheduler by broadcast; used at start-up and whenever the group fell more than one epoch behind;
* checkpoints: the optimizer state is sharded ZeRO-1 style over the ranks, ``state_dict()`` first gathers the shards
  so that the result is the complete Appendix-A.2 layout (loads into a single-GPU peer and vice versa).  The same
  gather runs after every global step so that the leader can serve state downloads from its own copy.

Every method that the training loop or the callback calls (``step``, ``state_dict``, ``load_state_dict``,
``load_state_from_peers``, ``all_ranks_agree``) is COLLECTIVE: call it on every rank.
"""
from __future__ import annotations

import logging
import threading
import time
from concurrent.futures import Future, ThreadPoolExecutor
from typing import Callable, List, Optional

import torch
import torch.distributed as dist

from ..dht import DHT, get_dht_time
from . import compression as C
from .averager import DecentralizedAverager
from .optimizer import GradientAverager, TrainingStateAverager, _Callbacks
from .progress_tracker import ProgressTracker

logger = logging.getLogger(__name__)


class _PerfMirror:
    samples_per_second = 0.0


class _TrackerMirror:
    """What non-leader ranks know about the swarm: refreshed by the leader's control broadcast of every step."""

    def __init__(self):
        self.global_epoch = 0
        self.performance_ema = _PerfMirror()


class _SamplesMirror:
    local_samples_accumulated = 0
    local_times_accumulated = 0
    reuse_grad_buffers = True


class _LocalState:
    """``opt.state_averager`` on non-leader ranks: the local optimizer / scheduler / epoch, no networking."""

    def __init__(self, optimizer, scheduler, main_parameters):
        self.optimizer, self.scheduler, self.main_parameters = optimizer, scheduler, main_parameters
        self.local_epoch = 0
        self.lock_optimizer = threading.Lock()
        self.state_sharing_priority = 0.0

    def shutdown(self):
        pass


class _PackedGradAverager(GradientAverager):
    """Leader's gradient averager: one packed device buffer and ONE host copy each way instead of a copy per tensor
    (the TTT model has 423 parameter tensors).  The gradients it is handed are already the group mean."""

    @torch.no_grad()
    def load_accumulators_into_averager_(self):
        grads = [p.grad.reshape(-1) for p in self.parameters]
        with self.get_tensors():
            if grads[0].is_cuda:
                packed = torch.cat(grads)
                self._flat.copy_(packed, non_blocking=False)
            else:
                torch.cat(grads, out=self._flat)

    @torch.no_grad()
    def put_averaged_grads_into_params_(self):
        with self.get_tensors():
            dev = self.parameters[0].grad.device
            packed = self._flat.to(dev, non_blocking=False)
            for p, piece in zip(self.parameters, packed.split(self._sizes)):
                p.grad.copy_(piece.view_as(p.grad))

    def reset_accumulated_grads_(self):
        self.local_samples_accumulated = self.local_times_accumulated = 0   # the group step zeroes the buffers


class _PackedStateAverager(TrainingStateAverager):
    """Leader's state averager with one packed copy each way between the device parameters and the host buffers."""

    @torch.no_grad()
    def _params_to_averager(self):
        with self.get_tensors():
            self._flat.copy_(torch.cat([p.detach().reshape(-1) for p in self.main_parameters]), non_blocking=False)

    @torch.no_grad()
    def _averager_to_params(self):
        with self.get_tensors():
            packed = self._flat.to(self.main_parameters[0].device, non_blocking=False)
            for p, piece in zip(self.main_parameters, packed.split(self._sizes)):
                p.copy_(piece.view_as(p))


class GroupOptimizer:
    def __init__(self, *, dht: Optional[DHT], run_id: str, target_batch_size: int, batch_size_per_step: int,
                 optimizer, params, scheduler=None, matchmaking_time: float = 15.0, averaging_timeout: float = 60.0,
                 allreduce_timeout: Optional[float] = None, load_state_timeout: float = 600.0,
                 reuse_grad_buffers: bool = True, offload_optimizer: bool = False, delay_optimizer_step: bool = False,
                 delay_grad_averaging: bool = False, average_state_every: int = 1, client_mode: Optional[bool] = None,
                 auxiliary: bool = False, grad_compression: C.CompressionBase = C.NoCompression(),
                 state_averaging_compression: C.CompressionBase = C.NoCompression(),
                 load_state_compression: C.CompressionBase = C.NoCompression(), averager_opts: Optional[dict] = None,
                 tracker_opts: Optional[dict] = None, verbose: bool = False, target_group_size: Optional[int] = None,
                 process_group=None, payload: Optional[str] = None, gather_state_every_epoch: bool = True):
        assert dist.is_initialized(), "GroupOptimizer runs under torchrun (one process per GPU)"
        assert reuse_grad_buffers, "a group accumulates in the model's own .grad buffers"
        assert not auxiliary, "auxiliary peers are single processes (run_aux_peer.py)"
        self.pg = process_group
        self.rank, self.world = dist.get_rank(process_group), dist.get_world_size(process_group)
        self.is_leader = self.rank == 0
        self.run_id, self.dht = run_id, dht
        self.target_batch_size, self.batch_size_per_step = target_batch_size, batch_size_per_step
        self.matchmaking_time, self.averaging_timeout = matchmaking_time, averaging_timeout
        self.load_state_timeout, self.average_state_every = load_state_timeout, average_state_every
        self.delay_optimizer_step, self.verbose = delay_optimizer_step, verbose
        self.gather_state_every_epoch = gather_state_every_epoch
        self._target_group_size = target_group_size
        self.callbacks = _Callbacks()
        self.auxiliary = False

        groups = params if isinstance(params, (list, tuple)) and params and isinstance(params[0], dict) \
            else [{"params": list(params)}]
        self.main_parameters: List[torch.nn.Parameter] = [p for g in groups for p in g["params"]]
        self.opt = optimizer(params) if callable(optimizer) and not isinstance(optimizer, torch.optim.Optimizer) \
            else optimizer
        self.sched = scheduler(self.opt) if scheduler is not None and not hasattr(scheduler, "step") else scheduler
        self.device = self.main_parameters[0].device
        self.fused = self.device.type == "cuda"
        self.group = None
        if self.fused:
            import os

            from ..parallel.group import FusedGroupAverager

            self.opt._build_fused()
            if self.world > 1:
                self.group = FusedGroupAverager(self.opt, process_group,
                                                payload=payload or os.environ.get("DALLE_B200_GROUP_PAYLOAD", "fp32"))
        # [do_global, multi_peer, global_epoch, samples/s, behind, apply_state_round]
        self._ctl = torch.zeros(6, dtype=torch.float64, device=self.device)
        self.samples = self.times = 0
        self.last_round_averaged = False

        if self.is_leader:
            assert dht is not None, "the group leader (rank 0) owns the DHT"
            self.client_mode = dht.client_mode if client_mode is None else client_mode
            opts = dict(min_matchmaking_time=matchmaking_time, allreduce_timeout=allreduce_timeout,
                        client_mode=self.client_mode, **(averager_opts or {}))
            self.state_averager = _PackedStateAverager(
                dht=dht, optimizer=self.opt, params=params, scheduler=self.sched, prefix=f"{run_id}_state_averager",
                compression=state_averaging_compression, state_compression=load_state_compression, **opts)
            self.grad_averager = _PackedGradAverager(
                self.main_parameters, dht=dht, prefix=f"{run_id}_grad_averager", reuse_grad_buffers=True,
                compression=grad_compression, allow_state_sharing=False, **opts)
            self.tracker = ProgressTracker(dht, run_id, target_batch_size, client_mode=self.client_mode,
                                           **(tracker_opts or {}))
        else:
            self.client_mode = bool(client_mode)
            self.state_averager = _LocalState(self.opt, self.sched, self.main_parameters)
            self.grad_averager = _SamplesMirror()
            self.tracker = _TrackerMirror()
        self._pool = ThreadPoolExecutor(max_workers=1)
        self._pending_state_round: Optional[Future] = None
        self._scheduled_grads = None       # leader: pre-scheduled gradient round (StepControl)
        self.scheduled_rounds_used = 0

    # ---- surface used by the reference ----------------------------------------------------------
    @property
    def local_epoch(self) -> int:
        return self.state_averager.local_epoch

    @property
    def param_groups(self):
        return self.opt.param_groups

    @property
    def scheduler(self):
        return self.sched

    def zero_grad(self, set_to_none: bool = False):
        raise ValueError("the group optimizer owns .grad (reuse_grad_buffers): gradients are reset after each "
                         "global step")

    # ---- small collectives ----------------------------------------------------------------------
    def _bcast(self, t: torch.Tensor):
        if self.world > 1:
            dist.broadcast(t, src=dist.get_global_rank(self.pg, 0) if self.pg is not None else 0, group=self.pg)
        return t

    def all_ranks_agree(self, ok: bool) -> bool:
        """True iff ``ok`` on EVERY rank (e.g. "my parameters are finite"): the group acts as one peer."""
        if self.world == 1:
            return bool(ok)
        t = torch.tensor([1.0 if ok else 0.0], device=self.device)
        dist.all_reduce(t, op=dist.ReduceOp.MIN, group=self.pg)
        return bool(t.item() > 0.5)

    def _layout(self):
        return self.opt._fused.layout if self.fused and self.opt._fused is not None else None

    def _refresh_shadows(self):
        L = self._layout()
        if L is not None and L.flat_bf16 is not None:
            L.flat_bf16.copy_(L.flat)
            L.mark_shadows_fresh()

    def _broadcast_parameters(self):
        L = self._layout()
        if L is not None:
            self._bcast(L.flat)
        else:
            for p in self.main_parameters:
                self._bcast(p.data)
        self._refresh_shadows()

    # ---- optimizer state: sharded inside the group, complete outside ----------------------------
    @torch.no_grad()
    def _gather_full_state(self):
        """Every rank ends with the complete optimizer state in its own arrays.  Each rank only ever updates the state
        of its shard of 4096-element blocks; what it holds for the other shards is stale.  Stale ranges are zeroed and
        a SUM all-reduce assembles the state (uint8 codes summed four at a time as int32: exactly one rank is non-zero
        per byte, so nothing carries)."""
        if self.group is None:
            return
        f = self.opt._fused
        import numpy as np

        from ..optim.lamb8bit import BLOCK, _BLOCK_DTYPE

        blocks = getattr(self, "_blocks_host", None)
        if blocks is None:
            blocks = self._blocks_host = f.blocks.cpu().numpy().view(_BLOCK_DTYPE)
        b0, b1 = self.group.blk0, self.group.blk1
        mine = blocks[b0:b1]
        q = mine[mine["quantized"] != 0]
        fp = mine[mine["quantized"] == 0]

        def keep_only(t: torch.Tensor, lo: int, hi: int):
            if lo > 0:
                t[:lo].zero_()
            if hi < t.numel():
                t[hi:].zero_()

        q_lo, q_hi = (int(q["state_off"].min()), int(q["state_off"].max()) + BLOCK) if len(q) else (0, 0)
        a_lo, a_hi = (int(q["absmax_idx"].min()), int(q["absmax_idx"].max()) + 1) if len(q) else (0, 0)
        f_lo, f_hi = (int(fp["state_off"].min()), int(fp["state_off"].max()) + BLOCK) if len(fp) else (0, 0)
        for t, lo, hi in ((f.q1, q_lo, q_hi), (f.q2, q_lo, q_hi), (f.absmax1, a_lo, a_hi), (f.absmax2, a_lo, a_hi),
                          (f.f1, f_lo, f_hi), (f.f2, f_lo, f_hi)):
            keep_only(t, lo, hi)
            view = t.view(torch.int32) if t.dtype == torch.uint8 else t
            dist.all_reduce(view, group=self.pg)

    def state_dict(self) -> dict:
        self._gather_full_state()
        sd = self.opt.state_dict()
        sd["state"]["local_epoch"] = self.local_epoch
        return sd

    def load_state_dict(self, state_dict: dict):
        state_dict = dict(state_dict, state=dict(state_dict["state"]))
        if "local_epoch" in state_dict["state"]:
            self.state_averager.local_epoch = state_dict["state"].pop("local_epoch")
        with self.state_averager.lock_optimizer:
            out = self.opt.load_state_dict(state_dict)
        self._refresh_shadows()
        return out

    def load_state_from_peers(self, **kwargs) -> bool:
        """Leader downloads the newest state in the swarm; every rank then receives it (collective)."""
        self._finish_pending()
        ok = False
        if self.is_leader:
            with self.tracker.pause_updates():
                try:
                    ok = self.state_averager.load_state_from_peers(timeout=self.load_state_timeout)
                except Exception as e:
                    logger.warning(f"load_state_from_peers failed: {e!r}")
        self._ctl[0] = float(ok)
        self._ctl[2] = float(self.local_epoch)
        self._bcast(self._ctl)
        ok = bool(self._ctl[0].item())
        if not ok:
            return False
        self._broadcast_state(int(self._ctl[2].item()))
        self.samples = self.times = 0
        self.grad_averager.local_samples_accumulated = self.grad_averager.local_times_accumulated = 0
        self._zero_grads()
        if self.is_leader:
            self.tracker.report_local_progress(self.local_epoch, 0)
        self.callbacks.on_load_state_from_peers.fire()
        return True

    @torch.no_grad()
    def _broadcast_state(self, epoch: int):
        """Parameters, optimizer state, scheduler and epoch of the leader -> every rank."""
        self._broadcast_parameters()
        if self.fused:
            f = self.opt._fused
            for t in (f.q1, f.q2, f.f1, f.f2, f.absmax1, f.absmax2):
                self._bcast(t)
            steps = torch.tensor([self.opt.state[p].get("step", 0) for p in self.main_parameters],
                                 dtype=torch.int64, device=self.device)
            self._bcast(steps)
            for p, s in zip(self.main_parameters, steps.tolist()):
                self.opt.state[p]["step"] = int(s)
        else:
            objs = [self.opt.state_dict() if self.is_leader else None]
            dist.broadcast_object_list(objs, src=0, group=self.pg)
            if not self.is_leader:
                self.opt.load_state_dict(objs[0])
        objs = [self.sched.state_dict() if (self.is_leader and self.sched is not None) else None]
        dist.broadcast_object_list(objs, src=0, group=self.pg)
        if not self.is_leader and self.sched is not None and objs[0] is not None:
            self.sched.load_state_dict(objs[0])
        self.state_averager.local_epoch = epoch

    def _zero_grads(self):
        L = self._layout()
        if L is not None:
            L.flat_grad.zero_()
        else:
            for p in self.main_parameters:
                if p.grad is not None:
                    p.grad.zero_()

    # ---- the step -------------------------------------------------------------------------------
    def step(self, closure=None, batch_size: Optional[int] = None, grad_scaler=None):
        """Call on EVERY rank after backward (gradients accumulate in ``.grad`` across calls)."""
        loss = closure() if closure is not None else None
        self.samples += batch_size if batch_size is not None else self.batch_size_per_step
        self.times += 1
        self.grad_averager.local_samples_accumulated = self.samples
        self.grad_averager.local_times_accumulated = self.times
        c = self._ctl
        if self.is_leader:
            tr = self.tracker
            behind = self.local_epoch < tr.global_epoch - 1
            if not behind:
                tr.report_local_progress(self.local_epoch, self.samples)
                self._maybe_schedule_gradient_averaging()
            fut = self._pending_state_round
            c[0] = float((not behind) and bool(tr.ready_to_update_epoch))
            c[1] = float(tr.global_progress.num_peers > 1 or bool(self._target_group_size))
            c[2] = float(tr.global_epoch)
            c[3] = float(tr.performance_ema.samples_per_second)
            c[4] = float(behind)
            c[5] = float(fut is not None and fut.done())
        self._bcast(c)
        do_global, multi_peer, behind, apply_state = (bool(c[0].item()), bool(c[1].item()), bool(c[4].item()),
                                                      bool(c[5].item()))
        if not self.is_leader:
            self.tracker.global_epoch = int(c[2].item())
            self.tracker.performance_ema.samples_per_second = float(c[3].item())
        if apply_state:
            self._apply_finished_state_round()
        if behind:
            if self.is_leader:
                logger.info(f"group is out of sync (local epoch {self.local_epoch}, global {self.tracker.global_epoch}): "
                            f"downloading state")
            self.load_state_from_peers()
            return loss
        if do_global:
            self._global_step(multi_peer)
        return loss

    def _maybe_schedule_gradient_averaging(self):
        """Leader only: start looking for a swarm group ``matchmaking_time`` before the tracker's ETA of the epoch
        (same rule as collab.Optimizer; reference arguments.py:66-68), so the search overlaps the last micro-batches."""
        ctl = self._scheduled_grads
        if ctl is not None and (ctl.done() or ctl.cancelled):
            self._scheduled_grads = ctl = None
        tr = self.tracker
        if ctl is not None or not (tr.global_progress.num_peers > 1 or self._target_group_size):
            return
        eta, now = tr.estimated_next_update_time, get_dht_time()
        if eta - now <= self.matchmaking_time:
            self._scheduled_grads = self.grad_averager.schedule_step(
                scheduled_time=max(eta, now + self.grad_averager.min_matchmaking_time),
                timeout=self.averaging_timeout + self.matchmaking_time, target_group_size=self._group_size_hint(),
                round_key=self.local_epoch)

    def _group_size_hint(self) -> Optional[int]:
        if self._target_group_size:
            return self._target_group_size
        n = self.tracker.global_progress.num_peers
        return n if n >= 2 else None

    def _reduce_group_gradients(self, scale: float):
        """Every rank's gradient buffer <- scale * sum over the ranks."""
        if self.group is not None:
            self.group.reduce_gradients(grad_scale=scale)
            return
        if self.world > 1:
            flat = torch.cat([p.grad.reshape(-1) for p in self.main_parameters])
            dist.all_reduce(flat, group=self.pg)
            flat.mul_(scale)
            for p, piece in zip(self.main_parameters, flat.split([p.numel() for p in self.main_parameters])):
                p.grad.copy_(piece.view_as(p.grad))
        else:
            for p in self.main_parameters:
                p.grad.mul_(scale)

    def _local_update(self, grads_are_reduced: bool, scale: float):
        """One LAMB step of the group on gradients that are either already the mean on every rank
        (``grads_are_reduced``) or still per-rank sums to be reduced with ``scale``."""
        if self.group is not None:
            if grads_are_reduced:
                if not self.is_leader:
                    self._layout().flat_grad.zero_()   # the fused step reduces "leader's gradient + zeros"
                self.group.step(grad_scale=1.0)
            else:
                self.group.step(grad_scale=scale)
            return
        if not grads_are_reduced:
            self._reduce_group_gradients(scale)
        elif self.world > 1:
            for p in self.main_parameters:   # the leader's (possibly swarm-averaged) gradients are everybody's
                self._bcast(p.grad)
        self.opt.step(zero_grad=True)

    def _global_step(self, multi_peer: bool):
        t0 = time.perf_counter()
        scale = 1.0 / (self.world * max(self.times, 1))
        ctx = self.tracker.pause_updates() if self.is_leader else _null()
        with ctx:
            averaged = False
            if multi_peer:
                self._reduce_group_gradients(scale)
                if self.is_leader:
                    if self.fused:
                        torch.cuda.synchronize()
                    ctl, self._scheduled_grads = self._scheduled_grads, None
                    if ctl is not None and not (ctl.done() or ctl.cancelled):
                        averaged = self.grad_averager.finish_scheduled_step(
                            ctl, weight=float(self.samples), timeout=self.averaging_timeout + self.matchmaking_time) is not None
                        self.scheduled_rounds_used += averaged
                    if not averaged:
                        try:
                            res = self.grad_averager.step(weight=float(self.samples), reset_accumulators=False,
                                                          timeout=self.averaging_timeout,
                                                          target_group_size=self._group_size_hint(),
                                                          round_key=self.local_epoch)
                            averaged = res is not None
                        except Exception as e:
                            logger.warning(f"gradient averaging failed: {e!r}")
                    if averaged:
                        self.grad_averager.put_averaged_grads_into_params_()
                    elif self.verbose:
                        logger.info("proceeding with the group's own gradients")
                    self._ctl[0] = float(averaged)
                self._bcast(self._ctl)   # also orders "the leader's H2D copy is done" before the group step
                averaged = bool(self._ctl[0].item())
                with self.state_averager.lock_optimizer:
                    self._local_update(grads_are_reduced=True, scale=1.0)
            else:
                with self.state_averager.lock_optimizer:
                    self._local_update(grads_are_reduced=False, scale=scale)
            self.last_round_averaged = averaged
            if self.sched is not None:
                self.sched.step()
            new_epoch = max(self.local_epoch + 1, int(self._ctl[2].item()))
            self.state_averager.local_epoch = new_epoch
            self.state_averager.state_sharing_priority = new_epoch
            samples = self.samples
            self.samples = self.times = 0
            self.grad_averager.local_samples_accumulated = self.grad_averager.local_times_accumulated = 0
            if self.gather_state_every_epoch and multi_peer:
                self._gather_full_state()   # the leader serves state downloads from its own (now complete) copy
            if multi_peer and self.average_state_every and new_epoch % self.average_state_every == 0:
                self._start_state_round()
            if self.is_leader:
                self.tracker.update_epoch(new_epoch)
        self.callbacks.on_after_global_step.fire()
        if self.is_leader and self.verbose:
            logger.info(f"global step done: epoch {self.local_epoch}, {samples} group samples, "
                        f"{'swarm-averaged' if averaged else 'group-local'} gradients, {time.perf_counter() - t0:.2f}s")

    # ---- parameter averaging between peers ------------------------------------------------------
    def _start_state_round(self):
        """Leader: parameters -> host buffers -> swarm round (background thread under ``delay_optimizer_step``); the
        averaged parameters reach every rank in ``_apply_finished_state_round`` of a later ``step()``."""
        if self.is_leader:
            self._finish_pending_leader_only()
            self.state_averager._params_to_averager()
            self._pending_state_round = self._pool.submit(self._state_round)
        if not self.delay_optimizer_step:
            if self.is_leader:
                self._pending_state_round.result()
            self._apply_finished_state_round()

    def _state_round(self) -> bool:
        try:
            out = DecentralizedAverager.step(self.state_averager, weight=1.0, timeout=self.averaging_timeout,
                                             gather={"epoch": self.local_epoch},
                                             target_group_size=self._group_size_hint())
            return out is not None
        except Exception as e:
            logger.warning(f"state averaging failed: {e!r}")
            return False

    def _finish_pending_leader_only(self):
        fut = self._pending_state_round
        if fut is not None:
            try:
                fut.result(self.averaging_timeout)
            except Exception:
                pass
            self._pending_state_round = None   # a result nobody applied is dropped (the next round supersedes it)

    def _apply_finished_state_round(self):
        """Collective: the leader's finished round (if it succeeded) becomes everybody's parameters."""
        ok = False
        if self.is_leader and self._pending_state_round is not None:
            try:
                ok = bool(self._pending_state_round.result(self.averaging_timeout))
            except Exception:
                ok = False
            self._pending_state_round = None
            if ok:
                with self.state_averager.lock_optimizer:
                    self.state_averager._averager_to_params()
        t = torch.tensor([float(ok)], device=self.device)
        self._bcast(t)
        if bool(t.item()):
            self._broadcast_parameters()

    def _finish_pending(self):
        """Collective: wait for and apply a state round that is still in flight."""
        t = torch.tensor([float(self.is_leader and self._pending_state_round is not None)], device=self.device)
        self._bcast(t)
        if bool(t.item()):
            self._apply_finished_state_round()

    def shutdown(self):
        if self._scheduled_grads is not None:
            self._scheduled_grads.cancel()
            self._scheduled_grads = None
        self._finish_pending()
        if self.is_leader:
            self.tracker.shutdown()
            self.grad_averager.shutdown()
            self.state_averager.shutdown()
        self._pool.shutdown(wait=False)


GroupCollaborativeOptimizer = GroupOptimizer   # earlier name


class _null:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False
