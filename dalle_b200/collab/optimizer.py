"""``Optimizer``: target-batch-size collaborative optimizer with the hivemind.Optimizer surface the reference pokes at
(constructed at reference task.py:127-134).

Every ``step()`` adds ``batch_size_per_step`` samples to the local counter and reports progress to the DHT; once the
swarm has collectively accumulated ``target_batch_size`` samples (or this peer fell behind) one *global step* runs:
gradients are averaged with the other peers (weighted by samples, compressed), the wrapped optimizer steps once,
the scheduler steps once, ``local_epoch += 1``, parameters are averaged (``average_state_every``), accumulators are
reset.  Averaging failures fall back to the local gradients.  A peer more than one epoch behind downloads the full
state from the best donor instead.

B200 placement: the wrapped optimizer is the fused LAMB on the GPU (or the CPU reference path for CPU models), the
averaged-gradient buffers are pinned host memory, and inside a multi-GPU box only the group leader talks to the
swarm (parallel/group.py reduces over NVLink first).  ``delay_optimizer_step`` keeps its meaning of hiding the slow
part - the parameter averaging round runs on a background thread and is applied at the next ``step()``.
"""
from __future__ import annotations

import logging
import os
import threading
import time
from concurrent.futures import Future, ThreadPoolExecutor
from typing import Callable, Dict, Iterable, List, Optional, Sequence, Union

import torch

from ..dht import DHT
from . import compression as C
from ..dht import get_dht_time
from .averager import DecentralizedAverager, StepControl
from .progress_tracker import ProgressTracker

logger = logging.getLogger(__name__)


class _Callbacks:
    """``opt.callbacks.on_after_global_step.add(fn)`` / ``on_load_state_from_peers.add(fn)`` (run_trainer_tpu.py:66)."""

    class _Hook:
        def __init__(self):
            self._fns: List[Callable] = []

        def add(self, fn):
            self._fns.append(fn)
            return fn

        def __call__(self, fn=None):
            # the reference calls ``on_load_state_from_peers(fn)`` instead of ``.add(fn)`` (a latent bug): accept it
            if fn is not None:
                return self.add(fn)
            for f in self._fns:
                f()

        def fire(self):
            for f in self._fns:
                f()

    def __init__(self):
        self.on_after_global_step = self._Hook()
        self.on_load_state_from_peers = self._Hook()


class GradientAverager(DecentralizedAverager):
    """Accumulates gradients over micro-steps and averages them across peers."""

    def __init__(self, parameters: Iterable[torch.nn.Parameter], *, dht: DHT, prefix: str,
                 reuse_grad_buffers: bool = False, **kwargs):
        self.parameters = [p for p in parameters]
        self.reuse_grad_buffers = reuse_grad_buffers
        self.local_samples_accumulated = 0
        self.local_times_accumulated = 0
        self._new_averaged_grads = False
        if not reuse_grad_buffers:
            self._accumulators = [torch.zeros_like(p, dtype=torch.float32) for p in self.parameters]
        super().__init__([tuple(p.shape) for p in self.parameters], dht, prefix=prefix,
                         pin_memory=any(p.is_cuda for p in self.parameters), **kwargs)

    def _grads(self):
        for p in self.parameters:
            yield p.grad if p.grad is not None else torch.zeros_like(p)

    @torch.no_grad()
    def accumulate_grads_(self, batch_size: int):
        self.local_samples_accumulated += batch_size
        self.local_times_accumulated += 1
        if not self.reuse_grad_buffers:
            for acc, g in zip(self._accumulators, self._grads()):
                acc.add_(g.to(acc.dtype))

    @torch.no_grad()
    def load_accumulators_into_averager_(self):
        src = self._grads() if self.reuse_grad_buffers else iter(self._accumulators)
        scale = 1.0 / max(self.local_times_accumulated, 1)
        with self.get_tensors() as avg:
            for a, g in zip(avg, src):
                a.copy_(g.detach().to("cpu", torch.float32, non_blocking=False)).mul_(scale)

    @torch.no_grad()
    def reset_accumulated_grads_(self):
        self.local_samples_accumulated = self.local_times_accumulated = 0
        if self.reuse_grad_buffers:
            for p in self.parameters:
                if p.grad is not None:
                    p.grad.zero_()
        else:
            for a in self._accumulators:
                a.zero_()

    def step(self, weight: Optional[float] = None, reset_accumulators: bool = True, timeout=None, **kwargs):
        self.load_accumulators_into_averager_()
        weight = float(self.local_samples_accumulated) if weight is None else weight
        out = super().step(weight=weight, timeout=timeout, gather={"samples": self.local_samples_accumulated}, **kwargs)
        self._new_averaged_grads = out is not None
        if reset_accumulators:
            self.reset_accumulated_grads_()
        return out

    def schedule_step(self, scheduled_time=None, timeout=None, **kwargs) -> StepControl:
        """Matchmaking now, all-reduce when ``finish_scheduled_step`` releases it (the accumulators keep filling)."""
        return super().schedule_step(scheduled_time=scheduled_time, timeout=timeout,
                                     gather={"samples": self.local_samples_accumulated}, **kwargs)

    def finish_scheduled_step(self, control: StepControl, weight: Optional[float] = None, timeout=None):
        self.load_accumulators_into_averager_()
        control.weight = float(self.local_samples_accumulated) if weight is None else float(weight)
        control.allow_allreduce()
        try:
            out = control.result(timeout)
        except Exception:
            control.cancel()
            out = None
        self._new_averaged_grads = out is not None
        return out

    @torch.no_grad()
    def put_averaged_grads_into_params_(self):
        """Make the (averaged or local) gradients the ``.grad`` the wrapped optimizer will see."""
        with self.get_tensors() as avg:
            for p, a in zip(self.parameters, avg):
                if p.grad is None:
                    p.grad = torch.zeros_like(p)
                p.grad.copy_(a.to(p.grad.device, non_blocking=True))


class TrainingStateAverager(DecentralizedAverager):
    """Owns the wrapped optimizer + scheduler + ``local_epoch``; averages parameters; serves / loads full state
    (SURVEY H3)."""

    def __init__(self, *, dht: DHT, optimizer, params, scheduler=None, prefix: str, **kwargs):
        self.main_parameters = [p for g in (params if isinstance(params, (list, tuple)) and params and isinstance(params[0], dict)
                                            else [{"params": list(params)}]) for p in g["params"]]
        self.optimizer = optimizer(params) if callable(optimizer) and not isinstance(optimizer, torch.optim.Optimizer) else optimizer
        self.scheduler = scheduler(self.optimizer) if callable(scheduler) and scheduler is not None and not hasattr(scheduler, "step") else scheduler
        self.local_epoch = 0
        self.lock_optimizer = threading.Lock()
        super().__init__([tuple(p.shape) for p in self.main_parameters], dht, prefix=prefix,
                         pin_memory=any(p.is_cuda for p in self.main_parameters), **kwargs)

    @torch.no_grad()
    def _params_to_averager(self):
        with self.get_tensors() as avg:
            for a, p in zip(avg, self.main_parameters):
                a.copy_(p.detach().to("cpu", torch.float32))

    @torch.no_grad()
    def _averager_to_params(self):
        with self.get_tensors() as avg:
            for a, p in zip(avg, self.main_parameters):
                p.copy_(a.to(p.device, non_blocking=True))

    def average_parameters(self, timeout=None, **kwargs) -> bool:
        self._params_to_averager()
        out = DecentralizedAverager.step(self, weight=1.0, timeout=timeout, gather={"epoch": self.local_epoch}, **kwargs)
        if out is not None:
            self._averager_to_params()
        return out is not None

    # ---- state sharing --------------------------------------------------------------------------
    def get_current_state(self):
        with self.lock_optimizer:
            sd = self.optimizer.state_dict()
            tensors, opt_meta = [p.detach().cpu().clone() for p in self.main_parameters], {}
            flat_state = []
            for idx, st in sd["state"].items():
                for k, v in st.items():
                    if isinstance(v, torch.Tensor) and v.dim() > 0:
                        flat_state.append((idx, k))
                        tensors.append(v.detach().cpu().clone())
                    else:
                        opt_meta.setdefault(idx, {})[k] = float(v) if isinstance(v, torch.Tensor) else v
            meta = {"epoch": self.local_epoch, "opt_keys": [[i, k] for i, k in flat_state], "opt_meta": opt_meta,
                    "param_groups": [{k: v for k, v in g.items() if k != "params"} for g in sd["param_groups"]],
                    "n_params": len(self.main_parameters),
                    "scheduler": self.scheduler.state_dict() if self.scheduler is not None else None}
        return meta, tensors

    @torch.no_grad()
    def load_state_from_peers(self, timeout: Optional[float] = None) -> bool:
        res = DecentralizedAverager.load_state_from_peers(self, wait=True, timeout=timeout)
        if res is None:
            return False
        meta, tensors = res
        if meta.get("epoch", -1) < self.local_epoch:
            return False
        n = meta["n_params"]
        if n != len(self.main_parameters):
            logger.warning("state from peers has a different parameter count; ignored")
            return False
        with self.lock_optimizer:
            for p, t in zip(self.main_parameters, tensors[:n]):
                p.copy_(t.view_as(p).to(p.device))
            sd = self.optimizer.state_dict()
            for (idx, k), t in zip(meta["opt_keys"], tensors[n:]):
                sd["state"].setdefault(idx, {})[k] = t
            for idx, kv in meta["opt_meta"].items():
                sd["state"].setdefault(int(idx), {}).update(kv)
            try:
                self.optimizer.load_state_dict(sd)
            except Exception as e:
                logger.warning(f"could not load optimizer state from peers: {e}")
            if self.scheduler is not None and meta.get("scheduler") is not None:
                self.scheduler.load_state_dict(meta["scheduler"])
            self.local_epoch = meta["epoch"]
        return True


class Optimizer:
    """See module docstring.  Constructor keywords = reference task.py:127-134 (+ CollaborativeArguments)."""

    def __init__(self, *, dht: DHT, run_id: str, target_batch_size: int, batch_size_per_step: Optional[int] = None,
                 optimizer, params=None, scheduler=None, matchmaking_time: float = 15.0, averaging_timeout: float = 60.0,
                 allreduce_timeout: Optional[float] = None, load_state_timeout: float = 600.0,
                 reuse_grad_buffers: bool = False, offload_optimizer: bool = False, delay_optimizer_step: bool = False,
                 delay_grad_averaging: bool = False, delay_state_averaging: bool = True, average_state_every: int = 1,
                 client_mode: Optional[bool] = None, auxiliary: bool = False,
                 grad_compression: C.CompressionBase = C.NoCompression(),
                 state_averaging_compression: C.CompressionBase = C.NoCompression(),
                 load_state_compression: C.CompressionBase = C.NoCompression(), averager_opts: Optional[dict] = None,
                 tracker_opts: Optional[dict] = None, verbose: bool = False, target_group_size: Optional[int] = None,
                 on_before_local_optimizer_step: Optional[Callable] = None):
        self.dht, self.run_id = dht, run_id
        self.target_batch_size, self.batch_size_per_step = target_batch_size, batch_size_per_step
        self.matchmaking_time, self.averaging_timeout = matchmaking_time, averaging_timeout
        self.load_state_timeout, self.average_state_every = load_state_timeout, average_state_every
        self.delay_optimizer_step, self.delay_grad_averaging = delay_optimizer_step, delay_grad_averaging
        self.offload_optimizer = offload_optimizer
        self.client_mode = dht.client_mode if client_mode is None else client_mode
        self.auxiliary, self.verbose = auxiliary, verbose
        self.callbacks = _Callbacks()
        self._target_group_size = target_group_size
        self._before_local_step = on_before_local_optimizer_step
        opts = dict(min_matchmaking_time=matchmaking_time, allreduce_timeout=allreduce_timeout,
                    client_mode=self.client_mode, auxiliary=auxiliary, **(averager_opts or {}))
        self.state_averager = TrainingStateAverager(
            dht=dht, optimizer=optimizer, params=params, scheduler=scheduler, prefix=f"{run_id}_state_averager",
            compression=state_averaging_compression, state_compression=load_state_compression, **opts)
        self.grad_averager = GradientAverager(
            self.state_averager.main_parameters, dht=dht, prefix=f"{run_id}_grad_averager",
            reuse_grad_buffers=reuse_grad_buffers, compression=grad_compression, allow_state_sharing=False, **opts)
        self.tracker = ProgressTracker(dht, run_id, target_batch_size, client_mode=self.client_mode,
                                       **(tracker_opts or {}))
        self._pool = ThreadPoolExecutor(max_workers=1)
        self._pending_state_round: Optional[Future] = None
        self._schema_hash = None
        # pre-scheduled gradient round: matchmaking starts `matchmaking_time` before the tracker's ETA of the epoch
        self._scheduled_grads: Optional[StepControl] = None
        self.scheduled_rounds_used = 0

    # ---- surface used by the reference ----------------------------------------------------------
    @property
    def local_epoch(self) -> int:
        return self.state_averager.local_epoch

    @property
    def param_groups(self):
        return self.state_averager.optimizer.param_groups

    @property
    def scheduler(self):
        return self.state_averager.scheduler

    @property
    def opt(self):
        return self.state_averager.optimizer

    def is_synchronized_with_peers(self) -> bool:
        return self.local_epoch >= self.tracker.global_epoch - (1 if self.delay_optimizer_step else 0) or \
            self.local_epoch >= self.tracker.global_epoch - 1

    def zero_grad(self, set_to_none: bool = False):
        if self.grad_averager.reuse_grad_buffers:
            raise ValueError("with reuse_grad_buffers=True the optimizer owns .grad: do not call zero_grad "
                             "(gradients are reset after each global step)")
        for p in self.state_averager.main_parameters:
            if p.grad is not None:
                p.grad = None if set_to_none else p.grad.zero_()

    def state_dict(self) -> dict:
        sd = self.state_averager.optimizer.state_dict()
        sd["state"]["local_epoch"] = self.local_epoch
        return sd

    def load_state_dict(self, state_dict: dict):
        state_dict = dict(state_dict, state=dict(state_dict["state"]))
        if "local_epoch" in state_dict["state"]:
            self.state_averager.local_epoch = state_dict["state"].pop("local_epoch")
        return self.state_averager.optimizer.load_state_dict(state_dict)

    def load_state_from_peers(self, **kwargs):
        self._finish_pending()
        ok = False
        with self.tracker.pause_updates():
            try:
                ok = self.state_averager.load_state_from_peers(timeout=self.load_state_timeout)
            except Exception as e:
                logger.warning(f"load_state_from_peers failed: {e!r}")
            if ok:
                self.grad_averager.reset_accumulated_grads_()
                self.tracker.report_local_progress(self.local_epoch, 0)
                self.callbacks.on_load_state_from_peers.fire()
        return ok

    # ---- the step -------------------------------------------------------------------------------
    def step(self, closure=None, batch_size: Optional[int] = None, grad_scaler=None):
        if self.auxiliary:
            return self._auxiliary_step()
        batch_size = batch_size if batch_size is not None else self.batch_size_per_step
        assert batch_size is not None, "pass batch_size or set batch_size_per_step"
        loss = closure() if closure is not None else None
        self._apply_finished_state_round()
        if grad_scaler is not None and not grad_scaler.are_grads_finite(self):
            logger.info("encountered non-finite gradients, skipping this micro-step")
            return loss
        if not self.is_synchronized_with_peers():
            logger.info(f"peer is out of sync (local epoch {self.local_epoch}, global {self.tracker.global_epoch}): "
                        f"downloading state")
            if self._scheduled_grads is not None:
                self._scheduled_grads.cancel()
                self._scheduled_grads = None
            self.load_state_from_peers()
            return loss
        self.grad_averager.accumulate_grads_(batch_size)
        self.tracker.report_local_progress(self.local_epoch, self.grad_averager.local_samples_accumulated)
        self._maybe_schedule_gradient_averaging()
        if self.tracker.ready_to_update_epoch:
            self._update_global_epoch(grad_scaler)
        return loss

    def _maybe_schedule_gradient_averaging(self):
        """hivemind.Optimizer._maybe_schedule_gradient_averaging: once the tracker expects the swarm to reach the target
        batch within ``matchmaking_time`` seconds, start looking for a group in the background so that the group is
        there when the last micro-batch is (reference arguments.py:66-68)."""
        ctl = self._scheduled_grads
        if ctl is not None and (ctl.done() or ctl.cancelled):
            self._scheduled_grads = ctl = None       # failed or gave up before its epoch came: schedule again
        if ctl is not None or not (self.tracker.global_progress.num_peers > 1 or self._target_group_size):
            return
        logger.debug(f"tracker: {self.tracker.global_progress}")
        eta = self.tracker.estimated_next_update_time
        now = get_dht_time()
        if eta - now <= self.matchmaking_time:
            # (never a shorter window than the un-scheduled path gives: peers notice the epoch at different moments)
            self._scheduled_grads = self.grad_averager.schedule_step(
                scheduled_time=max(eta, now + self.grad_averager.min_matchmaking_time),
                timeout=self.averaging_timeout + self.matchmaking_time,
                target_group_size=self._group_size_hint(), round_key=self.local_epoch)
            if self.verbose:
                logger.info(f"pre-scheduled gradient averaging, group due in {max(eta - now, 0.0):.1f}s")

    def _group_size_hint(self) -> Optional[int]:
        if self._target_group_size:
            return self._target_group_size
        n = self.tracker.global_progress.num_peers
        return n if n >= 2 else None

    def _update_global_epoch(self, grad_scaler=None):
        t0 = time.perf_counter()
        with self.tracker.pause_updates():
            samples = self.grad_averager.local_samples_accumulated
            averaged = None
            ctl, self._scheduled_grads = self._scheduled_grads, None
            if ctl is not None and (ctl.done() or ctl.cancelled):
                ctl = None                           # its matchmaking already ended without a group
            if ctl is not None:
                averaged = self.grad_averager.finish_scheduled_step(ctl, weight=float(samples),
                                                                    timeout=self.averaging_timeout + self.matchmaking_time)
                self.scheduled_rounds_used += averaged is not None
                if averaged is None:
                    logger.warning("the pre-scheduled gradient round found no group; looking again")
            if averaged is None and (self.tracker.global_progress.num_peers > 1 or self._target_group_size):
                try:
                    averaged = self.grad_averager.step(weight=float(samples), reset_accumulators=False,
                                                       timeout=self.averaging_timeout,
                                                       target_group_size=self._group_size_hint(),
                                                       round_key=self.local_epoch)
                except Exception as e:
                    logger.warning(f"gradient averaging failed: {e!r}")
            elif averaged is None:
                self.grad_averager.load_accumulators_into_averager_()
            if averaged is None and self.verbose:
                logger.info(f"proceeding with local gradients (the tracker sees {self.tracker.global_progress.num_peers} "
                            f"peer(s), global epoch {self.tracker.global_epoch}, local epoch {self.local_epoch})")
            # averaged (or local, normalised by micro-steps) gradients -> .grad -> wrapped optimizer
            self.grad_averager.put_averaged_grads_into_params_()
            if self._before_local_step is not None:
                self._before_local_step()
            with self.state_averager.lock_optimizer:
                if grad_scaler is not None:
                    grad_scaler.step_wrapped(self.state_averager.optimizer)
                else:
                    self.state_averager.optimizer.step()
                if self.state_averager.scheduler is not None:
                    self.state_averager.scheduler.step()
                self.state_averager.local_epoch = max(self.local_epoch + 1, self.tracker.global_epoch)
            self.grad_averager.reset_accumulated_grads_()
            self.state_averager.state_sharing_priority = self.local_epoch
            new_epoch = self.local_epoch
            should_average_state = (self.average_state_every and new_epoch % self.average_state_every == 0
                                    and (self.tracker.global_progress.num_peers > 1 or self._target_group_size))
            if should_average_state:
                if self.delay_optimizer_step:
                    self._pending_state_round = self._pool.submit(self._state_round)
                else:
                    self._state_round()
                    self._apply_finished_state_round(force=True)
            self.tracker.update_epoch(new_epoch)
        self.callbacks.on_after_global_step.fire()
        if self.verbose:
            logger.info(f"global step done: epoch {self.local_epoch}, {samples} local samples, "
                        f"{'averaged' if averaged else 'local'} gradients, {time.perf_counter() - t0:.2f}s")

    def _state_round(self) -> bool:
        try:
            self.state_averager._params_to_averager()
            out = DecentralizedAverager.step(self.state_averager, weight=1.0, timeout=self.averaging_timeout,
                                             gather={"epoch": self.local_epoch},
                                             target_group_size=self._group_size_hint())
            return out is not None
        except Exception as e:
            logger.warning(f"state averaging failed: {e!r}")
            return False

    def _apply_finished_state_round(self, force: bool = False):
        fut = self._pending_state_round
        if fut is None and not force:
            return
        if fut is not None:
            if not fut.done():
                return
            ok = fut.result()
            self._pending_state_round = None
        else:
            ok = True
        if ok:
            with self.state_averager.lock_optimizer:
                self.state_averager._averager_to_params()

    def _finish_pending(self):
        if self._pending_state_round is not None:
            try:
                self._pending_state_round.result(self.averaging_timeout)
            except Exception:
                pass
            self._apply_finished_state_round()

    def _auxiliary_step(self):
        """Aux peers help averaging with weight 0 (reference run_aux_peer.py:79-82, left unimplemented there)."""
        self.tracker.report_local_progress(self.tracker.global_epoch, 0)
        if self.tracker.global_progress.num_peers > 1 and self.tracker.ready_to_update_epoch is not None:
            self.grad_averager.step(weight=0.0, timeout=self.matchmaking_time + 1.0, allow_retries=False)
            DecentralizedAverager.step(self.state_averager, weight=0.0, timeout=self.matchmaking_time + 1.0,
                                       allow_retries=False)

    def shutdown(self):
        if self._scheduled_grads is not None:
            self._scheduled_grads.cancel()
        self._finish_pending()
        self.tracker.shutdown()
        self.grad_averager.shutdown()
        self.state_averager.shutdown()
        self._pool.shutdown(wait=False)


CollaborativeOptimizer = Optimizer  # legacy name imported for typing at reference hf_trainer.py:6


class HivemindGradScaler(torch.amp.GradScaler):
    """AMP grad scaler whose unscale / step / update happen only at global steps, so gradients accumulated over
    micro-steps stay consistently scaled (reference hf_trainer.py:26-28)."""

    def __init__(self, *args, **kwargs):
        kwargs.setdefault("device", "cuda" if torch.cuda.is_available() else "cpu")
        super().__init__(*args, **kwargs)
        self._is_running_global_step = False
        self._optimizer_states_to_reset = set()

    def unscale_(self, optimizer) -> bool:
        if isinstance(optimizer, Optimizer) and not self._is_running_global_step:
            return False  # deferred
        super().unscale_(optimizer.opt if isinstance(optimizer, Optimizer) else optimizer)
        return True

    def step(self, optimizer, *args, **kwargs):
        if isinstance(optimizer, Optimizer):
            return optimizer.step(*args, grad_scaler=self, **kwargs)
        return super().step(optimizer, *args, **kwargs)

    def step_wrapped(self, inner):
        self._is_running_global_step = True
        try:
            if self.is_enabled():
                super().unscale_(inner)
                super().step(inner)
                super().update()
            else:
                inner.step()
        finally:
            self._is_running_global_step = False

    def update(self, new_scale=None):
        if self._is_running_global_step or not self.is_enabled():
            return super().update(new_scale) if self.is_enabled() else None
        return None  # deferred to the global step

    def are_grads_finite(self, optimizer: Optimizer) -> bool:
        params = optimizer.state_averager.main_parameters
        return all(p.grad is None or bool(torch.isfinite(p.grad).all()) for p in params)
