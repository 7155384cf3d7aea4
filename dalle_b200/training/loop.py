"""The training loop (replaces ``CollaborativeHFTrainer`` = HF ``Trainer`` + the hacks of reference
lib/training/hf_trainer.py:15-85).

What the reference obtains by monkey-patching HF Trainer is stated directly here:
* the collaborative optimizer owns stepping and the LR schedule (``NoOpScheduler``);
* gradients accumulate in ``.grad`` across micro-steps and are NOT zeroed by the loop (``IgnoreGradManipulations``);
  a micro-batch that produced non-finite gradients is dropped by zeroing them (hf_trainer.py:73-78);
* no clipping in the loop - the optimizer clips by global norm (hf_trainer.py:80-85);
* each peer shuffles with its own seed (hf_trainer.py:30-33).
"""
from __future__ import annotations

import logging
import time
from typing import Iterable, Optional

import torch

logger = logging.getLogger(__name__)


class NoOpScheduler:
    """Placeholder scheduler object for code that expects one; the real schedule lives in the optimizer."""

    def __init__(self, optimizer):
        self.optimizer = optimizer
        self._last_lr = self.get_lr()

    def get_lr(self):
        return [g["lr"] for g in self.optimizer.param_groups]

    def get_last_lr(self):
        return self._last_lr

    def step(self):
        self._last_lr = self.get_lr()

    def state_dict(self):
        return {}

    def load_state_dict(self, *a, **k):
        pass


def grads_are_finite(model) -> bool:
    p0 = next(model.parameters())
    layout = getattr(p0, "_flat_layout", None)
    if layout is not None:
        return bool(torch.isfinite(layout.flat_grad).all())
    return all(p.grad is None or bool(torch.isfinite(p.grad).all()) for p in model.parameters())


def _failpoint_nan(model, micro: int):
    """Fault injection: ``DALLE_B200_INJECT_NAN="<rank>:<micro step>"`` poisons one parameter of that
    rank after that micro-step, so tests can watch the NaN guard roll the peer (all ranks of a group) back."""
    import os

    spec = os.environ.get("DALLE_B200_INJECT_NAN")
    if not spec:
        return
    rank_s, step_s = spec.split(":")
    my_rank = int(os.environ.get("RANK", "0"))
    if int(rank_s) == my_rank and int(step_s) == micro:
        with torch.no_grad():
            next(model.parameters()).view(-1)[0] = float("nan")
        logger.warning(f"failpoint: injected NaN into the parameters of rank {my_rank} after micro-step {micro}")


class CollaborativeTrainer:
    def __init__(self, *, model, args, data_collator, train_dataset, collaborative_optimizer, callbacks=(),
                 data_seed: int = 0, max_micro_steps: Optional[int] = None, check_finite_every_step: bool = True):
        self.model, self.args = model, args
        self.opt, self.callbacks = collaborative_optimizer, list(callbacks)
        self.collator, self.dataset, self.data_seed = data_collator, train_dataset, data_seed
        self.lr_scheduler = NoOpScheduler(collaborative_optimizer)
        self.max_micro_steps = max_micro_steps
        self.check_finite = check_finite_every_step
        self.device = next(model.parameters()).device

    def get_train_dataloader(self):
        torch.manual_seed(self.data_seed % 2 ** 31)
        return torch.utils.data.DataLoader(self.dataset, batch_size=self.args.per_device_train_batch_size,
                                           collate_fn=self.collator,
                                           num_workers=0 if self.device.type == "cpu" else self.args.dataloader_num_workers,
                                           pin_memory=self.device.type == "cuda")

    def train(self, model_path=None):
        for cb in self.callbacks:
            cb.on_train_begin()
        accum = max(1, self.args.gradient_accumulation_steps)
        micro, t_log, loss_log = 0, time.time(), 0.0
        inner = getattr(self.model, "model", self.model)
        if self.device.type == "cuda" and hasattr(inner, "engine"):
            inner.engine.expected_grad_out = 1.0 / accum   # this loop always calls (loss / accum).backward()
        for batch in self.get_train_dataloader():
            batch = {k: v.to(self.device, non_blocking=True) for k, v in batch.items()}
            loss = self.model(**batch)["loss"]
            (loss / accum).backward()
            micro += 1
            if micro % accum == 0:
                finite = (not self.check_finite) or grads_are_finite(self.model)
                if self.check_finite and hasattr(self.opt, "all_ranks_agree"):
                    finite = self.opt.all_ranks_agree(finite)   # a multi-GPU peer drops a micro-batch on every rank or none
                if not finite:
                    logger.warning("non-finite gradients: dropping this micro-batch")
                    for p in self.model.parameters():
                        if p.grad is not None:
                            p.grad.zero_()
                else:
                    self.opt.step()
                self.lr_scheduler.step()
                lval = float(loss.detach())
                loss_log += lval
                for cb in self.callbacks:
                    cb.on_step_end(loss=lval)
                if self.args.logging_steps and (micro // accum) % self.args.logging_steps == 0:
                    n = self.args.logging_steps
                    logger.info(f"step {micro // accum}: loss {loss_log / n:.4f}, lr {self.lr_scheduler.get_lr()[0]:.3g}, "
                                f"{n * self.args.per_device_train_batch_size * accum / (time.time() - t_log):.1f} samples/s")
                    t_log, loss_log = time.time(), 0.0
            _failpoint_nan(self.model, micro)
            if self.max_micro_steps is not None and micro >= self.max_micro_steps:
                break
            if micro >= self.args.max_steps * accum:
                break
        return micro
