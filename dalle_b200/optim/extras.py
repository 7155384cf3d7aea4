"""Alternative / auxiliary optimizers of the reference (lib/training/clipped_lamb.py, offload.py, wrapper.py; SURVEY
R14-R16).  None of them is on the reference's default path (``CPULAMB8Bit`` is), but users can select them.

* ``Lamb`` / ``LambWithGradientClipping`` - exact-state (fp32) LAMB with global-norm clipping.
* ``OptimizerWrapper``   - pure forwarding base class.
* ``OffloadOptimizer``   - keeps the optimizer's parameters, gradients and state on another device (CPU by default)
  and runs the update there: "optimizer state + update off the training device".  On B200 the fused GPU LAMB makes
  this unnecessary, but the capability is kept for memory-tight or CPU-optimizer experiments.
"""
from __future__ import annotations

import threading
from typing import Iterable, Optional

import torch


class Lamb(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-6, weight_decay=0.0, clamp_value=10.0,
                 debias=False):
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay))
        self.clamp_value, self.debias = clamp_value, debias

    @torch.no_grad()
    def step(self, closure=None):
        loss = closure() if closure is not None else None
        for group in self.param_groups:
            b1, b2 = group["betas"]
            for p in group["params"]:
                if p.grad is None:
                    continue
                st = self.state[p]
                if not st:
                    st["step"], st["exp_avg"], st["exp_avg_sq"] = 0, torch.zeros_like(p), torch.zeros_like(p)
                st["step"] += 1
                st["exp_avg"].mul_(b1).add_(p.grad, alpha=1 - b1)
                st["exp_avg_sq"].mul_(b2).addcmul_(p.grad, p.grad, value=1 - b2)
                bc = (1 - b2 ** st["step"]) ** 0.5 / (1 - b1 ** st["step"]) if self.debias else 1.0
                upd = st["exp_avg"] / (st["exp_avg_sq"].sqrt() + group["eps"])
                if group["weight_decay"]:
                    upd.add_(p, alpha=group["weight_decay"])
                wn, un = p.norm().clamp(0, self.clamp_value), upd.norm()
                trust = float(wn / un) if wn != 0 and un != 0 else 1.0
                p.add_(upd, alpha=-group["lr"] * bc * trust)
        return loss


class LambWithGradientClipping(Lamb):
    """LAMB + ``clip_grad_norm_`` over all parameters before every step (reference clipped_lamb.py:5-14)."""

    def __init__(self, *args, max_grad_norm: Optional[float] = None, **kwargs):
        self.max_grad_norm = max_grad_norm
        super().__init__(*args, **kwargs)

    def step(self, *args, **kwargs):
        if self.max_grad_norm is not None:
            torch.nn.utils.clip_grad_norm_([p for g in self.param_groups for p in g["params"]], self.max_grad_norm)
        return super().step(*args, **kwargs)


class OptimizerWrapper(torch.optim.Optimizer):
    """Forwards everything to the wrapped optimizer (reference wrapper.py:4-47)."""

    def __init__(self, optim: torch.optim.Optimizer):
        object.__setattr__(self, "optim", optim)

    @property
    def defaults(self):
        return self.optim.defaults

    @property
    def state(self):
        return self.optim.state

    @property
    def param_groups(self):
        return self.optim.param_groups

    @param_groups.setter
    def param_groups(self, value):
        self.optim.param_groups = value

    def __getstate__(self):
        return self.optim.__getstate__()

    def __setstate__(self, state):
        self.optim.__setstate__(state)

    def __repr__(self):
        return f"{self.__class__.__name__}({self.optim!r})"

    def state_dict(self):
        return self.optim.state_dict()

    def load_state_dict(self, state_dict):
        return self.optim.load_state_dict(state_dict)

    def step(self, *args, **kwargs):
        return self.optim.step(*args, **kwargs)

    def zero_grad(self, *args, **kwargs):
        return self.optim.zero_grad(*args, **kwargs)

    def add_param_group(self, param_group):
        return self.optim.add_param_group(param_group)


class OffloadOptimizer(OptimizerWrapper):
    """Runs ``optim_cls`` on copies of the parameters that live on ``offload_device`` (reference offload.py:10-93)."""

    def __init__(self, param_groups, optim_cls, *args, full_sync: bool = True, offload_device="cpu", **kwargs):
        param_groups = list(param_groups)
        if not isinstance(param_groups[0], dict):
            param_groups = [{"params": param_groups}]
        self.full_sync, self.lock = full_sync, threading.Lock()
        self._main = [[p for p in g["params"]] for g in param_groups]
        off_groups = []
        for g in param_groups:
            copies = [torch.nn.Parameter(p.detach().to(offload_device, copy=True), requires_grad=p.requires_grad)
                      for p in g["params"]]
            for c in copies:
                c.grad = torch.zeros_like(c)
            off_groups.append(dict(g, params=copies))
        self._off = [g["params"] for g in off_groups]
        super().__init__(optim_cls(off_groups, *args, **kwargs))

    @torch.no_grad()
    def step(self, closure=None, *args, **kwargs):
        assert closure is None, "closures are not supported with an offloaded optimizer"
        with self.lock:
            for mg, og in zip(self._main, self._off):
                for p, c in zip(mg, og):
                    if self.full_sync:
                        c.copy_(p, non_blocking=True)
                    if p.grad is not None:
                        c.grad.copy_(p.grad, non_blocking=True)
                    else:
                        c.grad.zero_()
            out = self.optim.step(*args, **kwargs)
            for mg, og in zip(self._main, self._off):
                for p, c in zip(mg, og):
                    p.copy_(c, non_blocking=True)
            return out

    def zero_grad(self, set_to_none: bool = False, *args, **kwargs):
        with self.lock:
            for mg in self._main:
                for p in mg:
                    if p.grad is not None:
                        p.grad = None if set_to_none else p.grad.zero_()
