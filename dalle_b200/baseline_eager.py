"""Baseline B1 of BASELINE.md: the SAME model stated naively in eager PyTorch (ATen / cuBLAS kernels, bf16
activations, python autograd with the reversible re-computation of dalle-pytorch) + the unfused per-tensor LAMB-8bit
written with torch ops.  This is the closest runnable stand-in for "the reference's own model on cuBLAS": the
reference itself cannot be installed offline.  ``bench.py --impl eager`` times it.  For N > 1 the
gradients are all-reduced with NCCL (DDP-style, one bucket)."""
from __future__ import annotations

import torch
import torch.distributed as dist

from .config import DalleConfig
from .models.dalle import DALLE
from .optim.lamb8bit import LAMB8Bit, get_linear_schedule_with_warmup
from .trainer import param_groups


class EagerTrainer:
    averaging = "nccl-allreduce"

    def __init__(self, cfg: DalleConfig, device, seed: int = 0, **kw):
        self.cfg, self.device = cfg, torch.device(device)
        self.world = dist.get_world_size() if dist.is_initialized() else 1
        self.model = DALLE(cfg, seed=seed).to(self.device)
        self.opt = LAMB8Bit(param_groups(self.model, 0.045), lr=2.5e-3, betas=(0.9, 0.96), eps=1e-6,
                            weight_decay=0.045, max_grad_norm=4.0, clamp_value=10000.0)
        self.sched = get_linear_schedule_with_warmup(self.opt, 3125, 31250)

    def step(self, text, image):
        loss = self.model.forward_reference(text, image, return_loss=True, build="sparse", dtype=torch.bfloat16,
                                            reversible_autograd=True)
        loss.backward()
        if self.world > 1:
            for p in self.model.parameters():
                if p.grad is not None:
                    dist.all_reduce(p.grad)
                    p.grad.div_(self.world)
        self.opt._step_reference()          # torch-op LAMB-8bit (per tensor), like the reference's CPU loop
        for p in self.model.parameters():
            p.grad = None
        self.sched.step()
        return loss.detach()

    def step_from_host(self, text_cpu, image_cpu) -> float:
        return float(self.step(text_cpu.to(self.device, non_blocking=True),
                               image_cpu.to(self.device, non_blocking=True)).item())
