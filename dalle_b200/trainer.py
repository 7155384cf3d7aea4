"""Local (one box) training step: the public API ``bench.py`` and ``run_trainer.py`` drive.

``GroupTrainer.step(batch)`` = forward + backward through the fused executor on every rank's micro-batch, gradient
averaging across the ranks of the box, one LAMB-8bit update, gradients zeroed.  It replaces what the reference does
with ``nn.DataParallel`` inside HF ``Trainer`` plus the CPU ``CPULAMB8Bit`` step.

Averaging back-ends inside the box:
* ``"fused"`` - the NVLink peer-memory kernel (parallel/group.py): reduce-scatter + LAMB on the shard + all-gather
  in device code, no NCCL call on the path;
* ``"nccl"``  - one ``all_reduce`` over the flat gradient buffer followed by the single-GPU LAMB kernels
  (the baseline arm, also the CPU/gloo plumbing path).
"""
from __future__ import annotations

import os
from typing import Dict, Optional

import torch
import torch.distributed as dist

from .config import DalleConfig
from .models.dalle import DALLE
from .optim.lamb8bit import LAMB8Bit, get_linear_schedule_with_warmup


def param_groups(model: DALLE, weight_decay: float):
    """Group 0: everything whose reference name has neither "bias" nor "LayerNorm.weight" (the latter never matches
    dalle-pytorch names, so norm weights and LayerScale ARE decayed - reference task.py:138-150);
    group 1: biases, no decay.  Order = reference named_parameters() order (optimizer-state indices, App. A.2)."""
    no_decay = ("bias", "LayerNorm.weight")
    named = list(model.reference_named_parameters())
    return [
        {"params": [p for n, p in named if not any(nd in n for nd in no_decay)], "weight_decay": weight_decay},
        {"params": [p for n, p in named if any(nd in n for nd in no_decay)], "weight_decay": 0.0},
    ]


class GroupTrainer:
    def __init__(self, cfg: DalleConfig, device, *, lr=2.5e-3, betas=(0.9, 0.96), eps=1e-6, weight_decay=0.045,
                 max_grad_norm=4.0, clamp_value=10000.0, warmup_steps=3125, total_steps=31250, seed=0,
                 averaging: str = "auto", process_group=None, payload: Optional[str] = None):
        self.cfg, self.device = cfg, torch.device(device)
        self.world = dist.get_world_size(process_group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(process_group) if dist.is_initialized() else 0
        self.pg = process_group
        self.model = DALLE(cfg, seed=seed)  # same seed on every rank -> identical initial weights
        self.model.to(self.device)
        self.opt = LAMB8Bit(param_groups(self.model, weight_decay), lr=lr, betas=betas, eps=eps,
                            weight_decay=weight_decay, max_grad_norm=max_grad_norm, clamp_value=clamp_value,
                            reuse_grad_buffers=True)
        self.sched = get_linear_schedule_with_warmup(self.opt, warmup_steps, total_steps)
        if self.device.type == "cuda":
            self.opt._build_fused()
        if averaging == "auto":
            averaging = "fused" if (self.world > 1 and self.device.type == "cuda") else "nccl"
        self.averaging = averaging
        self.group = None
        if self.world > 1 and averaging == "fused":
            from .parallel.group import FusedGroupAverager

            # what a rank exposes to its peers over NVLink: fp32 gradients, a bf16 copy, or uniform 8-bit codes
            payload = payload or os.environ.get("DALLE_B200_GROUP_PAYLOAD", "fp32")
            self.payload = payload
            self.group = FusedGroupAverager(self.opt, process_group, payload=payload)
        self.steps = 0
        # step() / accumulate() always run backward on the unscaled loss: the head's gradient is formed in the forward
        # pass (models/executor.py: expected_grad_out)
        if self.device.type == "cuda":
            self.model.engine.expected_grad_out = 1.0
        from .utils.phases import PhaseTrace

        self.trace = PhaseTrace()   # NVTX ranges + device-timed phases when DALLE_B200_TRACE=1

    # ---- one optimisation step on device-resident inputs -----------------------------------------
    def step(self, text: torch.Tensor, image: torch.Tensor) -> torch.Tensor:
        with self.trace.phase("forward"):
            loss = self.model(text, image, return_loss=True)
        with self.trace.phase("backward"):
            loss.backward()
        with self.trace.phase("averaging+optimizer"):
            self.optimizer_step()
        return loss.detach()

    def accumulate(self, text, image) -> torch.Tensor:
        """Forward/backward only (gradients accumulate in place) - for target-batch-size accumulation."""
        loss = self.model(text, image, return_loss=True)
        loss.backward()
        return loss.detach()

    def optimizer_step(self, grad_scale: float = 1.0):
        if self.world > 1:
            if self.group is not None:
                self.group.step(grad_scale / self.world)
                self.opt._opt_called = True   # the fused kernel was this step's optimizer.step() (LambdaLR's order check)
            else:
                flat = self._flat_grad()
                dist.all_reduce(flat, group=self.pg)
                self.opt.step(grad_scale=grad_scale / self.world, zero_grad=True)
        else:
            self.opt.step(grad_scale=grad_scale, zero_grad=True)
        self.sched.step()
        self.steps += 1

    def _flat_grad(self) -> torch.Tensor:
        if self.opt._fused is not None:
            return self.opt._fused.layout.flat_grad
        from torch._utils import _flatten_dense_tensors

        raise RuntimeError("flat gradient buffer requires the fused (CUDA) optimizer")

    # ---- end-to-end step from pinned host memory ---------------------------------------------------
    def step_from_host(self, text_cpu: torch.Tensor, image_cpu: torch.Tensor) -> float:
        text = text_cpu.to(self.device, non_blocking=True)
        image = image_cpu.to(self.device, non_blocking=True)
        loss = self.step(text, image)
        return float(loss.item())  # device -> host read of the step's result


def init_distributed_from_env(backend: Optional[str] = None):
    """One process per GPU; rendezvous from torchrun's environment (RANK / LOCAL_RANK / WORLD_SIZE / MASTER_*)."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world <= 1:
        return 0, 0, 1
    rank, local = int(os.environ["RANK"]), int(os.environ.get("LOCAL_RANK", os.environ["RANK"]))
    if backend is None:
        backend = "nccl" if torch.cuda.is_available() else "gloo"
    if torch.cuda.is_available():
        torch.cuda.set_device(local)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    if not dist.is_initialized():
        kw = {}
        if backend == "nccl":
            kw["device_id"] = torch.device("cuda", local)
        dist.init_process_group(backend, rank=rank, world_size=world, **kw)
    return rank, local, world
