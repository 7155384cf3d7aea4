"""Baseline B1 of BASELINE.md, the bar BASELINE.json names ("the reference's own NCCL + cuBLAS"): the same TTT DALL-E
written the way a competent PyTorch user would write it on this box, with library kernels only -

* bf16 autocast over fp32 master weights; cuBLAS GEMMs through ``F.linear``;
* attention through ``F.scaled_dot_product_attention`` with the dense ``static_mask`` of the inference build (the
  same sparsity as the sparse build; reference ``inference/run_inference.py:72``), rotary on q, k and v;
* the reversible net's function (two residual streams) with plain autograd - activations are stored, nothing is
  recomputed, which is also what the product does at this batch size; ``--checkpoint`` switches to per-layer
  activation checkpointing when the batch does not fit;
* segment logits (text rows against the text rows of the tied matrix, image rows against the image rows) - the
  logits mask of dalle-pytorch stated structurally, as the product does;
* ``DistributedDataParallel`` (bucketed NCCL all-reduce overlapped with backward, gradients as bucket views);
* LAMB with ``torch._foreach_*`` kernels (fp32 moments), global-norm clipping, same hyper-parameters.

None of this repository's kernels, executor or optimizer is on this path (only the parameter container and the
rotary / mask tables).  The unmodified reference cannot be installed offline, so this arm is what the
product is measured against at 1 / 2 / 4 / 8 GPUs: ``bench.py --impl eager``.
"""
from __future__ import annotations

import math

import torch
import torch.distributed as dist
import torch.nn as nn
import torch.nn.functional as F
from torch.utils.checkpoint import checkpoint

from .config import DalleConfig
from .models.dalle import DALLE
from .models.masks import static_mask
from .models.reference import embed, shift_tokens
from .models.rotary import apply_rotary


class EagerDalle(nn.Module):
    """Wraps the parameter container (so that initial weights equal the product's) in a plain autograd forward."""

    def __init__(self, cfg: DalleConfig, seed: int = 0, use_checkpoint: bool = False):
        super().__init__()
        self.cfg, self.use_checkpoint = cfg, use_checkpoint
        self.p = DALLE(cfg, seed=seed)
        for t in sorted(set(cfg.attn_types)):
            self.register_buffer(f"mask_{t}", static_mask(cfg, t, cfg.seq_len), persistent=False)

    def _attn(self, i: int, x):
        cfg, m = self.cfg, self.p
        a = m.attn_index[i]
        b, n, _ = x.shape
        h = F.layer_norm(x, (cfg.dim,), m.norm_f_w[i], m.norm_f_b[i], cfg.ln_eps)
        h = shift_tokens(h, cfg)
        qkv = F.linear(h, m.attn_qkv_w[a]).view(b, n, 3, cfg.heads, cfg.dim_head).permute(2, 0, 3, 1, 4)
        ang = m.rotary_angles[:n]
        q, k, v = apply_rotary(qkv[0], ang), apply_rotary(qkv[1], ang), apply_rotary(qkv[2], ang)
        o = F.scaled_dot_product_attention(q, k, v, attn_mask=getattr(self, f"mask_{cfg.attn_types[i]}")[:n, :n])
        o = F.linear(o.transpose(1, 2).reshape(b, n, cfg.inner_dim), m.attn_out_w[a], m.attn_out_b[a])
        return o * m.scale_f[i]

    def _ff(self, i: int, x):
        cfg, m = self.cfg, self.p
        f = m.ff_index[i]
        h = F.layer_norm(x, (cfg.dim,), m.norm_g_w[i], m.norm_g_b[i], cfg.ln_eps)
        h = shift_tokens(h, cfg)
        val, gate = F.linear(h, m.ff_w1[f], m.ff_b1[f]).chunk(2, dim=-1)
        return F.linear(val * F.gelu(gate), m.ff_w2[f], m.ff_b2[f]) * m.scale_g[i]

    def _layer(self, i: int, y1, y2):
        y1 = y1 + self._attn(i, y2)
        y2 = y2 + self._ff(i, y1)
        return y1, y2

    def forward(self, text, image):
        cfg, m = self.cfg, self.p
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=text.is_cuda):
            tokens, text_b = embed(m, text, image)
            y1 = y2 = tokens.float()
            for i in range(cfg.depth):
                if self.use_checkpoint:
                    y1, y2 = checkpoint(self._layer, i, y1, y2, use_reentrant=False)
                else:
                    y1, y2 = self._layer(i, y1, y2)
            h = F.layer_norm((y1 + y2) * 0.5, (cfg.dim,), m.final_norm_w, m.final_norm_b, cfg.ln_eps)
            nt, tt = cfg.text_seq_len, cfg.total_text_tokens
            labels = torch.cat((text_b[:, 1:], image), dim=1)
            lt = F.linear(h[:, :nt], m.emb_weight[:tt], m.emb_bias[:tt])
            li = F.linear(h[:, nt:], m.emb_weight[tt:], m.emb_bias[tt:])
        loss_t = F.cross_entropy(lt.float().flatten(0, 1), labels[:, :nt].flatten())
        loss_i = F.cross_entropy(li.float().flatten(0, 1), labels[:, nt:].flatten())
        return (loss_t + cfg.loss_img_weight * loss_i) / (cfg.loss_img_weight + 1)


class ForeachLamb:
    """LAMB (reference lib/training/lamb_8bit.py math, fp32 moments) on multi-tensor ``torch._foreach`` kernels."""

    def __init__(self, decay, no_decay, lr, betas, eps, weight_decay, max_grad_norm, clamp_value):
        self.groups = [(list(decay), weight_decay), (list(no_decay), 0.0)]
        self.lr, self.betas, self.eps = lr, betas, eps
        self.max_grad_norm, self.clamp_value = max_grad_norm, clamp_value
        self.m = [[torch.zeros_like(p) for p in ps] for ps, _ in self.groups]
        self.v = [[torch.zeros_like(p) for p in ps] for ps, _ in self.groups]
        self.param_groups = [{"lr": lr}]

    @torch.no_grad()
    def step(self):
        lr = self.param_groups[0]["lr"]
        b1, b2 = self.betas
        all_p = [p for ps, _ in self.groups for p in ps]
        torch.nn.utils.clip_grad_norm_(all_p, self.max_grad_norm, foreach=True)
        for (ps, wd), ms, vs in zip(self.groups, self.m, self.v):
            gs = [p.grad for p in ps]
            torch._foreach_mul_(ms, b1)
            torch._foreach_add_(ms, gs, alpha=1 - b1)
            torch._foreach_mul_(vs, b2)
            torch._foreach_addcmul_(vs, gs, gs, value=1 - b2)
            den = torch._foreach_sqrt(vs)
            torch._foreach_add_(den, self.eps)
            upd = torch._foreach_div(ms, den)
            if wd:
                torch._foreach_add_(upd, ps, alpha=wd)
            wn = torch._foreach_norm(ps)
            un = torch._foreach_norm(upd)
            wn_t = torch.stack(wn).clamp_(0, self.clamp_value)
            un_t = torch.stack(un)
            trust = torch.where((wn_t > 0) & (un_t > 0), wn_t / un_t, torch.ones_like(wn_t))
            torch._foreach_mul_(upd, list((-lr * trust).unbind(0)))
            torch._foreach_add_(ps, upd)

    def zero_grad(self):
        for ps, _ in self.groups:
            for p in ps:
                p.grad = None


class DDPTrainer:
    averaging = "nccl-ddp"
    payload = "fp32 buckets"

    def __init__(self, cfg: DalleConfig, device, seed: int = 0, use_checkpoint: bool = False, **kw):
        self.cfg, self.device = cfg, torch.device(device)
        self.world = dist.get_world_size() if dist.is_initialized() else 1
        self.net = EagerDalle(cfg, seed=seed, use_checkpoint=use_checkpoint).to(self.device)
        self.model = self.net
        if self.world > 1:
            self.model = nn.parallel.DistributedDataParallel(self.net, device_ids=[self.device.index],
                                                             gradient_as_bucket_view=True, static_graph=True)
        named = list(self.net.p.reference_named_parameters())
        decay = [p for n, p in named if "bias" not in n]
        no_decay = [p for n, p in named if "bias" in n]
        self.opt = ForeachLamb(decay, no_decay, lr=2.5e-3, betas=(0.9, 0.96), eps=1e-6, weight_decay=0.045,
                               max_grad_norm=4.0, clamp_value=10000.0)
        self.warmup, self.total, self.steps = 3125, 31250, 0

    def step(self, text, image):
        loss = self.model(text, image)
        loss.backward()
        e = self.steps
        self.opt.param_groups[0]["lr"] = self.opt.lr * min(e / self.warmup, max(0.0, (self.total - e) / (self.total - self.warmup)))
        self.opt.step()
        self.opt.zero_grad()
        self.steps += 1
        return loss.detach()

    def step_from_host(self, text_cpu, image_cpu) -> float:
        return float(self.step(text_cpu.to(self.device, non_blocking=True),
                               image_cpu.to(self.device, non_blocking=True)).item())
