"""Intra-box group runtime: symmetric NVLink memory + the fused averaging/optimizer kernel (SURVEY §5.8 level 1, §8
row 14).  One process per GPU; ``torch.distributed`` (NCCL) is used ONLY to bootstrap: exchange the peer-memory
handles (``torch.distributed._symmetric_memory`` as the allocator/rendezvous).  The step itself is
``b200_group_lamb_step`` - a cooperative kernel per rank that reduce-scatters gradients with P2P loads, runs LAMB-8bit
on its shard and all-gathers the updated parameters with P2P stores, synchronising with device-side flags.

Replaces ``nn.DataParallel`` inside a box: no per-step parameter broadcast (parameters are updated in
place on every rank), no gradient gather onto GPU 0, and the optimizer work is sharded 1/world per GPU.
"""
from __future__ import annotations

import ctypes
from typing import Optional

import torch
import torch.distributed as dist

from ..ops import lib
from ..optim.lamb8bit import BLOCK, LAMB8Bit


class FusedGroupAverager:
    def __init__(self, opt: LAMB8Bit, process_group=None, payload: str = "fp32"):
        import torch.distributed._symmetric_memory as symm

        self.opt = opt
        self.pg = process_group or dist.group.WORLD
        self.rank, self.world = dist.get_rank(self.pg), dist.get_world_size(self.pg)
        assert self.world <= 8, "one box: at most 8 ranks"
        f = opt._fused or opt._build_fused()
        L = f.layout
        dev = L.flat.device
        n = L.n_padded
        gname = self.pg.group_name

        def alloc(numel, dtype):
            t = symm.empty(numel, dtype=dtype, device=dev)
            t.zero_()
            h = symm.rendezvous(t, gname)
            return t, h

        self.grad, hg = alloc(n, torch.float32)
        self.param, hp = alloc(n, torch.float32)
        self.shadow, hs = alloc(n, torch.bfloat16)
        # what a rank exposes to its peers over NVLink.  fp32: the gradient buffer itself; bf16 / u8 (absmax int8 per
        # 4096 block): a compact copy; fp16 / u8h / adaptive: hivemind's averaging codecs (clamped fp16; uniform 8-bit
        # with a bucket-mean codebook per 131072-element chunk; `adaptive` = the size rule of reference
        # task.py:124-126: tensors of 65536 elements and more travel as 8-bit, smaller ones as fp16)
        self.payload_mode = {"fp32": 0, "bf16": 1, "u8": 2, "fp16": 3, "u8h": 3, "adaptive": 3}[payload]
        self.payload_name = payload
        # bf16: 2 B / element; u8: 1 B / element + one fp32 step per 4096-element block (fits the same 2 B / element)
        self.payload, hpl = alloc(n if self.payload_mode else 8, torch.bfloat16)
        self.xstride = (2 + 2 * f.n_tensors + 31) // 32 * 32
        self.xchg, hx = alloc(self.world * self.xstride, torch.float32)
        self.sig, hsig = alloc(64, torch.int32)
        self._handles = (hg, hp, hs, hpl, hx, hsig)
        L.rebind(self.param, self.grad, self.shadow)
        # shard = contiguous run of 4096-element blocks
        nb = f.n_blocks
        per = (nb + self.world - 1) // self.world
        self.blk0, self.blk1 = min(nb, self.rank * per), min(nb, (self.rank + 1) * per)
        self.stats = torch.zeros(4, device=dev, dtype=torch.float32)
        self.epoch = 0
        d = lib.B200Group()
        d.rank, d.world = self.rank, self.world
        for r in range(self.world):
            d.grad_ptrs[r], d.param_ptrs[r], d.shadow_ptrs[r] = hg.buffer_ptrs[r], hp.buffer_ptrs[r], hs.buffer_ptrs[r]
            d.payload_ptrs[r], d.xchg_ptrs[r], d.signal_ptrs[r] = hpl.buffer_ptrs[r], hx.buffer_ptrs[r], hsig.buffer_ptrs[r]
        d.xstride, d.n_tensors, d.n_blocks, d.blk0, d.blk1 = self.xstride, f.n_tensors, nb, self.blk0, self.blk1
        d.payload_mode = self.payload_mode
        self.chunk_codec = None
        if self.payload_mode == 3:
            import numpy as np

            from ..optim.lamb8bit import _BLOCK_DTYPE

            quant = f.blocks.cpu().numpy().view(_BLOCK_DTYPE)["quantized"] != 0
            n_chunks = (nb + 31) // 32
            big = np.array([quant[32 * c: 32 * c + 32].all() for c in range(n_chunks)])
            codec = {"fp16": np.full(n_chunks, 3), "u8h": np.full(n_chunks, 4),
                     "adaptive": np.where(big, 4, 3)}[payload].astype(np.uint8)
            self.chunk_codec = torch.from_numpy(codec).to(dev)
            d.chunk_codec = self.chunk_codec.data_ptr()
        d.q1, d.q2, d.f1, d.f2 = f.q1.data_ptr(), f.q2.data_ptr(), f.f1.data_ptr(), f.f2.data_ptr()
        d.absmax1, d.absmax2 = f.absmax1.data_ptr(), f.absmax2.data_ptr()
        d.code1, d.code2 = opt.name2qmap["dynamic"].data_ptr(), opt.name2qmap["udynamic"].data_ptr()
        d.blocks, d.wd_per_tensor = f.blocks.data_ptr(), f.wd.data_ptr()
        d.tensor_sums, d.stats, d.trust_out = f.sums.data_ptr(), self.stats.data_ptr(), f.trust.data_ptr()
        self._desc = d
        torch.cuda.synchronize()
        dist.barrier(self.pg)  # every rank's buffers are initialised before the first device-side barrier

    def reduce_gradients(self, grad_scale: float = 1.0):
        """Level-1 half of a two-level step: afterwards EVERY rank's gradient buffer holds
        ``grad_scale * sum_r grad_r`` (reduce-scatter + all-gather over NVLink, same kernel, no optimizer work)."""
        d = self._desc
        d.grad_scale, d.zero_grad, d.epoch, d.reduce_only = grad_scale, 0, self.epoch, 1
        self.epoch = (self.epoch + 4) & 0xFFFFFFFF   # wraps like the device's 32-bit signed-difference compare
        rc = lib.cuda_lib().b200_group_lamb_step(ctypes.byref(d), torch.cuda.current_stream().cuda_stream)
        lib.check(rc, "b200_group_lamb_step(reduce_only)")
        d.reduce_only = 0

    def step(self, grad_scale: float = 1.0, zero_grad: bool = True):
        opt, d = self.opt, self._desc
        g0 = opt.param_groups[0]
        d.beta1, d.beta2, d.eps = g0["betas"][0], g0["betas"][1], g0["eps"]
        d.max_grad_norm, d.clamp_value, d.lr = float(opt.max_grad_norm or 0.0), opt.clamp_value, g0["lr"]
        d.grad_scale, d.zero_grad, d.epoch = grad_scale, int(zero_grad), self.epoch
        self.epoch = (self.epoch + 4) & 0xFFFFFFFF   # wraps like the device's 32-bit signed-difference compare
        rc = lib.cuda_lib().b200_group_lamb_step(ctypes.byref(d), torch.cuda.current_stream().cuda_stream)
        lib.check(rc, "b200_group_lamb_step")
        opt._fused.layout.mark_shadows_fresh()
        for p in opt._fused.layout.params:
            opt.state[p]["step"] = opt.state[p].get("step", 0) + 1

    def step_was_skipped(self) -> bool:
        return bool(self.stats[3].item() != 0.0)

    def grad_norm(self) -> float:
        return float(self.stats[2].sqrt().item())
