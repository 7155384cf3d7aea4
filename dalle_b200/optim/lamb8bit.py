"""LAMB with block-wise 8-bit moments (replaces reference lib/training/lamb_8bit.py ``CPULAMB8Bit`` and the slice of
bitsandbytes it uses: ``Optimizer2State`` bookkeeping, ``create_dynamic_map``, ``quantize_blockwise`` /
``dequantize_blockwise``).

Two executors behind one ``torch.optim.Optimizer``:

* CUDA parameters: all parameters live in ONE flat fp32 buffer (each tensor starts on a 4096-element boundary, so
  quantisation blocks never straddle tensors and the state layout stays checkpoint compatible); ``step()`` is three
  kernel launches (csrc/lamb.cu) that also refresh the bf16 weight shadows used by the GEMMs, zero the gradient
  buffer and fold the per-step finite checks of the reference (hf_trainer.py:73-78, callback.py:95-100) into one
  device flag.
* CPU parameters: the per-tensor algorithm of the reference, using the C++ host quantiser
  (native/quant_host.cpp) when built, PyTorch otherwise.  This is the plumbing path of BASELINE config 1 and the
  oracle of the GPU tests.

``state_dict()`` has the reference layout (Appendix A.2): per parameter ``step, state1, state2`` (+ ``qmap1, qmap2,
absmax1, absmax2`` for tensors with >= ``min_8bit_size`` elements) and the diagnostics ``weight_norm, step_norm,
trust_ratio``.
"""
from __future__ import annotations

import math
from typing import Any, Dict, Iterable, List, Optional

import numpy as np
import torch

from ..ops import lib

BLOCK = 4096


# ------------------------------------------------------------------------------------------------
# quantisation maps and the block-wise quantiser (PyTorch statement = oracle)
# ------------------------------------------------------------------------------------------------
def create_dynamic_map(signed: bool = True, n: int = 7) -> torch.Tensor:
    """bitsandbytes 'dynamic tree' code: 256 sorted fp32 values in [-1, 1] (signed) or [0, 1] (unsigned)."""
    data: List[float] = []
    additional_items = 2 ** (7 - n) - 1
    if not signed:
        additional_items = 2 * additional_items
    i = 0
    for i in range(n):
        fraction_items = 2 ** (i + 7 - n) + 1 if signed else 2 ** (i + 7 - n + 1) + 1
        boundaries = torch.linspace(0.1, 1, fraction_items)
        means = (boundaries[:-1] + boundaries[1:]) / 2.0
        data += ((10 ** (-(n - 1) + i)) * means).tolist()
        if signed:
            data += (-(10 ** (-(n - 1) + i)) * means).tolist()
    if additional_items > 0:
        boundaries = torch.linspace(0.1, 1, additional_items + 1)
        means = (boundaries[:-1] + boundaries[1:]) / 2.0
        data += ((10 ** (-(n - 1) + i)) * means).tolist()
        if signed:
            data += (-(10 ** (-(n - 1) + i)) * means).tolist()
    data.append(0)
    data.append(1.0)
    data.sort()
    return torch.tensor(data, dtype=torch.float32)


def prepared_code(signed: bool) -> torch.Tensor:
    """The map as the quantiser actually uses it: bitsandbytes' CPU quantiser overwrites ``code[0] = -1.0`` in place
    on first use (for both maps), and the reference keeps using the modified tensor (lamb_8bit.py:192 note)."""
    code = create_dynamic_map(signed)
    code[0] = -1.0
    return code


def _nearest_code_index(code: torch.Tensor, x: torch.Tensor) -> torch.Tensor:
    idx = torch.searchsorted(code, x.contiguous(), right=True) - 1
    idx = idx.clamp_(0, 255)
    nxt = (idx + 1).clamp_(max=255)
    dl = (x - code[idx]).abs()
    dr = (x - code[nxt]).abs()
    return torch.where((idx < 255) & (dr < dl), nxt, idx)


def quantize_blockwise_ref(x: torch.Tensor, code: torch.Tensor, blocksize: int = BLOCK):
    flat = x.reshape(-1).float()
    n = flat.numel()
    nb = (n + blocksize - 1) // blocksize
    pad = nb * blocksize - n
    xp = torch.nn.functional.pad(flat, (0, pad)).view(nb, blocksize)
    absmax = xp.abs().amax(dim=1)
    inv = torch.where(absmax > 0, 1.0 / absmax, torch.zeros_like(absmax))
    idx = _nearest_code_index(code, (xp * inv[:, None]).reshape(-1)).view(nb, blocksize)
    return idx.reshape(-1)[:n].to(torch.uint8).view(x.shape), absmax


def dequantize_blockwise_ref(q: torch.Tensor, absmax: torch.Tensor, code: torch.Tensor, blocksize: int = BLOCK):
    flat = q.reshape(-1).long()
    n = flat.numel()
    vals = code[flat]
    scale = absmax.repeat_interleave(blocksize)[:n]
    return (vals * scale).view(q.shape)


def quantize_blockwise(x, code, blocksize=BLOCK):
    if not x.is_cuda and lib.host_lib_available():
        from ..native import host

        return host.quantize_blockwise(x, code, blocksize)
    return quantize_blockwise_ref(x, code, blocksize)


def dequantize_blockwise(q, absmax, code, blocksize=BLOCK):
    if not q.is_cuda and lib.host_lib_available():
        from ..native import host

        return host.dequantize_blockwise(q, absmax, code, blocksize)
    return dequantize_blockwise_ref(q, absmax, code, blocksize)


# ------------------------------------------------------------------------------------------------
# flat parameter storage
# ------------------------------------------------------------------------------------------------
class FlatLayout:
    """Re-homes parameters (and their gradients) into one flat fp32 buffer, 4096-aligned per tensor."""

    def __init__(self, params: List[torch.nn.Parameter]):
        self.params = params
        dev = params[0].device
        assert all(p.device == dev and p.dtype == torch.float32 for p in params)
        self.offsets, off = [], 0
        for p in params:
            self.offsets.append(off)
            off += (p.numel() + BLOCK - 1) // BLOCK * BLOCK
        self.n_padded = off
        self.flat = torch.zeros(off, device=dev, dtype=torch.float32)
        self.flat_grad = torch.zeros(off, device=dev, dtype=torch.float32)
        self.flat_bf16 = torch.zeros(off, device=dev, dtype=torch.bfloat16) if dev.type == "cuda" else None
        with torch.no_grad():
            for p, o in zip(params, self.offsets):
                n = p.numel()
                self.flat[o:o + n].copy_(p.detach().reshape(-1))
                old_grad = p.grad
                p.data = self.flat[o:o + n].view(p.shape)
                p.grad = self.flat_grad[o:o + n].view(p.shape)
                if old_grad is not None:
                    p.grad.copy_(old_grad)
                if self.flat_bf16 is not None:
                    p._bf16_shadow = self.flat_bf16[o:o + n].view(p.shape)
                    p._bf16_shadow.copy_(p.detach())
                    p._shadow_version = p._version

    @torch.no_grad()
    def rebind(self, flat: torch.Tensor, flat_grad: torch.Tensor, flat_bf16: Optional[torch.Tensor]):
        """Move the storage to caller-provided buffers (e.g. NVLink symmetric memory); views are re-pointed."""
        assert flat.numel() >= self.n_padded and flat_grad.numel() >= self.n_padded
        flat[: self.n_padded].copy_(self.flat)
        flat_grad[: self.n_padded].copy_(self.flat_grad)
        if flat_bf16 is not None and self.flat_bf16 is not None:
            flat_bf16[: self.n_padded].copy_(self.flat_bf16)
        self.flat, self.flat_grad = flat[: self.n_padded], flat_grad[: self.n_padded]
        self.flat_bf16 = flat_bf16[: self.n_padded] if flat_bf16 is not None else None
        for p, o in zip(self.params, self.offsets):
            n = p.numel()
            p.data = self.flat[o:o + n].view(p.shape)
            p.grad = self.flat_grad[o:o + n].view(p.shape)
            if self.flat_bf16 is not None:
                p._bf16_shadow = self.flat_bf16[o:o + n].view(p.shape)
                p._shadow_version = p._version

    def mark_shadows_fresh(self):
        for p in self.params:
            if self.flat_bf16 is not None:
                p._shadow_version = p._version


_BLOCK_DTYPE = np.dtype([("tensor", "<i4"), ("valid", "<i4"), ("state_off", "<i8"), ("quantized", "<i4"),
                         ("absmax_idx", "<i4")])


class LAMB8Bit(torch.optim.Optimizer):
    """Drop-in for the reference's ``CPULAMB8Bit`` (same constructor, same state layout), placed where the
    parameters are (GPU fused / CPU per-tensor)."""

    def __init__(self, params, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-6, weight_decay: float = 0,
                 clamp_value: float = 10, bias_correction: bool = False, min_8bit_size: int = 65536,
                 reuse_grad_buffers: bool = False, update_chunk_size: int = 2 ** 24,
                 max_grad_norm: Optional[float] = None, block_wise: int = BLOCK) -> None:
        if lr <= 0.0:
            raise ValueError(f"Invalid learning rate: {lr}")
        if eps < 0.0:
            raise ValueError(f"Invalid epsilon value: {eps}")
        if not 0.0 <= betas[0] < 1.0:
            raise ValueError(f"Invalid beta parameter at index 0: {betas[0]}")
        if not 0.0 <= betas[1] < 1.0:
            raise ValueError(f"Invalid beta parameter at index 1: {betas[1]}")
        if weight_decay < 0:
            raise ValueError(f"Invalid weight_decay value: {weight_decay}")
        if clamp_value < 0.0:
            raise ValueError(f"Invalid clamp value: {clamp_value}")
        assert block_wise == BLOCK
        self.clamp_value, self.bias_correction = clamp_value, bias_correction
        self.reuse_grad_buffers, self.update_chunk_size = reuse_grad_buffers, update_chunk_size
        self.max_grad_norm, self.min_8bit_size = max_grad_norm, min_8bit_size
        defaults = dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay)
        super().__init__(params, defaults)
        self.name2qmap: Dict[str, torch.Tensor] = {}
        self._fused = None
        self.last_step_skipped = False

    # ---- helpers -------------------------------------------------------------------------------
    def _all_params(self) -> List[torch.nn.Parameter]:
        return [p for g in self.param_groups for p in g["params"]]

    def fill_qmap(self, device):
        if "dynamic" not in self.name2qmap or self.name2qmap["dynamic"].device != device:
            self.name2qmap["dynamic"] = prepared_code(True).to(device)
            self.name2qmap["udynamic"] = prepared_code(False).to(device)

    def _is_quantized(self, p) -> bool:
        return p.numel() >= self.min_8bit_size and p.numel() >= BLOCK

    def init_state(self, p):
        state = self.state[p]
        state["step"] = 0
        if not self._is_quantized(p):
            state["state1"] = torch.zeros_like(p, dtype=torch.float32)
            state["state2"] = torch.zeros_like(p, dtype=torch.float32)
        else:
            self.fill_qmap(p.device)
            blocks = (p.numel() - 1) // BLOCK + 1
            state["state1"] = torch.zeros_like(p, dtype=torch.uint8)
            state["qmap1"] = self.name2qmap["dynamic"]
            state["state2"] = torch.zeros_like(p, dtype=torch.uint8)
            state["qmap2"] = self.name2qmap["udynamic"]
            state["absmax1"] = torch.zeros((blocks,), dtype=torch.float32, device=p.device)
            state["absmax2"] = torch.zeros((blocks,), dtype=torch.float32, device=p.device)

    # ---- CPU / reference executor ----------------------------------------------------------------
    @torch.no_grad()
    def _step_reference(self):
        params = [p for p in self._all_params() if p.grad is not None]
        if self.max_grad_norm is not None:
            torch.nn.utils.clip_grad_norm_(params, self.max_grad_norm)
        for group in self.param_groups:
            beta1, beta2 = group["betas"]
            for p in group["params"]:
                if p.grad is None:
                    continue
                if p not in self.state or "state1" not in self.state[p]:
                    self.init_state(p)
                state = self.state[p]
                step = state["step"] = state["step"] + 1
                g = p.grad.float()
                if state["state1"].dtype != torch.uint8:
                    m, v = state["state1"], state["state2"]
                    m.mul_(beta1).add_(g, alpha=1 - beta1)
                    v.mul_(beta2).addcmul_(g, g, value=1 - beta2)
                    delta = m / (v.sqrt() + group["eps"])
                else:
                    m = dequantize_blockwise(state["state1"], state["absmax1"], state["qmap1"])
                    v = dequantize_blockwise(state["state2"], state["absmax2"], state["qmap2"])
                    m.mul_(beta1).add_(g, alpha=1 - beta1)
                    v.mul_(beta2).addcmul_(g, g, value=1 - beta2)
                    q1, a1 = quantize_blockwise(m, state["qmap1"])
                    q2, a2 = quantize_blockwise(v, state["qmap2"])
                    state["state1"].copy_(q1); state["absmax1"].copy_(a1)
                    state["state2"].copy_(q2); state["absmax2"].copy_(a2)
                    delta = m / (v.sqrt() + group["eps"])
                if group["weight_decay"] != 0:
                    delta.add_(p, alpha=group["weight_decay"])
                step_norm = torch.norm(delta)
                weight_norm = p.norm().clamp(0, self.clamp_value)
                trust = weight_norm / step_norm if weight_norm != 0 and step_norm != 0 else 1.0
                state["weight_norm"], state["step_norm"], state["trust_ratio"] = weight_norm, step_norm, trust
                bc = math.sqrt(1 - beta2 ** step) / (1 - beta1 ** step) if self.bias_correction else 1
                p.add_(delta, alpha=-float(group["lr"] * bc * trust))

    # ---- fused CUDA executor ---------------------------------------------------------------------
    def _build_fused(self):
        params = self._all_params()
        dev = params[0].device
        layout = getattr(params[0], "_flat_layout", None)
        if layout is None or [id(p) for p in layout.params] != [id(p) for p in params]:
            layout = FlatLayout(params)
            for p in params:
                p._flat_layout = layout
        self.fill_qmap(dev)
        n_blocks = layout.n_padded // BLOCK
        blocks = np.zeros(n_blocks, dtype=_BLOCK_DTYPE)
        wd = np.zeros(len(params), dtype=np.float32)
        f_off = a_off = 0
        views = []
        pidx = 0
        for group in self.param_groups:
            for p in group["params"]:
                o, n = layout.offsets[pidx], p.numel()
                nb = (n + BLOCK - 1) // BLOCK
                quant = self._is_quantized(p)
                b0 = o // BLOCK
                blocks["tensor"][b0:b0 + nb] = pidx
                blocks["valid"][b0:b0 + nb] = BLOCK
                blocks["valid"][b0 + nb - 1] = n - (nb - 1) * BLOCK
                blocks["quantized"][b0:b0 + nb] = int(quant)
                if quant:
                    blocks["state_off"][b0:b0 + nb] = o + BLOCK * np.arange(nb)
                    blocks["absmax_idx"][b0:b0 + nb] = a_off + np.arange(nb)
                    views.append((p, True, o, a_off, nb))
                    a_off += nb
                else:
                    blocks["state_off"][b0:b0 + nb] = f_off + BLOCK * np.arange(nb)
                    views.append((p, False, f_off, 0, nb))
                    f_off += nb * BLOCK
                wd[pidx] = group["weight_decay"]
                pidx += 1
        f = _Fused()
        f.layout, f.n_blocks, f.n_tensors = layout, n_blocks, len(params)
        f.q1 = torch.zeros(layout.n_padded, device=dev, dtype=torch.uint8)
        f.q2 = torch.zeros(layout.n_padded, device=dev, dtype=torch.uint8)
        f.f1 = torch.zeros(max(f_off, BLOCK), device=dev, dtype=torch.float32)
        f.f2 = torch.zeros(max(f_off, BLOCK), device=dev, dtype=torch.float32)
        f.absmax1 = torch.zeros(max(a_off, 1), device=dev, dtype=torch.float32)
        f.absmax2 = torch.zeros(max(a_off, 1), device=dev, dtype=torch.float32)
        f.blocks = torch.from_numpy(blocks.view(np.uint8)).to(dev)
        f.wd = torch.from_numpy(wd).to(dev)
        f.sums = torch.zeros(2 * len(params), device=dev, dtype=torch.float32)
        f.trust = torch.zeros(3 * len(params), device=dev, dtype=torch.float32)
        f.stats = torch.zeros(2, device=dev, dtype=torch.float32)
        # per-parameter state entries are views into the flat state (checkpoint layout of Appendix A.2)
        for p, quant, so, ao, nb in views:
            st = self.state[p]
            old = dict(st)
            st["step"] = old.get("step", 0)
            n = p.numel()
            if quant:
                st["state1"] = f.q1[so:so + n].view(p.shape)
                st["state2"] = f.q2[so:so + n].view(p.shape)
                st["qmap1"], st["qmap2"] = self.name2qmap["dynamic"], self.name2qmap["udynamic"]
                st["absmax1"], st["absmax2"] = f.absmax1[ao:ao + nb], f.absmax2[ao:ao + nb]
            else:
                st["state1"] = f.f1[so:so + n].view(p.shape)
                st["state2"] = f.f2[so:so + n].view(p.shape)
            for k in ("state1", "state2", "absmax1", "absmax2"):
                if k in old and k in st and old[k].shape == st[k].shape:
                    st[k].copy_(old[k].to(st[k].device))
        self._fused = f
        return f

    @torch.no_grad()
    def _step_fused(self, grad_scale: float = 1.0, zero_grad: bool = False):
        f = self._fused or self._build_fused()
        g0 = self.param_groups[0]
        assert all(g["lr"] == g0["lr"] and g["betas"] == g0["betas"] and g["eps"] == g0["eps"]
                   for g in self.param_groups), "fused LAMB expects one lr/betas/eps for all groups"
        assert not self.bias_correction, "bias correction is off in the reference configuration"
        L = f.layout
        rc = lib.cuda_lib().b200_lamb_step(
            L.flat_grad.data_ptr(), L.flat.data_ptr(), L.flat_bf16.data_ptr(), L.n_padded, f.q1.data_ptr(),
            f.q2.data_ptr(), f.f1.data_ptr(), f.f2.data_ptr(), f.absmax1.data_ptr(), f.absmax2.data_ptr(),
            self.name2qmap["dynamic"].data_ptr(), self.name2qmap["udynamic"].data_ptr(), f.blocks.data_ptr(),
            f.n_blocks, f.wd.data_ptr(), f.sums.data_ptr(), f.n_tensors, f.trust.data_ptr(), f.stats.data_ptr(),
            g0["betas"][0], g0["betas"][1], g0["eps"], float(self.max_grad_norm or 0.0), self.clamp_value, g0["lr"],
            grad_scale, int(zero_grad), torch.cuda.current_stream().cuda_stream)
        lib.check(rc, "b200_lamb_step")
        lib.launch_count += 2
        L.mark_shadows_fresh()
        for p in L.params:
            self.state[p]["step"] = self.state[p].get("step", 0) + 1

    def step_was_skipped(self) -> bool:
        """True when the last fused step saw non-finite gradients and left parameters untouched (host sync)."""
        return bool(self._fused is not None and self._fused.stats[1].item() != 0.0)

    def grad_norm(self) -> float:
        return float(self._fused.stats[0].sqrt().item()) if self._fused is not None else float("nan")

    @torch.no_grad()
    def step(self, closure=None, grad_scale: float = 1.0, zero_grad: bool = False):
        loss = closure() if closure is not None else None
        params = self._all_params()
        if params and params[0].is_cuda:
            self._step_fused(grad_scale, zero_grad)
        else:
            if grad_scale != 1.0:
                for p in params:
                    if p.grad is not None:
                        p.grad.mul_(grad_scale)
            self._step_reference()
            if zero_grad:
                for p in params:
                    if p.grad is not None:
                        p.grad.zero_()
        return loss

    def zero_grad(self, set_to_none: bool = False):
        """Gradients are views of the flat gradient buffer: always zero in place, never drop them."""
        if self._fused is not None:
            self._fused.layout.flat_grad.zero_()
            return
        for p in self._all_params():
            if p.grad is not None:
                p.grad.zero_()

    def state_dict(self):
        if self._fused is not None:
            tr = self._fused.trust.view(-1, 3).cpu()
            for i, p in enumerate(self._fused.layout.params):
                st = self.state[p]
                st["weight_norm"], st["step_norm"], st["trust_ratio"] = tr[i, 0], tr[i, 1], tr[i, 2]
        return super().state_dict()

    def load_state_dict(self, state_dict):
        params = self._all_params()
        if self._fused is None and params and params[0].is_cuda:
            self._build_fused()
        views = {id(p): dict(self.state[p]) for p in params} if self._fused is not None else None
        super().load_state_dict(state_dict)
        if views is not None:  # keep the flat-buffer views, copy the loaded values into them
            for p in params:
                loaded, v = self.state[p], views[id(p)]
                for k in ("state1", "state2", "absmax1", "absmax2"):
                    if k in v:
                        if k in loaded and loaded[k] is not v[k]:
                            v[k].copy_(loaded[k].to(v[k].device).view_as(v[k]))
                        loaded[k] = v[k]
                for k in ("qmap1", "qmap2"):
                    if k in v:
                        loaded[k] = v[k]


class _Fused:
    pass


CPULAMB8Bit = LAMB8Bit


def get_linear_schedule_with_warmup(optimizer, num_warmup_steps: int, num_training_steps: int, last_epoch: int = -1):
    """lr * min(e / warmup, max(0, (total - e) / (total - warmup))) - stepped once per global epoch
    (reference task.py:163-165)."""

    def lr_lambda(step: int):
        if step < num_warmup_steps:
            return float(step) / float(max(1, num_warmup_steps))
        return max(0.0, float(num_training_steps - step) / float(max(1, num_training_steps - num_warmup_steps)))

    return torch.optim.lr_scheduler.LambdaLR(optimizer, lr_lambda, last_epoch)
