"""Sparse DALL-E attention core: (rotated, pre-scaled) qkv -> o, with its backward.

GPU tensors run the tcgen05 kernels: axial row / column layers go to csrc/attn_fwd4_sm100.cu and
attn_bwd4_sm100.cu (64-key chunks, P and dS consumed from TMEM), the convolution-like layer to attn_fwd2_sm100.cu and
attn_bwd_sm100.cu, and image grids wider than 32 to the first-generation forward in attn_fwd_sm100.cu
(``DALLE_B200_ATTN_FWD`` / ``DALLE_B200_ATTN_BWD`` select an older generation for A/B runs).  CPU tensors run the
dense-mask PyTorch statement below, which is also the oracle of the GPU tests.
"""
from __future__ import annotations

import torch

from ..config import DalleConfig
from ..models.masks import ATTN_TYPE_CODES, static_mask
from . import lib
from .gemm import rot_transposed

_MASK_CACHE = {}


def _mask(cfg: DalleConfig, attn_type: str, device) -> torch.Tensor:
    key = (cfg.text_seq_len, cfg.image_fmap_size, cfg.conv_kernel_size, cfg.conv_dilation, attn_type, str(device))
    m = _MASK_CACHE.get(key)
    if m is None:
        m = static_mask(cfg, attn_type, cfg.seq_len).to(device)
        _MASK_CACHE[key] = m
    return m


def attn_core_ref(qkv: torch.Tensor, cfg: DalleConfig, attn_type: str):
    """qkv [B*seq, 3*H*64] -> (o [B*seq, H*64], lse [B, H, seq] in log2 units). fp32 math."""
    H, dh, seq = cfg.heads, cfg.dim_head, cfg.seq_len
    B = qkv.shape[0] // seq
    x = qkv[: B * seq].float().reshape(B, seq, 3, H, dh).permute(2, 0, 3, 1, 4)
    q, k, v = x[0], x[1], x[2]
    dots = torch.einsum("bhid,bhjd->bhij", q, k)
    dots = dots.masked_fill(~_mask(cfg, attn_type, qkv.device), float("-inf"))
    lse = torch.logsumexp(dots, dim=-1)
    p = torch.exp(dots - lse[..., None])
    o = torch.einsum("bhij,bhjd->bhid", p, v).permute(0, 2, 1, 3).reshape(B * seq, H * dh)
    return o, lse * 1.4426950408889634


def attn_fwd(qkv: torch.Tensor, out: torch.Tensor, lse: torch.Tensor, cfg: DalleConfig, attn_type: str):
    """qkv must own at least one padding row beyond B*seq rows when it lives on the GPU (the image grid has one
    position more than there are tokens; the column-gather TMA map touches it)."""
    if not qkv.is_cuda:
        o, l = attn_core_ref(qkv[: out.shape[0]], cfg, attn_type)
        out.copy_(o.to(out.dtype))
        lse.copy_(l)
        return out
    assert cfg.dim_head == 64 and cfg.conv_dilation == 1
    B = out.shape[0] // cfg.seq_len
    rc = lib.cuda_lib().b200_attn_fwd(qkv.data_ptr(), out.data_ptr(), lse.data_ptr(), B, cfg.heads, cfg.seq_len,
                                      cfg.text_len, cfg.image_fmap_size, ATTN_TYPE_CODES[attn_type] + 1,
                                      cfg.conv_kernel_size, torch.cuda.current_stream().cuda_stream)
    lib.check(rc, "b200_attn_fwd")
    lib.launch_count += 1  # two kernels per call (text block + image block)
    return out


def inv_rotary_scale(d_rot: torch.Tensor, rot_pairs: torch.Tensor, cfg: DalleConfig, q_scale: float) -> torch.Tensor:
    """Transpose of the QKV epilogue (rotary on q,k,v + q scale): gradient w.r.t. the raw projection output."""
    M, C = d_rot.shape
    pos = torch.arange(M, device=d_rot.device) % cfg.seq_len
    cs = rot_pairs[pos].to(d_rot.device)
    cos, sin = cs[:, None, :, 0], cs[:, None, :, 1]
    x = d_rot.float().view(M, C // 64, 32, 2)
    x0, x1 = x[..., 0], x[..., 1]
    out = torch.stack((x0 * cos + x1 * sin, x1 * cos - x0 * sin), dim=-1).view(M, C).clone()
    out[:, : cfg.inner_dim] *= q_scale
    return out


def attn_bwd_ref(qkv, d_out, cfg: DalleConfig, attn_type: str, rot_pairs, q_scale):
    M = d_out.shape[0] // cfg.seq_len * cfg.seq_len  # padding rows (fewer than one sample) are ignored
    x = qkv[:M].detach().float().requires_grad_(True)
    with torch.enable_grad():
        o, _ = attn_core_ref(x, cfg, attn_type)
    (d_rot,) = torch.autograd.grad(o, x, d_out[:M].float())
    return inv_rotary_scale(d_rot, rot_pairs, cfg, q_scale)


def attn_bwd(qkv, out, d_out, lse, delta, dqkv, kv_acc, rot_pairs, cfg: DalleConfig, attn_type: str):
    """dqkv <- gradient w.r.t. the raw QKV projection.  GPU: d_out needs a padding row like qkv; kv_acc must be
    zero on entry (it is returned zeroed)."""
    q_scale = cfg.dim_head ** -0.5
    if not qkv.is_cuda:
        dqkv.copy_(attn_bwd_ref(qkv, d_out, cfg, attn_type, rot_pairs, q_scale).to(dqkv.dtype))
        return dqkv
    B = out.shape[0] // cfg.seq_len
    rc = lib.cuda_lib().b200_attn_bwd(qkv.data_ptr(), out.data_ptr(), d_out.data_ptr(), lse.data_ptr(),
                                      delta.data_ptr(), dqkv.data_ptr(), kv_acc.data_ptr(), rot_pairs.data_ptr(),
                                      rot_transposed(rot_pairs).data_ptr(), q_scale, B, cfg.heads, cfg.seq_len, cfg.text_len, cfg.image_fmap_size,
                                      ATTN_TYPE_CODES[attn_type] + 1, cfg.conv_kernel_size,
                                      torch.cuda.current_stream().cuda_stream)
    lib.check(rc, "b200_attn_bwd")
    lib.launch_count += 3  # delta + text blocks + image blocks + finish
    return dqkv
