"""Sparse DALL-E attention core: (rotated, pre-scaled) qkv -> o, with its backward.

GPU tensors run the tcgen05 kernels (csrc/attn_fwd_sm100.cu, attn_bwd_sm100.cu); CPU tensors run the dense-mask
PyTorch statement below, which is also the oracle of the GPU tests.
"""
from __future__ import annotations

import torch

from ..config import DalleConfig
from ..models.masks import ATTN_TYPE_CODES, static_mask
from . import lib

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
        o, l = attn_core_ref(qkv, cfg, attn_type)
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
