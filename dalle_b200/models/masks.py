"""Attention sparsity patterns (dalle-pytorch SparseAxialCausalAttention /
SparseConvCausalAttention, and the equivalent dense ``static_mask`` used by the inference build, component M9).

Positions t < text_len are text (<bos> + text tokens); image index p = t - text_len, (r, c) = divmod(p, fmap).
Text query t sees text keys <= t.  Image query p sees ALL text keys plus
  axial_row : {(r, c') : c' <= c}
  axial_col : {(r', c) : r' <= r}
  conv_like : {(r', c') : |r'-r| <= k//2 * dil, |c'-c| <= k//2 * dil (on the dilation grid), 32 r' + c' <= p}
One softmax over the union.
"""
from __future__ import annotations

import torch

from ..config import DalleConfig


def static_mask(cfg: DalleConfig, attn_type: str, n: int | None = None) -> torch.Tensor:
    """Boolean [n, n] (True = may attend), n defaults to seq_len + 1 (the padded length the sparse build uses)."""
    fmap, text_len = cfg.image_fmap_size, cfg.text_len
    total = text_len + cfg.image_seq_len
    idx = torch.arange(total)
    is_img = idx >= text_len
    p = (idx - text_len).clamp(min=0)
    r, c = p // fmap, p % fmap
    qi, kj = idx[:, None], idx[None, :]
    causal = kj <= qi
    q_img, k_img = is_img[:, None], is_img[None, :]
    rq, cq, rk, ck = r[:, None], c[:, None], r[None, :], c[None, :]
    if attn_type == "axial_row":
        local = rq == rk
    elif attn_type == "axial_col":
        local = cq == ck
    elif attn_type == "conv_like":
        half = (cfg.conv_kernel_size // 2) * cfg.conv_dilation
        dr, dc = rk - rq, ck - cq
        local = (dr.abs() <= half) & (dc.abs() <= half)
        if cfg.conv_dilation > 1:
            local &= (dr % cfg.conv_dilation == 0) & (dc % cfg.conv_dilation == 0)
    elif attn_type == "full":
        local = torch.ones(total, total, dtype=torch.bool)
    else:
        raise ValueError(f"attention type {attn_type!r} is not supported")
    mask = torch.where(q_img, ~k_img | (k_img & local), ~k_img) & causal
    n = total if n is None else n
    return mask[:n, :n].contiguous()


ATTN_TYPE_CODES = {"axial_row": 0, "axial_col": 1, "conv_like": 2, "full": 3}
