"""Rotary position tables (SURVEY Appendix D.4 / component M13).

dalle-pytorch builds one ``pos_emb[1, seq_len+1, 60]`` angle table: 20 dims of text position ('lang' frequencies,
image tokens pinned at position 8192) followed by 20+20 dims of image row/column position ('pixel' frequencies over
linspace(-1, 1, fmap), text pinned at -10).  Every frequency is repeated twice so that consecutive channel pairs
(2i, 2i+1) rotate together; the table is applied to the first ``rot`` dims of q, k AND v.
"""
from __future__ import annotations

import math

import torch

from ..config import DalleConfig


def rotary_angles(cfg: DalleConfig) -> torch.Tensor:
    """[seq_len + 1, rot] angles, rot = 6 * (dim_head // 3 // 2) (60 for dim_head 64)."""
    rot_dim = cfg.dim_head // 3
    nf = rot_dim // 2
    fmap, img_len, text_len = cfg.image_fmap_size, cfg.image_seq_len, cfg.text_len
    lang = 1.0 / (10000 ** (torch.arange(0, rot_dim, 2)[:nf].double() / rot_dim))
    pixel = torch.linspace(1.0, 10.0 / 2, nf, dtype=torch.float64) * math.pi

    def emb(t, freqs):
        f = t.double()[:, None] * freqs[None, :]
        return f.repeat_interleave(2, dim=-1)

    text_freqs = torch.cat((emb(torch.arange(text_len), lang), emb(torch.full((img_len,), 8192.0), lang)), dim=0)
    axial = emb(torch.linspace(-1, 1, fmap, dtype=torch.float64), pixel)  # [fmap, 2nf]
    img_freqs = torch.cat((axial[:, None, :].expand(fmap, fmap, -1), axial[None, :, :].expand(fmap, fmap, -1)), dim=-1)
    img_freqs = img_freqs.reshape(img_len, -1)
    text_axial = emb(torch.full((text_len,), -10.0), pixel)
    img_freqs = torch.cat((torch.cat((text_axial, text_axial), dim=-1), img_freqs), dim=0)
    return torch.cat((text_freqs, img_freqs), dim=-1).float()  # [text_len + img_len, 6nf]


def rotary_pair_table(cfg: DalleConfig) -> torch.Tensor:
    """[seq_len + 1, dim_head // 2, 2] (cos, sin) per channel pair; unrotated tail pairs are (1, 0).

    This is the layout consumed by the QKV GEMM epilogue and the attention backward kernels."""
    ang = rotary_angles(cfg)
    n, rot = ang.shape
    pairs = cfg.dim_head // 2
    a = torch.zeros(n, pairs, dtype=torch.float64)
    a[:, : rot // 2] = ang[:, 0::2].double()
    return torch.stack((a.cos(), a.sin()), dim=-1).float().contiguous()


def apply_rotary(t: torch.Tensor, angles: torch.Tensor) -> torch.Tensor:
    """Reference application: t [..., n, dim_head], angles [n, rot] (interleaved-pair rotate_half)."""
    rot = angles.shape[-1]
    tl, tr = t[..., :rot], t[..., rot:]
    x = tl.reshape(*tl.shape[:-1], rot // 2, 2)
    rh = torch.stack((-x[..., 1], x[..., 0]), dim=-1).reshape(tl.shape)
    tl = tl * angles.cos().to(t.dtype) + rh * angles.sin().to(t.dtype)
    return torch.cat((tl, tr), dim=-1)
