"""Deliberately naive PyTorch statement of the DALL-E transformer the reference trains.

This is the ORACLE: dense masks or literal sparse gathers, no fusion, works on CPU and GPU in any dtype.  Every
CUDA kernel and the fused layer executor are tested against these functions; ``bench.py --impl eager`` times this
same code in bf16 as the "cuBLAS + ATen" baseline B1 of BASELINE.md.

Behaviour follows dalle-pytorch (weight-sharing fork pinned by reference requirements.txt:2) as configured at
reference task.py:62-83.
"""
from __future__ import annotations

from typing import List, Optional

import torch
import torch.nn.functional as F

from ..config import DalleConfig
from .masks import static_mask
from .rotary import apply_rotary


# ------------------------------------------------------------------------------------------------
# token shift (PreShiftToken, M6)
# ------------------------------------------------------------------------------------------------
def shift_tokens(x: torch.Tensor, cfg: DalleConfig) -> torch.Tensor:
    """x [b, n<=seq_len, d]. Text: first half of channels from the previous position.  Image: first quarter from
    the token above, second quarter from the token to the left (zero at row start), rest unchanged."""
    b, n, d = x.shape
    text_len, fmap = cfg.text_len, cfg.image_fmap_size
    x_text, x_img = x[:, :text_len], x[:, text_len:]
    half = d // 2
    t_shift = F.pad(x_text[..., :half], (0, 0, 1, -1))
    x_text = torch.cat((t_shift, x_text[..., half:]), dim=-1)
    n_img = x_img.shape[1]
    if n_img == 0:
        return x_text
    pad = cfg.image_seq_len - n_img
    xi = F.pad(x_img, (0, 0, 0, pad)).reshape(b, fmap, fmap, d)
    q = d // 4
    top = F.pad(xi[..., :q], (0, 0, 0, 0, 1, -1))
    left = F.pad(xi[..., q:2 * q], (0, 0, 1, -1))
    xi = torch.cat((top, left, xi[..., 2 * q:]), dim=-1).reshape(b, fmap * fmap, d)[:, :n_img]
    return torch.cat((x_text, xi), dim=1)


# ------------------------------------------------------------------------------------------------
# attention (M7, M8, M9)
# ------------------------------------------------------------------------------------------------
def _qkv(x, w_qkv, cfg: DalleConfig, angles: Optional[torch.Tensor]):
    b, n, _ = x.shape
    h, dh = cfg.heads, cfg.dim_head
    qkv = F.linear(x, w_qkv).reshape(b, n, 3, h, dh).permute(2, 0, 3, 1, 4)  # [3, b, h, n, dh]
    q, k, v = qkv[0], qkv[1], qkv[2]
    if angles is not None:
        a = angles[:n].to(x.device)
        q, k, v = apply_rotary(q, a), apply_rotary(k, a), apply_rotary(v, a)
    q = q * (dh ** -0.5)
    return q, k, v


def attention_dense(x, w_qkv, w_out, b_out, cfg: DalleConfig, attn_type: str, angles) -> torch.Tensor:
    """Dense attention with the precomputed boolean pattern mask (the inference build of the reference)."""
    b, n, _ = x.shape
    q, k, v = _qkv(x, w_qkv, cfg, angles)
    mask = static_mask(cfg, attn_type, n).to(x.device)
    dots = torch.einsum("bhid,bhjd->bhij", q, k)
    dots = dots.masked_fill(~mask, -torch.finfo(dots.dtype).max)
    attn = dots.float().softmax(dim=-1).to(dots.dtype)
    out = torch.einsum("bhij,bhjd->bhid", attn, v)
    out = out.permute(0, 2, 1, 3).reshape(b, n, cfg.inner_dim)
    return F.linear(out, w_out, b_out)


def attention_sparse(x, w_qkv, w_out, b_out, cfg: DalleConfig, attn_type: str, angles) -> torch.Tensor:
    """Literal sparse formulation (training build): text->text causal block, image->(all text + local keys)."""
    b, n, _ = x.shape
    assert n == cfg.seq_len, "the sparse build expects a full-length sequence"
    h, dh, fmap, text_len = cfg.heads, cfg.dim_head, cfg.image_fmap_size, cfg.text_len
    x = F.pad(x, (0, 0, 0, 1))  # the reference pads to seq_len + 1 so that the image part is a full grid
    q, k, v = _qkv(x, w_qkv, cfg, angles)
    neg = -torch.finfo(q.dtype).max
    qt, kt, vt = q[:, :, :text_len], k[:, :, :text_len], v[:, :, :text_len]
    qi, ki, vi = q[:, :, text_len:], k[:, :, text_len:], v[:, :, text_len:]
    # text -> text
    dt = torch.einsum("bhid,bhjd->bhij", qt, kt)
    causal = torch.ones(text_len, text_len, dtype=torch.bool, device=x.device).triu_(1)
    dt = dt.masked_fill(causal, neg)
    out_text = torch.einsum("bhij,bhjd->bhid", dt.float().softmax(-1).to(dt.dtype), vt)
    # image -> text + local
    d_i2t = torch.einsum("bhid,bhjd->bhij", qi, kt)  # [b,h,P,text_len]
    P = fmap * fmap
    if attn_type in ("axial_row", "axial_col"):
        def split(t):
            t = t.reshape(b, h, fmap, fmap, dh)
            return t if attn_type == "axial_row" else t.transpose(2, 3)
        qg, kg, vg = split(qi), split(ki), split(vi)  # [b,h,x,i,dh]
        d_loc = torch.einsum("bhxid,bhxjd->bhxij", qg, kg)
        cm = torch.ones(fmap, fmap, dtype=torch.bool, device=x.device).triu_(1)
        d_loc = d_loc.masked_fill(cm, neg)
        d_txt = d_i2t.reshape(b, h, fmap, fmap, text_len)
        if attn_type == "axial_col":
            d_txt = d_txt.transpose(2, 3)
        dots = torch.cat((d_txt, d_loc), dim=-1)
        attn = dots.float().softmax(-1).to(dots.dtype)
        a_txt, a_loc = attn[..., :text_len], attn[..., text_len:]
        out_img = torch.einsum("bhxij,bhxjd->bhxid", a_loc, vg) + torch.einsum("bhxij,bhjd->bhxid", a_txt, vt)
        if attn_type == "axial_col":
            out_img = out_img.transpose(2, 3)
        out_img = out_img.reshape(b, h, P, dh)
    elif attn_type == "conv_like":
        ks, dil = cfg.conv_kernel_size, cfg.conv_dilation
        pad = ((ks - 1) * dil + 1) // 2

        def unfold(t):  # [b,h,P,dh] -> [b,h,P,ks*ks,dh]
            t = t.reshape(b * h, fmap, fmap, dh).permute(0, 3, 1, 2)
            u = F.unfold(t.float(), ks, padding=pad, dilation=dil).to(t.dtype)  # [bh, dh*ks*ks, P]
            return u.reshape(b, h, dh, ks * ks, P).permute(0, 1, 4, 3, 2)

        ku, vu = unfold(ki), unfold(vi)
        d_loc = torch.einsum("bhid,bhijd->bhij", qi, ku)
        kidx = torch.arange(P, dtype=torch.float32, device=x.device).reshape(1, 1, fmap, fmap)
        kidx = F.pad(kidx, (pad,) * 4, value=float(P))
        kidx = F.unfold(kidx, ks, dilation=dil)[0].t()  # [P, ks*ks]
        qidx = torch.arange(P, dtype=torch.float32, device=x.device)[:, None]
        d_loc = d_loc.masked_fill(qidx < kidx, neg)
        dots = torch.cat((d_i2t, d_loc), dim=-1)
        attn = dots.float().softmax(-1).to(dots.dtype)
        a_txt, a_loc = attn[..., :text_len], attn[..., text_len:]
        out_img = torch.einsum("bhij,bhijd->bhid", a_loc, vu) + torch.einsum("bhij,bhjd->bhid", a_txt, vt)
    else:
        raise ValueError(attn_type)
    out = torch.cat((out_text, out_img), dim=2)[:, :, :n]
    out = out.permute(0, 2, 1, 3).reshape(b, n, cfg.inner_dim)
    return F.linear(out, w_out, b_out)


# ------------------------------------------------------------------------------------------------
# blocks (M2, M4, M5, M14)
# ------------------------------------------------------------------------------------------------
def feed_forward(x, w1, b1, w2, b2):
    h = F.linear(x, w1, b1)
    val, gate = h.chunk(2, dim=-1)
    return F.linear(val * F.gelu(gate), w2, b2)


def block_f(model, i: int, x, build: str):
    """LayerScale(PreNorm(PreShiftToken(attn_i)))(x)."""
    cfg = model.cfg
    a = model.attn_index[i]
    h = F.layer_norm(x, (cfg.dim,), model.norm_f_w[i].to(x.dtype), model.norm_f_b[i].to(x.dtype), cfg.ln_eps)
    if cfg.shift_tokens:
        h = shift_tokens(h, cfg)
    fn = attention_sparse if build == "sparse" else attention_dense
    angles = model.rotary_angles if cfg.rotary_emb else None
    h = fn(h, model.attn_qkv_w[a].to(x.dtype), model.attn_out_w[a].to(x.dtype), model.attn_out_b[a].to(x.dtype), cfg,
           cfg.attn_types[i], angles)
    return h * model.scale_f[i].to(x.dtype)


def block_g(model, i: int, x):
    cfg = model.cfg
    f = model.ff_index[i]
    h = F.layer_norm(x, (cfg.dim,), model.norm_g_w[i].to(x.dtype), model.norm_g_b[i].to(x.dtype), cfg.ln_eps)
    if cfg.shift_tokens:
        h = shift_tokens(h, cfg)
    h = feed_forward(h, model.ff_w1[f].to(x.dtype), model.ff_b1[f].to(x.dtype), model.ff_w2[f].to(x.dtype),
                     model.ff_b2[f].to(x.dtype))
    return h * model.scale_g[i].to(x.dtype)


class _ReversibleOracle(torch.autograd.Function):
    """RevNet execution with input reconstruction (M3): O(1) activation memory in depth."""

    @staticmethod
    def forward(ctx, x, model, build, *params):
        ctx.model, ctx.build = model, build
        x1, x2 = x.chunk(2, dim=-1)
        with torch.no_grad():
            for i in range(model.cfg.depth):
                x1 = x1 + block_f(model, i, x2, build)
                x2 = x2 + block_g(model, i, x1)
        ctx.save_for_backward(x1.detach(), x2.detach())
        return torch.cat((x1, x2), dim=-1)

    @staticmethod
    def backward(ctx, dy):
        model, build = ctx.model, ctx.build
        y1, y2 = ctx.saved_tensors
        dy1, dy2 = dy.chunk(2, dim=-1)
        params = [p for p in model.parameters() if p.requires_grad]
        grads = [torch.zeros_like(p) for p in params]
        for i in reversed(range(model.cfg.depth)):
            with torch.enable_grad():
                y1_ = y1.detach().requires_grad_(True)
                gy1 = block_g(model, i, y1_)
            gs = torch.autograd.grad(gy1, [y1_] + params, dy2, allow_unused=True)
            with torch.no_grad():
                x2 = y2 - gy1
                dx1 = dy1 + gs[0]
                for g, acc in zip(gs[1:], grads):
                    if g is not None:
                        acc += g
            with torch.enable_grad():
                x2_ = x2.detach().requires_grad_(True)
                fx2 = block_f(model, i, x2_, build)
            gs = torch.autograd.grad(fx2, [x2_] + params, dx1, allow_unused=True)
            with torch.no_grad():
                x1 = y1 - fx2
                dx2 = dy2 + gs[0]
                for g, acc in zip(gs[1:], grads):
                    if g is not None:
                        acc += g
            y1, y2, dy1, dy2 = x1, x2, dx1, dx2
        return (torch.cat((dy1, dy2), dim=-1), None, None, *grads)


def transformer(model, x, build: str = "sparse", reversible_autograd: bool = False):
    cfg = model.cfg
    if cfg.reversible:
        if reversible_autograd:
            params = [p for p in model.parameters() if p.requires_grad]
            y = _ReversibleOracle.apply(torch.cat((x, x), dim=-1), model, build, *params)
            y1, y2 = y.chunk(2, dim=-1)
        else:
            y1, y2 = x, x
            for i in range(cfg.depth):
                y1 = y1 + block_f(model, i, y2, build)
                y2 = y2 + block_g(model, i, y1)
        return (y1 + y2) * 0.5
    for i in range(cfg.depth):
        x = x + block_f(model, i, x, build)
        x = x + block_g(model, i, x)
    return x


# ------------------------------------------------------------------------------------------------
# model level (M1)
# ------------------------------------------------------------------------------------------------
def embed(model, text: torch.Tensor, image: Optional[torch.Tensor], dtype=None):
    """Unique-pad remap, <bos>, tied-embedding gather, concat, trim to seq_len.  Returns (tokens, text_with_bos)."""
    cfg = model.cfg
    text = text[:, : cfg.text_seq_len]
    pad_ids = torch.arange(cfg.text_seq_len, device=text.device) + cfg.num_text_tokens
    text = torch.where(text == 0, pad_ids, text)
    text = F.pad(text, (1, 0), value=0)
    w = model.emb_weight if dtype is None else model.emb_weight.to(dtype)
    tokens = F.embedding(text, w[: cfg.total_text_tokens])
    if image is not None and image.numel() > 0:
        tokens = torch.cat((tokens, F.embedding(image, w[cfg.total_text_tokens:])), dim=1)
    if tokens.shape[1] > cfg.seq_len:
        tokens = tokens[:, : cfg.seq_len]
    return tokens, text


def logits_mask(cfg: DalleConfig, n: int, device) -> torch.Tensor:
    """True where a logit must be masked: text positions may only predict text ids, image positions image ids."""
    seq = torch.arange(n, device=device)[:, None]
    voc = torch.arange(cfg.total_tokens, device=device)[None, :]
    return ((seq >= cfg.text_seq_len) & (voc < cfg.total_text_tokens)) | (
        (seq < cfg.text_seq_len) & (voc >= cfg.total_text_tokens))


def to_logits(model, h: torch.Tensor) -> torch.Tensor:
    cfg = model.cfg
    h = F.layer_norm(h, (cfg.dim,), model.final_norm_w.to(h.dtype), model.final_norm_b.to(h.dtype), cfg.ln_eps)
    logits = F.linear(h, model.emb_weight.to(h.dtype), model.emb_bias.to(h.dtype))
    return logits.masked_fill(logits_mask(cfg, logits.shape[1], logits.device), -torch.finfo(logits.dtype).max)


def forward(model, text, image=None, return_loss: bool = False, build: str = "sparse", dtype=None,
            reversible_autograd: bool = False):
    cfg = model.cfg
    tokens, text_b = embed(model, text, image, dtype)
    h = transformer(model, tokens, build, reversible_autograd)
    logits = to_logits(model, h)
    if not return_loss:
        return logits
    labels = torch.cat((text_b[:, 1:], image + cfg.total_text_tokens), dim=1)
    lt = F.cross_entropy(logits[:, : cfg.text_seq_len].float().transpose(1, 2), labels[:, : cfg.text_seq_len])
    li = F.cross_entropy(logits[:, cfg.text_seq_len:].float().transpose(1, 2), labels[:, cfg.text_seq_len:])
    return (lt + cfg.loss_img_weight * li) / (cfg.loss_img_weight + 1)
