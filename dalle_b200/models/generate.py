"""KV-cached autoregressive image sampling (``DALLE.generate_images(..., use_cache=True)`` as
called at reference inference/run_inference.py:87-91).

Two executors with identical semantics:
* ``decode_step_reference`` - plain PyTorch over explicit caches (CPU/GPU): per layer a K/V cache, and per wrapped
  block the token-shift cache (the last ``fmap`` normalised tokens' shifted channel halves).  It is the oracle: its
  logits equal a full re-forward at every position (tests/test_generate.py).
* the GPU path (``DecodeEngine`` in models/decode_engine.py) - the same step as sm_100a kernels captured in a CUDA
  graph; used automatically for CUDA tensors when ``dim_head == 64``.

Sampling: logits of the last position -> (image segment only) -> top-k (if > 0) -> nucleus top-p (if < 1), both on
the raw logits -> divide the kept logits by the temperature -> multinomial draw (dalle-pytorch's order; the CUDA
sampler in ops/csrc/decode.cu does the same).
"""
from __future__ import annotations

from typing import List, Optional

import torch
import torch.nn.functional as F

from ..config import DalleConfig
from . import reference as R
from .masks import static_mask
from .rotary import apply_rotary


def filter_logits(logits: torch.Tensor, top_k: int = 0, top_p: float = 1.0) -> torch.Tensor:
    """Keep the k largest logits (k > 0) and/or the smallest set whose probability mass reaches top_p."""
    neg = float("-inf")
    if top_k and top_k > 0:
        k = min(top_k, logits.shape[-1])
        kth = logits.topk(k, dim=-1).values[..., -1:]
        logits = logits.masked_fill(logits < kth, neg)
    if top_p < 1.0:
        sorted_logits, idx = logits.sort(dim=-1, descending=True)
        probs = sorted_logits.softmax(dim=-1)
        cum = probs.cumsum(dim=-1)
        remove = cum - probs > top_p          # always keeps the most likely token
        sorted_logits = sorted_logits.masked_fill(remove, neg)
        logits = torch.full_like(logits, neg).scatter(-1, idx, sorted_logits)
    return logits


def sample_from_logits(logits: torch.Tensor, temperature: float = 1.0, top_k: int = 0, top_p: float = 1.0,
                       generator: Optional[torch.Generator] = None) -> torch.Tensor:
    logits = filter_logits(logits.float(), top_k, top_p) / max(temperature, 1e-6)
    probs = logits.softmax(dim=-1)
    return torch.multinomial(probs, 1, generator=generator).squeeze(-1)


class DecodeCache:
    """Per-layer K/V (rotated, q-scale NOT applied) and token-shift caches for incremental decoding."""

    def __init__(self, cfg: DalleConfig, batch: int, device, dtype):
        self.pos = 0
        n, h, dh = cfg.seq_len, cfg.heads, cfg.dim_head
        self.k = [torch.zeros(batch, h, n, dh, device=device, dtype=dtype) for _ in range(cfg.depth)]
        self.v = [torch.zeros(batch, h, n, dh, device=device, dtype=dtype) for _ in range(cfg.depth)]
        # normalised (pre-shift) inputs of every wrapped block: [depth][f|g] -> [batch, n, dim/2] (shifted halves)
        self.norm = [[torch.zeros(batch, n, cfg.dim // 2, device=device, dtype=dtype) for _ in range(2)]
                     for _ in range(cfg.depth)]


def _shift_one(cfg: DalleConfig, cache_norm: torch.Tensor, n_new: torch.Tensor, pos: int) -> torch.Tensor:
    """Token shift of the single position ``pos`` given the cache of earlier normalised tokens."""
    b, d = n_new.shape
    half, q = d // 2, d // 4
    cache_norm[:, pos] = n_new[:, :half]
    out = n_new.clone()
    T, Fm = cfg.text_len, cfg.image_fmap_size
    if pos < T:
        out[:, :half] = cache_norm[:, pos - 1] if pos >= 1 else 0
    else:
        p = pos - T
        r, c = divmod(p, Fm)
        out[:, :q] = cache_norm[:, pos - Fm, :q] if r >= 1 else 0
        out[:, q:half] = cache_norm[:, pos - 1, q:half] if c >= 1 else 0
    return out


def _allowed_keys(cfg: DalleConfig, attn_type: str, device) -> torch.Tensor:
    return static_mask(cfg, attn_type, cfg.seq_len).to(device)


@torch.no_grad()
def decode_step_reference(model, x: torch.Tensor, cache: DecodeCache) -> torch.Tensor:
    """x: embeddings of the token at position cache.pos, [batch, dim].  Returns the hidden state (mean of streams)."""
    cfg: DalleConfig = model.cfg
    pos, h, dh = cache.pos, cfg.heads, cfg.dim_head
    dt = x.dtype
    x1 = x2 = x
    ang = model.rotary_angles[pos: pos + 1].to(x.device)
    for i in range(cfg.depth):
        a, f = model.attn_index[i], model.ff_index[i]
        # ---- f: attention ----
        n = F.layer_norm(x2, (cfg.dim,), model.norm_f_w[i].to(dt), model.norm_f_b[i].to(dt), cfg.ln_eps)
        hcur = _shift_one(cfg, cache.norm[i][0], n, pos)
        qkv = F.linear(hcur, model.attn_qkv_w[a].to(dt)).view(-1, 3, h, dh)
        q, k, v = (apply_rotary(t.unsqueeze(2), ang).squeeze(2) for t in (qkv[:, 0], qkv[:, 1], qkv[:, 2]))
        cache.k[i][:, :, pos], cache.v[i][:, :, pos] = k, v
        allowed = _allowed_keys(cfg, cfg.attn_types[i], x.device)[pos, : pos + 1]
        dots = torch.einsum("bhd,bhjd->bhj", q * dh ** -0.5, cache.k[i][:, :, : pos + 1]).float()
        dots = dots.masked_fill(~allowed, float("-inf"))
        o = torch.einsum("bhj,bhjd->bhd", dots.softmax(-1).to(dt), cache.v[i][:, :, : pos + 1]).reshape(-1, h * dh)
        o = F.linear(o, model.attn_out_w[a].to(dt), model.attn_out_b[a].to(dt))
        x1 = x1 + o * model.scale_f[i].view(-1).to(dt)
        # ---- g: feed-forward ----
        n = F.layer_norm(x1, (cfg.dim,), model.norm_g_w[i].to(dt), model.norm_g_b[i].to(dt), cfg.ln_eps)
        hcur = _shift_one(cfg, cache.norm[i][1], n, pos)
        o = R.feed_forward(hcur, model.ff_w1[f].to(dt), model.ff_b1[f].to(dt), model.ff_w2[f].to(dt), model.ff_b2[f].to(dt))
        x2 = x2 + o * model.scale_g[i].view(-1).to(dt)
    cache.pos += 1
    return (x1 + x2) * 0.5


@torch.no_grad()
def logits_from_hidden(model, hidden: torch.Tensor, image_only: bool) -> torch.Tensor:
    cfg = model.cfg
    dt = hidden.dtype
    hn = F.layer_norm(hidden, (cfg.dim,), model.final_norm_w.to(dt), model.final_norm_b.to(dt), cfg.ln_eps)
    W, b = model.emb_weight.to(dt), model.emb_bias.to(dt)
    if image_only:
        return F.linear(hn, W[cfg.total_text_tokens:], b[cfg.total_text_tokens:])
    return F.linear(hn, W, b)


@torch.no_grad()
def generate_codes_reference(model, text: torch.Tensor, *, temperature=1.0, top_k=0, top_p=1.0, use_cache=True,
                             generator=None, dtype=None, return_logits: bool = False):
    """-> image codes [batch, image_seq_len] (int64).  PyTorch executor (oracle; also the eager baseline B3)."""
    cfg: DalleConfig = model.cfg
    dev = text.device
    dt = dtype or (torch.bfloat16 if text.is_cuda else torch.float32)
    text = text[:, : cfg.text_seq_len]
    B = text.shape[0]
    image = torch.zeros(B, 0, dtype=torch.int64, device=dev)
    all_logits: List[torch.Tensor] = []
    if use_cache:
        cache = DecodeCache(cfg, B, dev, dt)
        tokens, _ = R.embed(model, text, None, dt)          # [B, text_len, dim] (<bos> + text)
        hidden = None
        for t in range(cfg.text_len):
            hidden = decode_step_reference(model, tokens[:, t], cache)
    for step in range(cfg.image_seq_len):
        if use_cache:
            logits = logits_from_hidden(model, hidden, image_only=True)
        else:
            full = R.forward(model, text, image, False, "dense", dt)
            logits = full[:, -1, cfg.total_text_tokens:]
        if return_logits:
            all_logits.append(logits.float().cpu())
        nxt = sample_from_logits(logits, temperature, top_k, top_p, generator)
        image = torch.cat((image, nxt[:, None]), dim=1)
        if use_cache and step + 1 < cfg.image_seq_len:
            emb = model.emb_weight[cfg.total_text_tokens + nxt].to(dt)
            hidden = decode_step_reference(model, emb, cache)
    return (image, all_logits) if return_logits else image


@torch.no_grad()
def generate_images(model, text, *, temperature=1.0, top_k=0, top_p=1.0, use_cache=True, generator=None,
                    return_codes=False):
    """``DALLE.generate_images``: sample 1024 image tokens, then decode them with the attached VAE
    (``model.vae.decode``) into images [b, 3, H, W] in [0, 1].  Without a real VAE the codes are returned."""
    codes = None
    if text.is_cuda and use_cache and model.cfg.dim_head == 64:
        try:
            from .decode_engine import DecodeEngine, generate_many

            B = text.shape[0]
            if B <= 32:
                eng = getattr(model, "_decode_engine", None)
                if eng is None or eng.batch != B:
                    eng = model._decode_engine = DecodeEngine(model, B)
                codes = eng.generate(text, temperature=temperature, top_k=top_k, top_p=top_p, generator=generator)
            else:
                # larger batches: slices, one engine and one stream each, decoded concurrently.  Measured on the TTT
                # model: 64 prompts as 2 x 32 take 4.0 s (one engine with a 64-row tile: 4.6 s), 128 prompts as 2 x 64
                # 7.2 s (4 x 32: 7.4 s) - hence slices of 32 up to 64 prompts and of 64 beyond.
                cut = 32 if B <= 64 else DecodeEngine.MAX_BATCH
                sizes = [min(cut, B - o) for o in range(0, B, cut)]
                engs = getattr(model, "_decode_engines", None)
                if engs is None or [e.batch for e in engs] != sizes:
                    engs = model._decode_engines = [DecodeEngine(model, n) for n in sizes]
                parts = generate_many(engs, list(text.split(cut)), temperature=temperature,
                                      top_k=top_k, top_p=top_p, generator=generator)
                codes = torch.cat(parts, dim=0)
        except ImportError:
            codes = None
    if codes is None:
        codes = generate_codes_reference(model, text, temperature=temperature, top_k=top_k, top_p=top_p,
                                         use_cache=use_cache, generator=generator)
    vae = getattr(model, "vae", None)
    if return_codes or vae is None or not hasattr(vae, "decode"):
        return codes
    return vae.decode(codes)
