"""DALLE - the model object users hold (replaces ``dalle_pytorch.DALLE`` as configured at reference task.py:62-83).

One set of parameters, stored per UNIQUE module (5 shared attention + 5 shared feed-forward blocks for the TTT
config; LayerNorm / LayerScale are per layer; one tied embedding/logits matrix), three ways to run them:

* ``forward``            - the product path: fused sm_100a kernels driven by the reversible layer executor
                           (CPU tensors run the same executor on the PyTorch oracle ops).
* ``forward_reference``  - the naive PyTorch oracle (models/reference.py), sparse or dense-mask build.
* ``generate_images``    - KV-cached autoregressive sampling (models/generate.py).

``state_dict()`` / ``load_state_dict()`` speak the dalle-pytorch key schema, including the
aliases under every layer index and both the training and the inference spelling.
"""
from __future__ import annotations

import math
import re
from collections import OrderedDict
from typing import Dict, Iterator, List, Optional, Tuple

import torch
import torch.nn as nn

from ..config import DalleConfig
from . import reference
from .rotary import rotary_angles, rotary_pair_table


# inference build: two extra CachedAs/NonCached wrappers -> 4-5 "fn." levels instead of 3
_DEEP_FN = re.compile(r"\.net\.(?:fn\.){3,}(to_qkv|to_out|net\.)")


class DALLE(nn.Module):
    def __init__(self, cfg: DalleConfig, device=None, seed: Optional[int] = None):
        super().__init__()
        self.cfg = cfg
        dim, inner = cfg.dim, cfg.inner_dim
        gen = torch.Generator().manual_seed(seed) if seed is not None else None

        def linear_w(out_f, in_f):
            w = torch.empty(out_f, in_f)
            bound = 1.0 / math.sqrt(in_f)  # == kaiming_uniform(a=sqrt(5)), the nn.Linear default
            return nn.Parameter(w.uniform_(-bound, bound, generator=gen))

        def linear_b(out_f, in_f):
            bound = 1.0 / math.sqrt(in_f)
            return nn.Parameter(torch.empty(out_f).uniform_(-bound, bound, generator=gen))

        self.attn_ids, self.ff_ids = list(cfg.unique_attn_ids), list(cfg.unique_ff_ids)
        self.attn_index = [self.attn_ids.index(i) for i in cfg.shared_attn_ids]
        self.ff_index = [self.ff_ids.index(i) for i in cfg.shared_ff_ids]

        # tied input embedding / output projection (SharedEmbedding over to_logits[1])
        self.emb_weight = linear_w(cfg.total_tokens, dim)
        self.emb_bias = linear_b(cfg.total_tokens, dim)
        self.final_norm_w = nn.Parameter(torch.ones(dim))
        self.final_norm_b = nn.Parameter(torch.zeros(dim))

        self.attn_qkv_w = nn.ParameterList([linear_w(3 * inner, dim) for _ in self.attn_ids])
        self.attn_out_w = nn.ParameterList([linear_w(dim, inner) for _ in self.attn_ids])
        self.attn_out_b = nn.ParameterList([linear_b(dim, inner) for _ in self.attn_ids])
        hid = dim * cfg.ff_mult
        self.ff_w1 = nn.ParameterList([linear_w(2 * hid, dim) for _ in self.ff_ids])
        self.ff_b1 = nn.ParameterList([linear_b(2 * hid, dim) for _ in self.ff_ids])
        self.ff_w2 = nn.ParameterList([linear_w(dim, hid) for _ in self.ff_ids])
        self.ff_b2 = nn.ParameterList([linear_b(dim, hid) for _ in self.ff_ids])

        def per_layer(make):
            return nn.ParameterList([nn.Parameter(make(i)) for i in range(cfg.depth)])

        self.scale_f = per_layer(lambda i: torch.full((1, 1, dim), cfg.layer_scale_init(i)))
        self.norm_f_w = per_layer(lambda i: torch.ones(dim))
        self.norm_f_b = per_layer(lambda i: torch.zeros(dim))
        self.scale_g = per_layer(lambda i: torch.full((1, 1, dim), cfg.layer_scale_init(i)))
        self.norm_g_w = per_layer(lambda i: torch.ones(dim))
        self.norm_g_b = per_layer(lambda i: torch.zeros(dim))

        self.register_buffer("rotary_angles", rotary_angles(cfg), persistent=False)       # [seq_len+1, 60]
        self.register_buffer("rotary_pairs", rotary_pair_table(cfg), persistent=False)    # [seq_len+1, 32, 2]
        self._engine = None
        self.vae = None
        if device is not None:
            self.to(device)

    # ------------------------------------------------------------------------------------------
    # parameter naming: dalle-pytorch order and spelling (Appendix A.1 / A.2)
    # ------------------------------------------------------------------------------------------
    def reference_named_parameters(self) -> Iterator[Tuple[str, nn.Parameter]]:
        """(name, param) in the order ``named_parameters()`` of the reference model yields them (shared tensors
        once, at first occurrence).  This order defines the optimizer-state indices of optimizer_state.pt."""
        cfg = self.cfg
        yield "text_emb.linear.weight", self.emb_weight
        yield "text_emb.linear.bias", self.emb_bias
        seen_a, seen_f = set(), set()
        for i in range(cfg.depth):
            p = f"transformer.layers.blocks.{i}."
            a, f = self.attn_index[i], self.ff_index[i]
            yield p + "f.net.scale", self.scale_f[i]
            yield p + "f.net.fn.norm.weight", self.norm_f_w[i]
            yield p + "f.net.fn.norm.bias", self.norm_f_b[i]
            if a not in seen_a:
                seen_a.add(a)
                yield p + "f.net.fn.fn.fn.to_qkv.weight", self.attn_qkv_w[a]
                yield p + "f.net.fn.fn.fn.to_out.0.weight", self.attn_out_w[a]
                yield p + "f.net.fn.fn.fn.to_out.0.bias", self.attn_out_b[a]
            yield p + "g.net.scale", self.scale_g[i]
            yield p + "g.net.fn.norm.weight", self.norm_g_w[i]
            yield p + "g.net.fn.norm.bias", self.norm_g_b[i]
            if f not in seen_f:
                seen_f.add(f)
                yield p + "g.net.fn.fn.fn.net.0.weight", self.ff_w1[f]
                yield p + "g.net.fn.fn.fn.net.0.bias", self.ff_b1[f]
                yield p + "g.net.fn.fn.fn.net.3.weight", self.ff_w2[f]
                yield p + "g.net.fn.fn.fn.net.3.bias", self.ff_b2[f]
        yield "to_logits.0.weight", self.final_norm_w
        yield "to_logits.0.bias", self.final_norm_b

    def _all_keys(self) -> "OrderedDict[str, torch.Tensor]":
        """Every key of the dalle-pytorch training-build state dict, aliases included."""
        cfg = self.cfg
        sd: "OrderedDict[str, torch.Tensor]" = OrderedDict()
        for nm in ("text_emb", "image_emb"):
            sd[f"{nm}.linear.weight"] = self.emb_weight
            sd[f"{nm}.linear.bias"] = self.emb_bias
        if cfg.rotary_emb:
            sd["transformer.pos_emb"] = self.rotary_angles[None]
        for i in range(cfg.depth):
            p = f"transformer.layers.blocks.{i}."
            a, f = self.attn_index[i], self.ff_index[i]
            sd[p + "f.net.scale"] = self.scale_f[i]
            sd[p + "f.net.fn.norm.weight"] = self.norm_f_w[i]
            sd[p + "f.net.fn.norm.bias"] = self.norm_f_b[i]
            sd[p + "f.net.fn.fn.fn.to_qkv.weight"] = self.attn_qkv_w[a]
            sd[p + "f.net.fn.fn.fn.to_out.0.weight"] = self.attn_out_w[a]
            sd[p + "f.net.fn.fn.fn.to_out.0.bias"] = self.attn_out_b[a]
            sd[p + "g.net.scale"] = self.scale_g[i]
            sd[p + "g.net.fn.norm.weight"] = self.norm_g_w[i]
            sd[p + "g.net.fn.norm.bias"] = self.norm_g_b[i]
            sd[p + "g.net.fn.fn.fn.net.0.weight"] = self.ff_w1[f]
            sd[p + "g.net.fn.fn.fn.net.0.bias"] = self.ff_b1[f]
            sd[p + "g.net.fn.fn.fn.net.3.weight"] = self.ff_w2[f]
            sd[p + "g.net.fn.fn.fn.net.3.bias"] = self.ff_b2[f]
        sd["to_logits.0.weight"] = self.final_norm_w
        sd["to_logits.0.bias"] = self.final_norm_b
        sd["to_logits.1.weight"] = self.emb_weight
        sd["to_logits.1.bias"] = self.emb_bias
        return sd

    def state_dict(self, *args, destination=None, prefix: str = "", keep_vars: bool = False):
        out = OrderedDict() if destination is None else destination
        for k, v in self._all_keys().items():
            out[prefix + k] = v if keep_vars else v.detach()
        return out

    @staticmethod
    def canonical_key(key: str) -> str:
        """Map the inference-build spelling (reference inference/run_inference.py:117) back to the training one."""
        return _DEEP_FN.sub(r".net.fn.fn.fn.\1", key)

    def load_state_dict(self, state_dict, strict: bool = True, assign: bool = False):
        own = self._all_keys()
        missing, unexpected, loaded = [], [], set()
        with torch.no_grad():
            for k, v in state_dict.items():
                ck = self.canonical_key(k)
                if ck not in own:
                    unexpected.append(k)
                    continue
                tgt = own[ck]
                if ck == "transformer.pos_emb":
                    loaded.add(ck)
                    continue  # deterministic buffer; regenerated from the config
                if tuple(tgt.shape) != tuple(v.shape):
                    raise RuntimeError(f"size mismatch for {k}: checkpoint {tuple(v.shape)} vs model {tuple(tgt.shape)}")
                if id(tgt) not in loaded:
                    tgt.copy_(v)
                loaded.add(ck)
                loaded.add(id(tgt))
        tensors_loaded = {i for i in loaded if isinstance(i, int)}
        for k, t in own.items():
            if id(t) not in tensors_loaded and k != "transformer.pos_emb":
                missing.append(k)
        if strict and (missing or unexpected):
            raise RuntimeError(f"load_state_dict: missing {missing[:5]}... unexpected {unexpected[:5]}...")
        if self._engine is not None:
            self._engine.params_changed()
        return torch.nn.modules.module._IncompatibleKeys(missing, unexpected)

    # ------------------------------------------------------------------------------------------
    # execution
    # ------------------------------------------------------------------------------------------
    def forward_reference(self, text, image=None, mask=None, return_loss: bool = False, build: str = "sparse",
                          dtype=None, reversible_autograd: bool = False):
        """Oracle path. ``mask`` is accepted and ignored exactly like the reference model does (task.py:41)."""
        return reference.forward(self, text, image, return_loss, build, dtype, reversible_autograd)

    @property
    def engine(self):
        if self._engine is None:
            from .executor import Engine

            self._engine = Engine(self)
        return self._engine

    def forward(self, text, image=None, mask=None, return_loss: bool = False):
        """Product path: fused kernels + reversible executor.  ``mask`` accepted and ignored (reference quirk)."""
        return self.engine.forward(text, image, return_loss)

    @torch.no_grad()
    def generate_images(self, text, *, temperature: float = 1.0, top_k: int = 0, top_p: float = 1.0,
                        use_cache: bool = True, generator: Optional[torch.Generator] = None, return_codes=False):
        from .generate import generate_images

        return generate_images(self, text, temperature=temperature, top_k=top_k, top_p=top_p, use_cache=use_cache,
                               generator=generator, return_codes=return_codes)

    def num_unique_parameters(self) -> int:
        return sum(p.numel() for p in self.parameters())


class ModelWrapper(nn.Module):
    """``{'loss': ...}`` adapter with the ``model.`` key prefix of the reference checkpoints (task.py:35-42)."""

    def __init__(self, model: DALLE):
        super().__init__()
        self.model = model

    def forward(self, input_ids, attention_mask=None, image=None):
        loss = self.model(text=input_ids, image=image, mask=attention_mask, return_loss=True)
        return {"loss": loss}

    def state_dict(self, *args, destination=None, prefix: str = "", keep_vars: bool = False):
        return self.model.state_dict(destination=destination, prefix=prefix + "model.", keep_vars=keep_vars)

    def load_state_dict(self, state_dict, strict: bool = True, assign: bool = False):
        stripped = OrderedDict((k[len("model."):] if k.startswith("model.") else k, v) for k, v in state_dict.items())
        return self.model.load_state_dict(stripped, strict=strict)
