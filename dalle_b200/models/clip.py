"""CLIP ViT-B/32 scorer (``clip.load("ViT-B/32")`` + ``logits_per_text.softmax`` at reference
inference/run_inference.py:126, 135-138).  Restated architecture with OpenAI's state-dict key names so the public
checkpoint loads; the BPE tokenizer needs OpenAI's vocabulary file (``bpe_path``).  Not on the transformer hot path.
"""
from __future__ import annotations

import gzip
import html
import re
from collections import OrderedDict
from functools import lru_cache
from typing import List, Optional, Union

import torch
import torch.nn as nn
import torch.nn.functional as F


class QuickGELU(nn.Module):
    def forward(self, x):
        return x * torch.sigmoid(1.702 * x)


class ResidualAttentionBlock(nn.Module):
    def __init__(self, d_model: int, n_head: int, attn_mask: Optional[torch.Tensor] = None):
        super().__init__()
        self.attn = nn.MultiheadAttention(d_model, n_head)
        self.ln_1 = nn.LayerNorm(d_model)
        self.mlp = nn.Sequential(OrderedDict([("c_fc", nn.Linear(d_model, d_model * 4)), ("gelu", QuickGELU()),
                                              ("c_proj", nn.Linear(d_model * 4, d_model))]))
        self.ln_2 = nn.LayerNorm(d_model)
        self.attn_mask = attn_mask

    def forward(self, x):
        m = self.attn_mask.to(dtype=x.dtype, device=x.device) if self.attn_mask is not None else None
        y = self.ln_1(x)
        x = x + self.attn(y, y, y, need_weights=False, attn_mask=m)[0]
        return x + self.mlp(self.ln_2(x))


class Transformer(nn.Module):
    def __init__(self, width, layers, heads, attn_mask=None):
        super().__init__()
        self.resblocks = nn.Sequential(*[ResidualAttentionBlock(width, heads, attn_mask) for _ in range(layers)])

    def forward(self, x):
        return self.resblocks(x)


class VisionTransformer(nn.Module):
    def __init__(self, input_resolution=224, patch_size=32, width=768, layers=12, heads=12, output_dim=512):
        super().__init__()
        self.conv1 = nn.Conv2d(3, width, patch_size, patch_size, bias=False)
        scale = width ** -0.5
        self.class_embedding = nn.Parameter(scale * torch.randn(width))
        self.positional_embedding = nn.Parameter(scale * torch.randn((input_resolution // patch_size) ** 2 + 1, width))
        self.ln_pre, self.transformer, self.ln_post = nn.LayerNorm(width), Transformer(width, layers, heads), nn.LayerNorm(width)
        self.proj = nn.Parameter(scale * torch.randn(width, output_dim))

    def forward(self, x):
        x = self.conv1(x).flatten(2).permute(0, 2, 1)
        cls = self.class_embedding.to(x.dtype).expand(x.shape[0], 1, -1)
        x = self.ln_pre(torch.cat([cls, x], dim=1) + self.positional_embedding.to(x.dtype))
        x = self.transformer(x.permute(1, 0, 2)).permute(1, 0, 2)
        return self.ln_post(x[:, 0]) @ self.proj


class CLIP(nn.Module):
    def __init__(self, embed_dim=512, image_resolution=224, vision_layers=12, vision_width=768, vision_patch_size=32,
                 context_length=77, vocab_size=49408, transformer_width=512, transformer_heads=8, transformer_layers=12):
        super().__init__()
        self.context_length = context_length
        self.visual = VisionTransformer(image_resolution, vision_patch_size, vision_width, vision_layers,
                                        vision_width // 64, embed_dim)
        mask = torch.full((context_length, context_length), float("-inf")).triu_(1)
        self.transformer = Transformer(transformer_width, transformer_layers, transformer_heads, mask)
        self.token_embedding = nn.Embedding(vocab_size, transformer_width)
        self.positional_embedding = nn.Parameter(torch.empty(context_length, transformer_width).normal_(std=0.01))
        self.ln_final = nn.LayerNorm(transformer_width)
        self.text_projection = nn.Parameter(torch.empty(transformer_width, embed_dim).normal_(std=transformer_width ** -0.5))
        self.logit_scale = nn.Parameter(torch.ones([]) * 2.6592)

    def encode_image(self, image):
        return self.visual(image)

    def encode_text(self, text):
        x = self.token_embedding(text) + self.positional_embedding
        x = self.ln_final(self.transformer(x.permute(1, 0, 2)).permute(1, 0, 2))
        return x[torch.arange(x.shape[0]), text.argmax(dim=-1)] @ self.text_projection

    def forward(self, image, text):
        i = F.normalize(self.encode_image(image), dim=-1)
        t = F.normalize(self.encode_text(text), dim=-1)
        logits_per_image = self.logit_scale.exp() * i @ t.t()
        return logits_per_image, logits_per_image.t()


_MEAN = (0.48145466, 0.4578275, 0.40821073)
_STD = (0.26862954, 0.26130258, 0.27577711)


def preprocess(images: torch.Tensor, resolution: int = 224) -> torch.Tensor:
    """images [n, h, w, 3] or [n, 3, h, w] floats in [0, 1] -> normalised [n, 3, res, res] (bicubic + centre crop)."""
    if images.shape[-1] == 3:
        images = images.permute(0, 3, 1, 2)
    h, w = images.shape[-2:]
    s = resolution / min(h, w)
    images = F.interpolate(images.float(), size=(round(h * s), round(w * s)), mode="bicubic", align_corners=False)
    top, left = (images.shape[-2] - resolution) // 2, (images.shape[-1] - resolution) // 2
    images = images[..., top: top + resolution, left: left + resolution].clamp(0, 1)
    mean, std = (torch.tensor(v, device=images.device).view(1, 3, 1, 1) for v in (_MEAN, _STD))
    return (images - mean) / std


@lru_cache()
def _bytes_to_unicode():
    bs = list(range(ord("!"), ord("~") + 1)) + list(range(ord("¡"), ord("¬") + 1)) + list(range(ord("®"), ord("ÿ") + 1))
    cs, n = bs[:], 0
    for b in range(256):
        if b not in bs:
            bs.append(b)
            cs.append(256 + n)
            n += 1
    return dict(zip(bs, (chr(c) for c in cs)))


class SimpleTokenizer:
    """OpenAI CLIP byte-level BPE (needs bpe_simple_vocab_16e6.txt.gz)."""

    def __init__(self, bpe_path: str):
        self.byte_encoder = _bytes_to_unicode()
        merges = gzip.open(bpe_path).read().decode("utf-8").split("\n")[1: 49152 - 256 - 2 + 1]
        merges = [tuple(m.split()) for m in merges]
        vocab = list(self.byte_encoder.values())
        vocab += [v + "</w>" for v in vocab] + ["".join(m) for m in merges] + ["<|startoftext|>", "<|endoftext|>"]
        self.encoder = dict(zip(vocab, range(len(vocab))))
        self.bpe_ranks = dict(zip(merges, range(len(merges))))
        self.cache = {"<|startoftext|>": "<|startoftext|>", "<|endoftext|>": "<|endoftext|>"}
        self.pat = re.compile(r"<\|startoftext\|>|<\|endoftext\|>|'s|'t|'re|'ve|'m|'ll|'d|[a-zA-Z]+|[0-9]|[^\sa-zA-Z0-9]+", re.I)

    def bpe(self, token):
        if token in self.cache:
            return self.cache[token]
        word = tuple(token[:-1]) + (token[-1] + "</w>",)
        while len(word) > 1:
            pairs = set(zip(word[:-1], word[1:]))
            bigram = min(pairs, key=lambda p: self.bpe_ranks.get(p, float("inf")))
            if bigram not in self.bpe_ranks:
                break
            first, second = bigram
            new, i = [], 0
            while i < len(word):
                if i < len(word) - 1 and word[i] == first and word[i + 1] == second:
                    new.append(first + second)
                    i += 2
                else:
                    new.append(word[i])
                    i += 1
            word = tuple(new)
        self.cache[token] = " ".join(word)
        return self.cache[token]

    def encode(self, text: str) -> List[int]:
        text = re.sub(r"\s+", " ", html.unescape(html.unescape(text))).strip().lower()
        out = []
        for tok in re.findall(self.pat, text):
            tok = "".join(self.byte_encoder[b] for b in tok.encode("utf-8"))
            out.extend(self.encoder[t] for t in self.bpe(tok).split(" "))
        return out


def tokenize(tokenizer: SimpleTokenizer, texts: Union[str, List[str]], context_length: int = 77) -> torch.Tensor:
    texts = [texts] if isinstance(texts, str) else texts
    sot, eot = tokenizer.encoder["<|startoftext|>"], tokenizer.encoder["<|endoftext|>"]
    out = torch.zeros(len(texts), context_length, dtype=torch.long)
    for i, t in enumerate(texts):
        ids = [sot] + tokenizer.encode(t)[: context_length - 2] + [eot]
        out[i, : len(ids)] = torch.tensor(ids)
    return out


def load(checkpoint_path: str, device="cuda") -> CLIP:
    """Load OpenAI's ViT-B/32 weights (a state dict, or the TorchScript archive's state dict)."""
    try:
        sd = torch.jit.load(checkpoint_path, map_location="cpu").state_dict()
    except Exception:
        sd = torch.load(checkpoint_path, map_location="cpu", weights_only=False)
        sd = sd.get("state_dict", sd)
    model = CLIP()
    own = model.state_dict()
    model.load_state_dict({k: v.float() for k, v in sd.items() if k in own}, strict=False)
    return model.to(device).eval()
