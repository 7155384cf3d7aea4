"""VQGAN (GumbelVQ, f8) decoder: image codes -> RGB (dalle-pytorch ``VQGanVAE.decode`` over
taming-transformers' ``GumbelVQ``, instantiated at reference inference/run_inference.py:122).

The architecture is restated here with taming's state-dict key names, so an upstream ``*.ckpt`` + ``*.yaml`` pair
loads directly (``VQGanVAE(ckpt_path, config_path)``).  Plain PyTorch modules (cuDNN convolutions): this runs once
per 1024 sampled tokens and is outside the transformer hot path.
"""
from __future__ import annotations

import math
from typing import Optional, Sequence

import torch
import torch.nn as nn
import torch.nn.functional as F


def _norm(c):
    return nn.GroupNorm(num_groups=32, num_channels=c, eps=1e-6, affine=True)


class ResnetBlock(nn.Module):
    def __init__(self, cin, cout):
        super().__init__()
        self.norm1, self.conv1 = _norm(cin), nn.Conv2d(cin, cout, 3, padding=1)
        self.norm2, self.conv2 = _norm(cout), nn.Conv2d(cout, cout, 3, padding=1)
        if cin != cout:
            self.nin_shortcut = nn.Conv2d(cin, cout, 1)

    def forward(self, x):
        h = self.conv1(F.silu(self.norm1(x)))
        h = self.conv2(F.silu(self.norm2(h)))
        return (self.nin_shortcut(x) if hasattr(self, "nin_shortcut") else x) + h


class AttnBlock(nn.Module):
    def __init__(self, c):
        super().__init__()
        self.norm = _norm(c)
        self.q, self.k, self.v, self.proj_out = (nn.Conv2d(c, c, 1) for _ in range(4))

    def forward(self, x):
        b, c, h, w = x.shape
        n = self.norm(x)
        q, k, v = (m(n).reshape(b, c, h * w) for m in (self.q, self.k, self.v))
        attn = torch.softmax(torch.bmm(q.transpose(1, 2), k) * c ** -0.5, dim=2)
        out = torch.bmm(v, attn.transpose(1, 2)).reshape(b, c, h, w)
        return x + self.proj_out(out)


class Upsample(nn.Module):
    def __init__(self, c):
        super().__init__()
        self.conv = nn.Conv2d(c, c, 3, padding=1)

    def forward(self, x):
        return self.conv(F.interpolate(x, scale_factor=2.0, mode="nearest"))


class Decoder(nn.Module):
    def __init__(self, *, ch=128, out_ch=3, ch_mult=(1, 1, 2, 4), num_res_blocks=2, attn_resolutions=(32,),
                 resolution=256, z_channels=256, **unused):
        super().__init__()
        self.num_resolutions, self.num_res_blocks = len(ch_mult), num_res_blocks
        block_in = ch * ch_mult[-1]
        curr_res = resolution // 2 ** (self.num_resolutions - 1)
        self.conv_in = nn.Conv2d(z_channels, block_in, 3, padding=1)
        self.mid = nn.Module()
        self.mid.block_1, self.mid.attn_1, self.mid.block_2 = ResnetBlock(block_in, block_in), AttnBlock(block_in), ResnetBlock(block_in, block_in)
        self.up = nn.ModuleList()
        for i_level in reversed(range(self.num_resolutions)):
            level = nn.Module()
            level.block, level.attn = nn.ModuleList(), nn.ModuleList()
            block_out = ch * ch_mult[i_level]
            for _ in range(num_res_blocks + 1):
                level.block.append(ResnetBlock(block_in, block_out))
                block_in = block_out
                if curr_res in attn_resolutions:
                    level.attn.append(AttnBlock(block_in))
            if i_level != 0:
                level.upsample = Upsample(block_in)
                curr_res *= 2
            self.up.insert(0, level)
        self.norm_out, self.conv_out = _norm(block_in), nn.Conv2d(block_in, out_ch, 3, padding=1)

    def forward(self, z):
        h = self.conv_in(z)
        h = self.mid.block_2(self.mid.attn_1(self.mid.block_1(h)))
        for i_level in reversed(range(self.num_resolutions)):
            level = self.up[i_level]
            for j in range(self.num_res_blocks + 1):
                h = level.block[j](h)
                if len(level.attn) > 0:
                    h = level.attn[j](h)
            if i_level != 0:
                h = level.upsample(h)
        return self.conv_out(F.silu(self.norm_out(h)))


class _Quantize(nn.Module):
    def __init__(self, n_embed, embed_dim):
        super().__init__()
        self.embed = nn.Embedding(n_embed, embed_dim)


class VQGanVAE(nn.Module):
    """Decoder half of GumbelVQ f8 with the attributes dalle-pytorch reads (``num_layers``, ``image_size``,
    ``num_tokens``, ``is_gumbel``)."""

    def __init__(self, vqgan_model_path: Optional[str] = None, vqgan_config_path: Optional[str] = None, *,
                 n_embed: int = 8192, embed_dim: int = 256, ddconfig: Optional[dict] = None):
        super().__init__()
        if vqgan_config_path is not None:
            import yaml

            conf = yaml.safe_load(open(vqgan_config_path))["model"]["params"]
            n_embed, embed_dim, ddconfig = conf["n_embed"], conf["embed_dim"], conf["ddconfig"]
        ddconfig = dict(ddconfig or dict(ch=128, out_ch=3, ch_mult=(1, 1, 2, 4), num_res_blocks=2,
                                         attn_resolutions=(32,), resolution=256, z_channels=256))
        self.quantize = _Quantize(n_embed, embed_dim)
        self.post_quant_conv = nn.Conv2d(embed_dim, ddconfig["z_channels"], 1)
        self.decoder = Decoder(**ddconfig)
        f = ddconfig["resolution"] // (ddconfig["resolution"] // 2 ** (len(ddconfig["ch_mult"]) - 1))
        self.num_layers = int(math.log2(f))
        self.image_size, self.num_tokens, self.is_gumbel = ddconfig["resolution"], n_embed, True
        if vqgan_model_path is not None:
            sd = torch.load(vqgan_model_path, map_location="cpu", weights_only=False)
            sd = sd.get("state_dict", sd)
            own = self.state_dict()
            missing = [k for k in own if k not in sd]
            if missing:
                raise RuntimeError(f"VQGAN checkpoint lacks keys: {missing[:5]}...")
            self.load_state_dict({k: sd[k] for k in own})

    @torch.no_grad()
    def decode(self, img_seq: torch.Tensor) -> torch.Tensor:
        b, n = img_seq.shape
        side = int(math.sqrt(n))
        z = self.quantize.embed(img_seq).view(b, side, side, -1).permute(0, 3, 1, 2).contiguous()
        img = self.decoder(self.post_quant_conv(z))
        return (img.clamp(-1.0, 1.0) + 1) * 0.5
