"""Averaging payload compression (hivemind.compression as used at reference task.py:125-126).

* ``Float16Compression``: clamp to +-65504, cast fp16.
* ``Uniform8BitQuantization``: per serialized part: mu = mean, sigma = ||x-mu|| / sqrt(n-1), scale = 6 sigma / 256,
  q = clamp(round_half_even((x-mu)/scale) + 128, 0, 255); codebook L[k] = mean of the ORIGINAL x in bucket k;
  wire = int64(256) | L (256 x fp32) | q (n x uint8); decode x^_i = L[q_i].
* ``SizeAdaptiveCompression(threshold, less, greater_equal)``: chosen by the whole tensor's numel.

The PyTorch code below is the oracle and the wire-format owner; CPU tensors go through the C++ host codec
(native/quant_host.cpp, ``dalle_b200.native.host``) when it is built.  On the GPU the same two codecs run inside the
fused averaging kernel (ops/csrc/group_avg.cu, payload mode 3) and never produce a separate serialized tensor.
"""
from __future__ import annotations

import math
from typing import Any, Dict, Optional

import numpy as np
import torch


class CompressionBase:
    name = "none"

    def compress(self, tensor: torch.Tensor) -> Dict[str, Any]:
        raise NotImplementedError

    def extract(self, msg: Dict[str, Any]) -> torch.Tensor:
        raise NotImplementedError

    def estimate_compression_ratio(self, numel: int) -> float:
        return 1.0


class NoCompression(CompressionBase):
    name = "none"

    def compress(self, tensor):
        t = tensor.detach().to(torch.float32).contiguous().cpu()
        return {"c": self.name, "n": t.numel(), "d": t.numpy().tobytes()}

    def extract(self, msg):
        return torch.from_numpy(np.frombuffer(msg["d"], dtype=np.float32).copy())


class Float16Compression(CompressionBase):
    name = "fp16"
    FP16_MAX = 65504.0

    def compress(self, tensor):
        t = tensor.detach().to(torch.float32).clamp(-self.FP16_MAX, self.FP16_MAX).to(torch.float16).contiguous().cpu()
        return {"c": self.name, "n": t.numel(), "d": t.numpy().tobytes()}

    def extract(self, msg):
        return torch.from_numpy(np.frombuffer(msg["d"], dtype=np.float16).astype(np.float32))

    def estimate_compression_ratio(self, numel):
        return 0.5


def uniform8_quantize_ref(x: torch.Tensor, n_bins: int = 256, range_in_sigmas: float = 6.0):
    """-> (q uint8 [n], lookup fp32 [n_bins])  -- the oracle statement of Appendix D.2."""
    x = x.detach().to(torch.float32).reshape(-1)
    n = x.numel()
    mu = x.mean()
    centered = x - mu
    std = torch.norm(centered) / math.sqrt(max(n - 1, 1))
    scale = range_in_sigmas * std / n_bins
    scale = scale if float(scale) > 0 else torch.tensor(1.0)
    q = torch.round(centered / scale).add_(n_bins // 2).clamp_(0, n_bins - 1).to(torch.int64)
    sums = torch.zeros(n_bins, dtype=torch.float32).scatter_add_(0, q, x)
    counts = torch.bincount(q, minlength=n_bins).clamp_(min=1).to(torch.float32)
    return q.to(torch.uint8), sums / counts


class Uniform8BitQuantization(CompressionBase):
    name = "u8"
    N_BINS = 256

    def compress(self, tensor):
        t = tensor.detach().to(torch.float32).reshape(-1)
        q, lookup = _uniform8(t)
        return {"c": self.name, "n": t.numel(), "l": lookup.numpy().tobytes(), "d": q.numpy().tobytes()}

    def extract(self, msg):
        lookup = torch.from_numpy(np.frombuffer(msg["l"], dtype=np.float32).copy())
        q = torch.from_numpy(np.frombuffer(msg["d"], dtype=np.uint8).copy())
        return lookup[q.long()]

    def estimate_compression_ratio(self, numel):
        return 0.25 + 1024.0 / max(4 * numel, 1)


def _uniform8(t: torch.Tensor):
    if not t.is_cuda:
        try:
            from ..native import host

            if host.available():
                return host.uniform8_quantize(t)
        except Exception:
            pass
    return uniform8_quantize_ref(t.cpu())


class SizeAdaptiveCompression(CompressionBase):
    def __init__(self, threshold: int, less: CompressionBase, greater_equal: CompressionBase):
        self.threshold, self.less, self.greater_equal = threshold, less, greater_equal

    def choose(self, tensor_numel: int) -> CompressionBase:
        return self.greater_equal if tensor_numel >= self.threshold else self.less

    def compress(self, tensor, tensor_numel: Optional[int] = None):
        return self.choose(tensor.numel() if tensor_numel is None else tensor_numel).compress(tensor)

    def extract(self, msg):
        return extract(msg)

    def estimate_compression_ratio(self, numel):
        return self.choose(numel).estimate_compression_ratio(numel)


_BY_NAME = {c.name: c() for c in (NoCompression, Float16Compression, Uniform8BitQuantization)}


def extract(msg: Dict[str, Any]) -> torch.Tensor:
    """Decode any serialized part (the codec name travels with the payload)."""
    return _BY_NAME[msg["c"]].extract(msg)
