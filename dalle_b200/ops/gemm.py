"""Python front-end of the tcgen05 GEMM family (csrc/gemm_sm100.cu).

``gemm(...)`` launches on the current CUDA stream for CUDA tensors.  For CPU tensors the same contract is
evaluated with plain PyTorch (``gemm_ref``) - that path is the oracle used by the CPU tests of the layer
executor and is never taken for GPU tensors.
"""
from __future__ import annotations

import ctypes
import math
from typing import Optional

import torch

from . import lib

NT, NN, TN = 0, 1, 2
EPI_BF16, EPI_F32, EPI_ATOMIC_F32, EPI_ROTARY, EPI_RESID, EPI_GEGLU, EPI_GEGLU_BWD = range(7)


_ROT_T = {}


def rot_transposed(rot: torch.Tensor) -> torch.Tensor:
    """[n, 32, 2] -> pair-major [32, n, 2] copy (cached): consecutive rows become contiguous for the epilogue."""
    key = (rot.data_ptr(), tuple(rot.shape), str(rot.device))
    t = _ROT_T.get(key)
    if t is None:
        t = rot.permute(1, 0, 2).contiguous()
        _ROT_T[key] = t
    return t


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else t.data_ptr()


def _ld(t: Optional[torch.Tensor]) -> int:
    if t is None:
        return 0
    assert t.dim() == 2 and t.stride(1) == 1, "row-major 2-D tensors only"
    return t.stride(0)


def gelu_grad(x: torch.Tensor) -> torch.Tensor:
    cdf = 0.5 * (1.0 + torch.erf(x * 0.7071067811865476))
    pdf = 0.3989422804014327 * torch.exp(-0.5 * x * x)
    return cdf + x * pdf


def apply_rotary_pairs(x: torch.Tensor, rot: torch.Tensor, seq_len: int) -> torch.Tensor:
    """x: [M, C] with C % 64 == 0, rows are (batch, position) flattened; rot: [seq_len, 32, 2] (cos, sin)."""
    M, C = x.shape
    pos = torch.arange(M, device=x.device) % seq_len
    cs = rot[pos]  # [M, 32, 2]
    xv = x.view(M, C // 64, 32, 2)
    cos, sin = cs[:, None, :, 0], cs[:, None, :, 1]
    x0, x1 = xv[..., 0], xv[..., 1]
    out = torch.stack((x0 * cos - x1 * sin, x1 * cos + x0 * sin), dim=-1)
    return out.view(M, C)


def gemm_ref(mode, epi, A, B, out, *, out2=None, bias=None, scale=None, res=None, rot=None, seq_len=0, q_cols=0,
             q_scale=1.0, aux=None, n_half=0, alpha=1.0):
    """PyTorch statement of the kernel contract (fp32 math on the given operands)."""
    a, b = A.float(), B.float()
    if mode == NT:
        acc = a @ b.t()
    elif mode == NN:
        acc = a @ b
    else:
        acc = a.t() @ b
    if epi == EPI_BF16:
        if bias is not None:
            acc = acc + bias
        out.copy_((alpha * acc).to(out.dtype))
    elif epi == EPI_F32:
        out.copy_(alpha * acc)
    elif epi == EPI_ATOMIC_F32:
        out.add_(alpha * acc)
    elif epi == EPI_ROTARY:
        if bias is not None:
            acc = acc + bias
        r = apply_rotary_pairs(acc, rot, seq_len)
        r[:, :q_cols] *= q_scale
        out.copy_(r.to(out.dtype))
    elif epi == EPI_RESID:
        if bias is not None:
            acc = acc + bias
        out.copy_(res + alpha * scale * acc)
        if out2 is not None:
            out2.copy_(acc.to(out2.dtype))
    elif epi == EPI_GEGLU:
        if bias is not None:
            acc = acc + bias
        val, gate = acc[:, :n_half], acc[:, n_half:]
        out.copy_((val * torch.nn.functional.gelu(gate)).to(out.dtype))
        if out2 is not None:
            out2.copy_(acc.to(out2.dtype))
    elif epi == EPI_GEGLU_BWD:
        pre = aux.float()
        val, gate = pre[:, :n_half], pre[:, n_half:]
        out[:, :n_half].copy_((acc * torch.nn.functional.gelu(gate)).to(out.dtype))
        out[:, n_half:].copy_((acc * val * gelu_grad(gate)).to(out.dtype))
    else:
        raise ValueError(epi)
    return out


def gemm(mode: int, epi: int, A: torch.Tensor, B: torch.Tensor, out: torch.Tensor, *, out2=None, bias=None,
         scale=None, res=None, rot=None, seq_len: int = 0, q_cols: int = 0, q_scale: float = 1.0, aux=None,
         n_half: int = 0, alpha: float = 1.0, splits: int = 0, pos_ptr=None, a_colsum=None):
    """out = epilogue(A (x) B).  Shapes by mode: NT A[M,K] B[N,K]; NN A[M,K] B[K,N]; TN A[K,M] B[K,N].
    ``a_colsum`` (TN + EPI_ATOMIC_F32): fp32 [M] that additionally receives the column sums of A."""
    if not A.is_cuda:
        if a_colsum is not None:
            a_colsum.view(-1).add_(A.float().sum(0))
        return gemm_ref(mode, epi, A, B, out, out2=out2, bias=bias, scale=scale, res=res, rot=rot, seq_len=seq_len,
                        q_cols=q_cols, q_scale=q_scale, aux=aux, n_half=n_half, alpha=alpha)
    assert A.dtype == torch.bfloat16 and B.dtype == torch.bfloat16
    if mode == NT:
        M, K = A.shape
        N = B.shape[0]
        assert B.shape[1] == K
    elif mode == NN:
        M, K = A.shape
        N = B.shape[1]
        assert B.shape[0] == K
    else:
        K, M = A.shape
        N = B.shape[1]
        assert B.shape[0] == K
    if epi == EPI_GEGLU:
        assert mode == NT and N == 2 * n_half
        N = n_half
    d = lib.B200Gemm()
    d.mode, d.epi, d.M, d.N, d.K = mode, epi, M, N, K
    d.A, d.lda, d.B, d.ldb = A.data_ptr(), _ld(A), B.data_ptr(), _ld(B)
    d.out, d.ldo = out.data_ptr(), _ld(out)
    d.out2, d.ldo2 = _ptr(out2), _ld(out2)
    d.bias, d.scale = _ptr(bias), _ptr(scale)
    d.res, d.ldres = _ptr(res), _ld(res)
    if rot is not None:
        rot_t = rot_transposed(rot)
        d.rot, d.rot_len = rot_t.data_ptr(), rot_t.shape[1]
    d.seq_len, d.q_cols, d.q_scale = seq_len, q_cols, q_scale
    d.pos_ptr = _ptr(pos_ptr)
    d.aux, d.ldaux = _ptr(aux), _ld(aux)
    d.n_half, d.alpha, d.splits = n_half, alpha, splits
    if a_colsum is not None:
        assert mode == TN and epi == EPI_ATOMIC_F32 and a_colsum.dtype == torch.float32 and a_colsum.numel() == M
    d.a_colsum = _ptr(a_colsum)
    rc = lib.cuda_lib().b200_gemm_bf16(ctypes.byref(d), torch.cuda.current_stream().cuda_stream)
    lib.check(rc, "b200_gemm_bf16")
    return out
