"""Python front-ends of csrc/elementwise.cu.  CUDA tensors launch the kernels; CPU tensors evaluate the same
contract in PyTorch (oracle for CPU tests of the executor and for the GPU numerics tests)."""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn.functional as F

from ..config import DalleConfig
from ..models.reference import shift_tokens
from . import lib


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _p(t: Optional[torch.Tensor]):
    return None if t is None else t.data_ptr()


# ------------------------------------------------------------------------------------------------
# LayerNorm + token shift
# ------------------------------------------------------------------------------------------------
def _permute_rows(n, cfg: DalleConfig, split: int):
    """[B*seq, D] -> [all samples' first `split` tokens | all remaining tokens]."""
    M, D = n.shape
    v = n.view(-1, cfg.seq_len, D)
    return torch.cat((v[:, :split].reshape(-1, D), v[:, split:].reshape(-1, D)), dim=0)


def _unpermute_rows(n, cfg: DalleConfig, split: int):
    M, D = n.shape
    B = M // cfg.seq_len
    a = n[: B * split].view(B, split, D)
    c = n[B * split:].view(B, cfg.seq_len - split, D)
    return torch.cat((a, c), dim=1).reshape(M, D)


def _ln_shift_ref(xin, w, b, cfg: DalleConfig, shift: bool):
    M, D = xin.shape
    mean = xin.mean(-1)
    var = xin.var(-1, unbiased=False)
    rstd = torch.rsqrt(var + cfg.ln_eps)
    n = (xin - mean[:, None]) * rstd[:, None] * w + b
    if shift:
        n = shift_tokens(n.view(-1, cfg.seq_len, D), cfg).reshape(M, D)
    return n, mean, rstd


def ln_shift_fwd(x, x2, w, b, out, mean, rstd, cfg: DalleConfig, shift: bool, row_split: int = 0):
    """out(bf16) = shift(LN(x or 0.5(x+x2))); fills mean/rstd.  row_split: see include/b200_api.h."""
    M, D = x.shape
    if not x.is_cuda:
        xin = x if x2 is None else 0.5 * (x + x2)
        n, mu, rs = _ln_shift_ref(xin.float(), w.detach().float(), b.detach().float(), cfg, shift)
        if row_split:
            n = _permute_rows(n, cfg, row_split)
        out.copy_(n.to(out.dtype)); mean.copy_(mu); rstd.copy_(rs)
        return out
    rc = lib.cuda_lib().b200_ln_shift_fwd(x.data_ptr(), _p(x2), w.data_ptr(), b.data_ptr(), out.data_ptr(),
                                          mean.data_ptr(), rstd.data_ptr(), M, D, cfg.seq_len, cfg.text_len,
                                          cfg.image_fmap_size, int(shift), row_split, cfg.ln_eps, _stream())
    lib.check(rc, "b200_ln_shift_fwd")
    lib.launch_count += 1
    return out


def ln_shift_bwd(dh, x, x2, mean, rstd, w, dx, dx2, dw, db, cfg: DalleConfig, shift: bool, mode: int,
                 row_split: int = 0, next_prep=None):
    """mode 0: dx += grad; mode 1: dx = dx2 = 0.5 grad.  dw, db accumulate.

    ``next_prep = (o, scale, d_o, dscale, dbias)`` (mode 0 only) additionally runs ``resid_bwd_prep`` of the block
    the backward walk visits next on the finished ``dx`` inside the same kernel."""
    M, D = x.shape
    if next_prep is not None:
        assert mode == 0 and x2 is None
        if x.is_cuda:
            o, scale, d_o, dscale, dbias = next_prep
            rc = lib.cuda_lib().b200_ln_shift_bwd_prep(
                dh.data_ptr(), x.data_ptr(), mean.data_ptr(), rstd.data_ptr(), w.data_ptr(), dx.data_ptr(),
                dw.data_ptr(), db.data_ptr(), o.data_ptr(), scale.data_ptr(), d_o.data_ptr(), dscale.data_ptr(),
                dbias.data_ptr(), M, D, cfg.seq_len, cfg.text_len, cfg.image_fmap_size, int(shift), row_split,
                _stream())
            lib.check(rc, "b200_ln_shift_bwd_prep")
            return
        ln_shift_bwd(dh, x, x2, mean, rstd, w, dx, dx2, dw, db, cfg, shift, mode, row_split)
        resid_bwd_prep(dx, *next_prep)
        return
    if not x.is_cuda:
        if row_split:
            dh = _unpermute_rows(dh, cfg, row_split)
        xin = (x if x2 is None else 0.5 * (x + x2)).detach().float().requires_grad_(True)
        wl = w.detach().float().requires_grad_(True)
        bl = torch.zeros(D, requires_grad=True)
        with torch.enable_grad():
            n, _, _ = _ln_shift_ref(xin, wl, bl, cfg, shift)
        gx, gw, gb = torch.autograd.grad(n, (xin, wl, bl), dh.float())
        if mode == 0:
            dx.add_(gx)
        else:
            dx.copy_(0.5 * gx); dx2.copy_(0.5 * gx)
        dw.add_(gw.view_as(dw)); db.add_(gb.view_as(db))
        return
    rc = lib.cuda_lib().b200_ln_shift_bwd(dh.data_ptr(), x.data_ptr(), _p(x2), mean.data_ptr(), rstd.data_ptr(),
                                          w.data_ptr(), dx.data_ptr(), _p(dx2), dw.data_ptr(), db.data_ptr(), M, D,
                                          cfg.seq_len, cfg.text_len, cfg.image_fmap_size, int(shift), row_split, mode,
                                          _stream())
    lib.check(rc, "b200_ln_shift_bwd")


def resid_bwd_prep(dy, o, scale, d_o, dscale, dbias):
    M, D = dy.shape
    if not dy.is_cuda:
        d = dy * scale.view(1, D)
        d_o.copy_(d.to(d_o.dtype))
        dscale.view(-1).add_((dy * o.float()).sum(0))
        dbias.view(-1).add_(d.sum(0))
        return
    rc = lib.cuda_lib().b200_resid_bwd_prep(dy.data_ptr(), o.data_ptr(), scale.data_ptr(), d_o.data_ptr(),
                                            dscale.data_ptr(), dbias.data_ptr(), M, D, _stream())
    lib.check(rc, "b200_resid_bwd_prep")


def colsum(x, out, alpha: float = 1.0):
    M, N = x.shape
    if not x.is_cuda:
        out.view(-1).add_(alpha * x.float().sum(0))
        return
    assert x.stride(1) == 1
    rc = lib.cuda_lib().b200_colsum_bf16(x.data_ptr(), x.stride(0), out.data_ptr(), M, N, alpha, _stream())
    lib.check(rc, "b200_colsum_bf16")


def embed_fwd(ids, W, x1, x2):
    M, D = x1.shape
    if not x1.is_cuda:
        v = W[ids.view(-1)]
        x1.copy_(v)
        if x2 is not None:
            x2.copy_(v)
        return
    rc = lib.cuda_lib().b200_embed_fwd(ids.data_ptr(), W.data_ptr(), x1.data_ptr(), _p(x2), M, D, _stream())
    lib.check(rc, "b200_embed_fwd")


def embed_bwd(ids, d1, d2, dW):
    M, D = d1.shape
    if not d1.is_cuda:
        g = d1 if d2 is None else d1 + d2
        dW.index_add_(0, ids.view(-1), g)
        return
    rc = lib.cuda_lib().b200_embed_bwd(ids.data_ptr(), d1.data_ptr(), _p(d2), dW.data_ptr(), M, D, _stream())
    lib.check(rc, "b200_embed_bwd")


def ce_fwd_bwd(logits, labels, n_classes: int, row_weight: float, grad_scale, loss_acc, write_grad=True):
    """In place on logits [rows, ld>=n_classes]: logits[:, :n_classes] <- row_weight*grad_scale*(softmax-onehot);
    loss_acc += row_weight * sum_rows NLL."""
    rows = logits.shape[0]
    if not logits.is_cuda:
        z = logits[:, :n_classes].float()
        lse = torch.logsumexp(z, -1)
        nll = lse - z.gather(1, labels.view(-1, 1)).squeeze(1)
        loss_acc.add_(row_weight * nll.sum())
        if write_grad:
            g = torch.softmax(z, -1)
            g[torch.arange(rows), labels.view(-1)] -= 1.0
            gs = 1.0 if grad_scale is None else grad_scale.float()
            logits[:, :n_classes] = (row_weight * gs * g).to(logits.dtype)
        return
    rc = lib.cuda_lib().b200_ce_fwd_bwd(logits.data_ptr(), logits.stride(0), labels.data_ptr(), rows, n_classes,
                                        row_weight, _p(grad_scale), loss_acc.data_ptr(), int(write_grad), _stream())
    lib.check(rc, "b200_ce_fwd_bwd")


def cast_bf16(src, dst):
    if not src.is_cuda:
        dst.copy_(src.to(dst.dtype))
        return
    rc = lib.cuda_lib().b200_cast_f32_bf16(src.data_ptr(), dst.data_ptr(), src.numel(), _stream())
    lib.check(rc, "b200_cast_f32_bf16")
