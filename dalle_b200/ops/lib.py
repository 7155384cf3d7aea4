"""ctypes loader for the in-tree native libraries (C ABI declared in include/b200_api.h).

The CUDA library is mandatory on a GPU box: ``cuda_lib()`` raises if it is missing so that nothing silently
falls back to eager PyTorch (the CPU oracle ops live in :mod:`dalle_b200.ops.ref` and are only used for
CPU tensors).
"""
from __future__ import annotations

import ctypes
import os
import pathlib
from pathlib import Path

_HERE = Path(__file__).resolve().parent
_cuda = None
_host = None
launch_count = 0  # number of kernel-launching C-ABI calls made by this process (bench.py reports it)


class B200Gemm(ctypes.Structure):
    _fields_ = [
        ("mode", ctypes.c_int), ("epi", ctypes.c_int),
        ("M", ctypes.c_int), ("N", ctypes.c_int), ("K", ctypes.c_int),
        ("A", ctypes.c_void_p), ("lda", ctypes.c_longlong),
        ("B", ctypes.c_void_p), ("ldb", ctypes.c_longlong),
        ("out", ctypes.c_void_p), ("ldo", ctypes.c_longlong),
        ("out2", ctypes.c_void_p), ("ldo2", ctypes.c_longlong),
        ("bias", ctypes.c_void_p), ("scale", ctypes.c_void_p),
        ("res", ctypes.c_void_p), ("ldres", ctypes.c_longlong),
        ("rot", ctypes.c_void_p),
        ("seq_len", ctypes.c_int), ("rot_len", ctypes.c_int), ("q_cols", ctypes.c_int), ("q_scale", ctypes.c_float),
        ("pos_ptr", ctypes.c_void_p),
        ("aux", ctypes.c_void_p), ("ldaux", ctypes.c_longlong),
        ("n_half", ctypes.c_int), ("alpha", ctypes.c_float), ("splits", ctypes.c_int),
        ("a_colsum", ctypes.c_void_p),
    ]


class B200Group(ctypes.Structure):
    _fields_ = [
        ("rank", ctypes.c_int), ("world", ctypes.c_int),
        ("grad_ptrs", ctypes.c_void_p * 8), ("param_ptrs", ctypes.c_void_p * 8), ("shadow_ptrs", ctypes.c_void_p * 8),
        ("payload_ptrs", ctypes.c_void_p * 8), ("signal_ptrs", ctypes.c_void_p * 8), ("xchg_ptrs", ctypes.c_void_p * 8),
        ("xstride", ctypes.c_int), ("n_tensors", ctypes.c_int), ("n_blocks", ctypes.c_int),
        ("blk0", ctypes.c_int), ("blk1", ctypes.c_int), ("epoch", ctypes.c_uint), ("payload_mode", ctypes.c_int),
        ("zero_grad", ctypes.c_int), ("reduce_only", ctypes.c_int),
        ("q1", ctypes.c_void_p), ("q2", ctypes.c_void_p), ("f1", ctypes.c_void_p), ("f2", ctypes.c_void_p),
        ("absmax1", ctypes.c_void_p), ("absmax2", ctypes.c_void_p), ("code1", ctypes.c_void_p),
        ("code2", ctypes.c_void_p), ("blocks", ctypes.c_void_p), ("wd_per_tensor", ctypes.c_void_p),
        ("tensor_sums", ctypes.c_void_p), ("stats", ctypes.c_void_p), ("trust_out", ctypes.c_void_p),
        ("beta1", ctypes.c_float), ("beta2", ctypes.c_float), ("eps", ctypes.c_float),
        ("max_grad_norm", ctypes.c_float), ("clamp_value", ctypes.c_float), ("lr", ctypes.c_float),
        ("grad_scale", ctypes.c_float),
        ("chunk_codec", ctypes.c_void_p),
    ]


def _try_build():
    if os.environ.get("DALLE_B200_NO_AUTOBUILD"):
        return
    try:
        from . import build

        build.build_all()
    except Exception as e:  # pragma: no cover - build errors surface when the library is then missing
        import warnings

        warnings.warn(f"dalle_b200: native build failed: {e}")


def cuda_lib() -> ctypes.CDLL:
    global _cuda
    if _cuda is None:
        path = _HERE / "libdalle_b200.so"
        alt = os.environ.get("DALLE_B200_LIB")   # developer switch: A/B two builds inside one benchmark job
        if alt:
            path = pathlib.Path(alt)
        if not path.exists():
            _try_build()
        if not path.exists():
            raise RuntimeError(
                f"{path} is missing: build it with `python -m dalle_b200.ops.build` "
                "(the CUDA ops have no eager fallback on GPU tensors)")
        _cuda = ctypes.CDLL(str(path))
        for name in dir(_Sigs):
            if name.startswith("b200_"):
                fn = getattr(_cuda, name, None)
                if fn is not None:
                    fn.restype = ctypes.c_int
                    fn.argtypes = getattr(_Sigs, name)
    return _cuda


def host_lib() -> ctypes.CDLL:
    global _host
    if _host is None:
        path = _HERE / "libdalle_b200_host.so"
        if not path.exists():
            _try_build()
        if not path.exists():
            raise RuntimeError(f"{path} is missing: build it with `python -m dalle_b200.ops.build`")
        _host = ctypes.CDLL(str(path))
    return _host


def host_lib_available() -> bool:
    try:
        host_lib()
        return True
    except Exception:
        return False


_p = ctypes.c_void_p
_i = ctypes.c_int
_l = ctypes.c_longlong
_f = ctypes.c_float


class _Sigs:
    b200_gemm_bf16 = [ctypes.POINTER(B200Gemm), _p]
    b200_group_lamb_step = [ctypes.POINTER(B200Group), _p]
    b200_attn_fwd = [_p, _p, _p, _i, _i, _i, _i, _i, _i, _i, _p]
    b200_ln_shift_fwd = [_p, _p, _p, _p, _p, _p, _p, _i, _i, _i, _i, _i, _i, _i, _f, _p]
    b200_ln_shift_bwd = [_p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _i, _i, _i, _i, _i, _i, _i, _i, _p]
    b200_ln_shift_bwd_prep = [_p] * 13 + [_i] * 7 + [_p]
    b200_resid_bwd_prep = [_p, _p, _p, _p, _p, _p, _i, _i, _p]
    b200_colsum_bf16 = [_p, _l, _p, _i, _i, _f, _p]
    b200_embed_fwd = [_p, _p, _p, _p, _i, _i, _p]
    b200_embed_bwd = [_p, _p, _p, _p, _i, _i, _p]
    b200_ce_fwd_bwd = [_p, _l, _p, _i, _i, _f, _p, _p, _i, _p]
    b200_cast_f32_bf16 = [_p, _p, _l, _p]
    b200_lamb_step = [_p, _p, _p, _l, _p, _p, _p, _p, _p, _p, _p, _p, _p, _i, _p, _p, _i, _p, _p,
                      _f, _f, _f, _f, _f, _f, _f, _i, _p]
    b200_decode_embed = [_p, _i, _p, _p, _p, _p, _i, _i, _p]
    b200_decode_ln_shift = [_p, _p, _p, _p, _p, _p, _p, _i, _i, _i, _i, _i, _f, _i, _p]
    b200_decode_attn = [_p, _p, _p, _p, _p, _i, _i, _i, _i, _i, _i, _i, _p]
    b200_decode_sample = [_p, _i, _i, _f, _i, _f, _p, _p, _i, _l, _p, _i, _p, _i, _i, _p]
    b200_decode_advance = [_p, _p]
    b200_decode2_gemm = [_i, _p, _i, _i, _p, _p, _p, _p, _p, _p, _p, _i, _i, _f, _p, _p, _p, _p, _p, _i, _p, _i, _f, _i, _p]
    b200_decode2_attn = [_p, _p, _p, _p, _p, _i, _i, _i, _i, _i, _i, _i, _p]
    b200_decode2_embed = [_p, _i, _p, _p, _p, _p, _i, _p, _i, _i, _i, _p]
    b200_decode2_set_pdl = [_i]
    b200_attn_bwd = [_p, _p, _p, _p, _p, _p, _p, _p, _p, _f, _i, _i, _i, _i, _i, _i, _i, _p]


def check(rc: int, what: str):
    global launch_count
    launch_count += 1
    if rc != 0:
        raise RuntimeError(f"{what} failed with code {rc}")
