"""ctypes front-end of libdalle_b200_host.so (native/*.cpp)."""
from __future__ import annotations

import ctypes
from typing import Optional, Tuple

import torch

from ..ops import lib as _lib

_c = ctypes
_configured = False


class AllreduceDesc(_c.Structure):
    """include/b200_host_api.h B200hAllreduce."""
    _fields_ = [("n_members", _c.c_int), ("my_index", _c.c_int), ("modes", _c.POINTER(_c.c_int)),
                ("hosts", _c.POINTER(_c.c_char_p)), ("ports", _c.POINTER(_c.c_int)),
                ("span_begin", _c.POINTER(_c.c_int)), ("span_end", _c.POINTER(_c.c_int)), ("n_parts", _c.c_int),
                ("part_start", _c.POINTER(_c.c_longlong)), ("part_end", _c.POINTER(_c.c_longlong)),
                ("part_codec", _c.POINTER(_c.c_int)), ("flat", _c.c_void_p), ("weight", _c.c_float),
                ("group_id", _c.POINTER(_c.c_ubyte)), ("listener", _c.c_void_p), ("timeout_ms", _c.c_int),
                ("failed_out", _c.POINTER(_c.c_int))]


def allreduce_round(desc: "AllreduceDesc") -> int:
    """Blocking; ctypes releases the GIL for the duration of the call."""
    return int(_h().b200h_allreduce_round(_c.byref(desc)))


def available() -> bool:
    return _lib.host_lib_available()


def _h():
    global _configured
    h = _lib.host_lib()
    if not _configured:
        p, ll, i, f = _c.c_void_p, _c.c_longlong, _c.c_int, _c.c_float
        h.b200h_quantize_blockwise.argtypes = [p, p, p, p, ll, i]
        h.b200h_dequantize_blockwise.argtypes = [p, p, p, p, ll, i]
        h.b200h_uniform8_quantize.argtypes = [p, ll, p, p]
        h.b200h_uniform8_dequantize.argtypes = [p, p, p, ll]
        h.b200h_weighted_add.argtypes = [p, p, f, ll]
        h.b200h_delta.argtypes = [p, f, p, p, ll]
        h.b200h_listen.argtypes, h.b200h_listen.restype = [_c.c_char_p, i, _c.POINTER(i)], p
        h.b200h_accept.argtypes, h.b200h_accept.restype = [p, i], i
        h.b200h_close_listener.argtypes = [p]
        h.b200h_connect.argtypes, h.b200h_connect.restype = [_c.c_char_p, i, i], i
        h.b200h_send_frame.argtypes, h.b200h_send_frame.restype = [i, _c.c_ulonglong, p, ll], i
        h.b200h_recv_header.argtypes, h.b200h_recv_header.restype = [i, _c.POINTER(_c.c_ulonglong), i], ll
        h.b200h_recv_payload.argtypes, h.b200h_recv_payload.restype = [i, p, ll, i], i
        h.b200h_close.argtypes = [i]
        h.b200h_allreduce_round.argtypes, h.b200h_allreduce_round.restype = [_c.POINTER(AllreduceDesc)], i
        _configured = True
    return h


def quantize_blockwise(x: torch.Tensor, code: torch.Tensor, blocksize: int = 4096):
    xf = x.detach().reshape(-1).to(torch.float32).contiguous()
    n = xf.numel()
    q = torch.empty(n, dtype=torch.uint8)
    absmax = torch.empty((n + blocksize - 1) // blocksize, dtype=torch.float32)
    code = code.to(torch.float32).contiguous()
    _h().b200h_quantize_blockwise(code.data_ptr(), xf.data_ptr(), absmax.data_ptr(), q.data_ptr(), n, blocksize)
    return q.view(x.shape), absmax


def dequantize_blockwise(q: torch.Tensor, absmax: torch.Tensor, code: torch.Tensor, blocksize: int = 4096):
    qf = q.reshape(-1).contiguous()
    out = torch.empty(qf.numel(), dtype=torch.float32)
    code = code.to(torch.float32).contiguous()
    _h().b200h_dequantize_blockwise(code.data_ptr(), qf.data_ptr(), absmax.contiguous().data_ptr(), out.data_ptr(),
                                    qf.numel(), blocksize)
    return out.view(q.shape)


def uniform8_quantize(x: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    xf = x.detach().reshape(-1).to(torch.float32).contiguous()
    q, lookup = torch.empty(xf.numel(), dtype=torch.uint8), torch.empty(256, dtype=torch.float32)
    _h().b200h_uniform8_quantize(xf.data_ptr(), xf.numel(), q.data_ptr(), lookup.data_ptr())
    return q, lookup


def weighted_add_(acc: torch.Tensor, x: torch.Tensor, w: float):
    assert acc.is_contiguous() and x.is_contiguous() and acc.dtype == x.dtype == torch.float32
    _h().b200h_weighted_add(acc.data_ptr(), x.data_ptr(), float(w), acc.numel())
    return acc
