"""Python façade of the C++ bulk TCP transport (native/transport.cpp): framed tensor transfer with the GIL released.

Used for moving whole tensors / state between peers (``send_tensor`` / ``recv_tensor``); the msgpack streams of
``transport.p2p`` remain the control plane.  Falls back to nothing: callers check ``available()``.
"""
from __future__ import annotations

import ctypes
from typing import Optional, Tuple

import torch

from ..native import host


def available() -> bool:
    return host.available()


class BulkListener:
    def __init__(self, bind_host: str = "127.0.0.1", port: int = 0):
        h = host._h()
        out = ctypes.c_int(0)
        self._h = h
        self._l = h.b200h_listen(bind_host.encode(), port, ctypes.byref(out))
        if not self._l:
            raise OSError("bulk listener could not bind")
        self.port = out.value

    def accept(self, timeout: float = 10.0) -> Optional["BulkConnection"]:
        fd = self._h.b200h_accept(self._l, int(timeout * 1000))
        return BulkConnection(fd) if fd >= 0 else None

    def close(self):
        if self._l:
            self._h.b200h_close_listener(self._l)
            self._l = None


class BulkConnection:
    def __init__(self, fd: int):
        self.fd, self._h = fd, host._h()

    @classmethod
    def connect(cls, addr: Tuple[str, int], timeout: float = 10.0) -> "BulkConnection":
        fd = host._h().b200h_connect(addr[0].encode(), addr[1], int(timeout * 1000))
        if fd < 0:
            raise ConnectionError(f"cannot connect to {addr}")
        return cls(fd)

    def send_tensor(self, t: torch.Tensor, tag: int = 0):
        t = t.detach().contiguous()
        if self._h.b200h_send_frame(self.fd, tag, t.data_ptr(), t.numel() * t.element_size()) != 0:
            raise ConnectionError("send failed")

    def recv_into(self, t: torch.Tensor, timeout: float = 30.0) -> int:
        tag = ctypes.c_ulonglong(0)
        n = self._h.b200h_recv_header(self.fd, ctypes.byref(tag), int(timeout * 1000))
        if n < 0 or n != t.numel() * t.element_size():
            raise ConnectionError(f"bad frame ({n} bytes)")
        if self._h.b200h_recv_payload(self.fd, t.data_ptr(), n, int(timeout * 1000)) != 0:
            raise ConnectionError("recv failed")
        return tag.value

    def close(self):
        if self.fd >= 0:
            self._h.b200h_close(self.fd)
            self.fd = -1
