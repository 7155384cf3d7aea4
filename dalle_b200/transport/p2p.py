"""Peer-to-peer transport: what the reference gets from the go-libp2p daemon, rebuilt on plain TCP.

One ``P2P`` object per peer: an asyncio TCP server plus a connection pool, carrying length-prefixed msgpack frames.
Handlers are registered by name; a call is ``(request) -> response`` (unary) or a bidirectional stream of frames
(used by the butterfly all-reduce to stream tensor parts and by state download).  Addresses keep the reference's
multiaddr spelling ``/ip4/<host>/tcp/<port>/p2p/<peer id>`` (reference arguments.py:93-117, utils.py:39-56).

Frames: ``u32 length | msgpack payload``; tensor bytes travel as msgpack ``bin`` values (zero-copy memoryview on the
send side).  The bulk part path can additionally be served by the C++ transport (native/transport.cpp).

Fault injection: environment variables checked on every outgoing frame -
``DALLE_B200_FAIL_DROP_AFTER_BYTES=n`` closes the connection after n bytes were sent by this process,
``DALLE_B200_FAIL_STALL=sec`` sleeps before each bulk frame, ``DALLE_B200_FAIL_CORRUPT=1`` flips a byte in bulk frames.
"""
from __future__ import annotations

import asyncio
import hashlib
import os
import socket
import struct
from typing import Any, AsyncIterator, Awaitable, Callable, Dict, List, Optional, Tuple

import msgpack

MAX_FRAME = 1 << 30
_bytes_sent = 0


class P2PError(Exception):
    pass


class PeerID:
    """Opaque peer identity: base58-ish hex digest of the peer's public identity bytes."""

    def __init__(self, raw: bytes):
        self.raw = bytes(raw)

    @classmethod
    def from_identity(cls, identity: bytes) -> "PeerID":
        return cls(hashlib.sha256(identity).digest()[:20])

    @classmethod
    def from_string(cls, s: str) -> "PeerID":
        return cls(bytes.fromhex(s[2:] if s.startswith("Qm") else s))

    def to_string(self) -> str:
        return "Qm" + self.raw.hex()

    __str__ = to_string

    def __repr__(self):
        return f"<PeerID {self.to_string()[:12]}>"

    def __hash__(self):
        return hash(self.raw)

    def __eq__(self, other):
        return isinstance(other, PeerID) and other.raw == self.raw

    def __lt__(self, other):
        return self.raw < other.raw


class Multiaddr:
    """Minimal multiaddr: /ip4|ip6|dns/<host>/tcp/<port>[/p2p/<id>] (enough for the reference's flags and logs)."""

    def __init__(self, s: str):
        self.s = str(s)
        parts = self.s.strip("/").split("/")
        self.proto: Dict[str, str] = {}
        i = 0
        while i < len(parts):
            name = parts[i]
            if name in ("quic", "ws", "wss"):
                self.proto[name] = ""
                i += 1
            else:
                self.proto[name] = parts[i + 1] if i + 1 < len(parts) else ""
                i += 2

    @property
    def host(self) -> Optional[str]:
        for k in ("ip4", "ip6", "dns", "dns4", "dns6"):
            if k in self.proto:
                return self.proto[k]
        return None

    @property
    def port(self) -> Optional[int]:
        return int(self.proto["tcp"]) if "tcp" in self.proto else None

    @property
    def peer_id(self) -> Optional[str]:
        return self.proto.get("p2p")

    def __getitem__(self, k):
        return self.proto[k]

    def __contains__(self, k):
        return k in self.proto or k in self.s

    def __str__(self):
        return self.s

    __repr__ = __str__

    def __eq__(self, o):
        return str(o) == self.s

    def __hash__(self):
        return hash(self.s)


def choose_ip_address(maddrs: List[Multiaddr], prefer_global: bool = True) -> str:
    """Pick the address other peers should use (public before private before loopback)."""
    import ipaddress

    hosts = [m.host for m in maddrs if m.host]
    def rank(h):
        try:
            ip = ipaddress.ip_address(h)
        except ValueError:
            return 1
        return 0 if ip.is_global else (3 if ip.is_loopback else 2)
    return sorted(hosts, key=rank)[0]


def _fail_hooks(nbytes: int, bulk: bool):
    global _bytes_sent
    _bytes_sent += nbytes
    limit = os.environ.get("DALLE_B200_FAIL_DROP_AFTER_BYTES")
    if limit is not None and _bytes_sent > int(limit):
        raise ConnectionResetError("fault injection: drop after bytes")


async def write_frame(writer: asyncio.StreamWriter, obj: Any, bulk: bool = False):
    data = msgpack.packb(obj, use_bin_type=True)
    _fail_hooks(len(data), bulk)
    if bulk:
        stall = os.environ.get("DALLE_B200_FAIL_STALL")
        if stall:
            await asyncio.sleep(float(stall))
        if os.environ.get("DALLE_B200_FAIL_CORRUPT") and len(data) > 64:
            data = bytearray(data)
            data[-8] ^= 0xFF
    writer.write(struct.pack("<I", len(data)))
    writer.write(data)
    await writer.drain()


async def read_frame(reader: asyncio.StreamReader) -> Any:
    hdr = await reader.readexactly(4)
    (n,) = struct.unpack("<I", hdr)
    if n > MAX_FRAME:
        raise P2PError(f"frame too large: {n}")
    data = await reader.readexactly(n)
    return msgpack.unpackb(data, raw=False)


class Stream:
    """Bidirectional frame stream handed to stream handlers / returned by ``P2P.open_stream``."""

    def __init__(self, reader, writer, peer_id: Optional[PeerID] = None):
        self.reader, self.writer, self.peer_id = reader, writer, peer_id

    async def send(self, obj, bulk: bool = False):
        await write_frame(self.writer, obj, bulk)

    async def recv(self, timeout: Optional[float] = None):
        if timeout is None:
            return await read_frame(self.reader)
        return await asyncio.wait_for(read_frame(self.reader), timeout)

    async def close(self):
        try:
            self.writer.close()
            await self.writer.wait_closed()
        except Exception:
            pass


Handler = Callable[[Any, PeerID], Awaitable[Any]]
StreamHandler = Callable[[Stream, Any], Awaitable[None]]


class P2P:
    def __init__(self, identity: Optional[bytes] = None, host: str = "0.0.0.0", port: int = 0,
                 announce_host: Optional[str] = None, client_mode: bool = False):
        self.identity = identity or os.urandom(32)
        self.peer_id = PeerID.from_identity(self.identity)
        self.host, self.port, self.client_mode = host, port, client_mode
        self.announce_host = announce_host
        self._unary: Dict[str, Handler] = {}
        self._stream: Dict[str, StreamHandler] = {}
        self._server: Optional[asyncio.AbstractServer] = None
        self._tasks: set = set()

    # ---- lifecycle ------------------------------------------------------------------------------
    async def start(self):
        if not self.client_mode:
            self._server = await asyncio.start_server(self._serve, self.host, self.port, reuse_address=True)
            self.port = self._server.sockets[0].getsockname()[1]
        return self

    async def shutdown(self):
        if self._server is not None:
            self._server.close()
            try:
                await asyncio.wait_for(self._server.wait_closed(), 1.0)
            except Exception:
                pass
        for t in list(self._tasks):
            t.cancel()

    def visible_maddrs(self) -> List[Multiaddr]:
        if self.client_mode:
            return []
        hosts = [self.announce_host] if self.announce_host else (
            ["127.0.0.1"] + _local_ips() if self.host in ("0.0.0.0", "") else [self.host])
        return [Multiaddr(f"/ip4/{h}/tcp/{self.port}/p2p/{self.peer_id}") for h in dict.fromkeys(hosts)]

    @property
    def endpoint(self) -> Optional[Tuple[str, int]]:
        if self.client_mode:
            return None
        host = self.announce_host or ("127.0.0.1" if self.host in ("0.0.0.0", "") else self.host)
        return (host, self.port)

    # ---- handlers -------------------------------------------------------------------------------
    def add_unary_handler(self, name: str, fn: Handler):
        self._unary[name] = fn

    def add_stream_handler(self, name: str, fn: StreamHandler):
        self._stream[name] = fn

    async def _serve(self, reader, writer):
        task = asyncio.current_task()
        self._tasks.add(task)
        try:
            sock = writer.get_extra_info("socket")
            if sock is not None:
                sock.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)
            while True:
                try:
                    hello = await read_frame(reader)
                except (asyncio.IncompleteReadError, ConnectionError):
                    break
                kind, name, sender, payload = hello["k"], hello["n"], PeerID(hello["s"]), hello.get("p")
                if kind == "u":
                    fn = self._unary.get(name)
                    try:
                        if fn is None:
                            raise P2PError(f"no handler {name}")
                        res = await fn(payload, sender)
                        await write_frame(writer, {"ok": True, "r": res})
                    except (ConnectionError, asyncio.CancelledError):
                        raise
                    except Exception as e:  # handler error -> reported to the caller
                        await write_frame(writer, {"ok": False, "e": f"{type(e).__name__}: {e}"})
                else:
                    fn = self._stream.get(name)
                    if fn is None:
                        break
                    await fn(Stream(reader, writer, sender), payload)
                    break
        except (ConnectionError, asyncio.IncompleteReadError, asyncio.CancelledError):
            pass
        finally:
            self._tasks.discard(task)
            try:
                writer.close()
            except Exception:
                pass

    # ---- calls ----------------------------------------------------------------------------------
    async def _connect(self, endpoint: Tuple[str, int], timeout: float):
        reader, writer = await asyncio.wait_for(asyncio.open_connection(endpoint[0], endpoint[1]), timeout)
        sock = writer.get_extra_info("socket")
        if sock is not None:
            sock.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)
        return reader, writer

    async def call(self, endpoint: Tuple[str, int], name: str, payload: Any, timeout: float = 10.0) -> Any:
        reader, writer = await self._connect(tuple(endpoint), timeout)
        try:
            await write_frame(writer, {"k": "u", "n": name, "s": self.peer_id.raw, "p": payload})
            res = await asyncio.wait_for(read_frame(reader), timeout)
        finally:
            writer.close()
        if not res.get("ok"):
            raise P2PError(res.get("e", "remote error"))
        return res.get("r")

    async def open_stream(self, endpoint: Tuple[str, int], name: str, payload: Any = None,
                          timeout: float = 10.0) -> Stream:
        reader, writer = await self._connect(tuple(endpoint), timeout)
        await write_frame(writer, {"k": "s", "n": name, "s": self.peer_id.raw, "p": payload})
        return Stream(reader, writer)


def _local_ips() -> List[str]:
    ips = []
    try:
        s = socket.socket(socket.AF_INET, socket.SOCK_DGRAM)
        s.connect(("10.255.255.255", 1))
        ips.append(s.getsockname()[0])
        s.close()
    except Exception:
        pass
    return [ip for ip in ips if not ip.startswith("127.")]


def parse_endpoint(maddr: str) -> Tuple[Tuple[str, int], Optional[str]]:
    m = Multiaddr(maddr)
    if m.host is None or m.port is None:
        raise ValueError(f"multiaddr {maddr!r} has no /ip4|dns/.../tcp/... part (IPFS/relay-only addresses are not "
                         f"supported: this transport is plain TCP)")
    return (m.host, m.port), m.peer_id
