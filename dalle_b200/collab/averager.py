"""Decentralized averaging between peers: DHT-keyed matchmaking + butterfly all-reduce + state download
(hivemind DecentralizedAverager / Matchmaking / AllReduceRunner as driven by hivemind.Optimizer).

Matchmaking: every peer that wants to average declares ``(endpoint, deadline)`` under ``{prefix}.matchmaking``; the
declared peer with the earliest deadline leads, the others ask it to ``join``; the leader closes the group when
``target_group_size`` peers are present or its deadline passes (``min_group_size`` = 2) and sends everybody the
ordered member list.

All-reduce ("butterfly"): the flat fp32 buffer is cut into parts of <= ``part_size_bytes``; contiguous runs of
parts are owned by the non-client members.  Every member streams the parts of owner k to k (compressed: fp16 for
tensors below the size threshold, uniform 8-bit above; reference task.py:125-126), k accumulates
``sum_i w_i x_i``, divides by ``sum_i w_i`` and streams back the DELTA ``avg - x_i`` (compressed); the sender applies
``x += delta``.  A sender that fails or stalls is dropped from the denominator of the remaining parts; an owner that
fails leaves its span un-averaged on the survivors (they keep local values) - the round still completes.
"""
from __future__ import annotations

import asyncio
import contextlib
import enum
import hashlib
import logging
import os
import threading
import time
from typing import Any, Dict, List, Optional, Sequence, Tuple

import msgpack
import numpy as np
import torch

from ..dht import DHT, get_dht_time
from ..transport.p2p import P2PError, PeerID, Stream
from . import compression as C

logger = logging.getLogger(__name__)


def _bulk_available() -> bool:
    try:
        from ..transport import bulk

        return bulk.available()
    except Exception:
        return False


class AveragingMode(enum.IntEnum):
    NODE = 0      # owns a span and contributes data
    CLIENT = 1    # contributes data, cannot accept connections (owns nothing)
    AUX = 2       # owns a span, contributes nothing (weight 0)


class AllreduceException(Exception):
    pass


class MatchmakingException(Exception):
    pass


class StepControl:
    """Handle of a pre-scheduled averaging round (hivemind.averaging.control.StepControl): matchmaking starts at once
    and runs until ``scheduled_time``; the all-reduce itself begins only after ``allow_allreduce()`` - by then the
    caller has put its final data into the averager and set ``weight``."""

    def __init__(self, weight: float, scheduled_time: Optional[float]):
        self.weight, self.scheduled_time = float(weight), scheduled_time
        self._allow, self._cancelled = threading.Event(), threading.Event()
        self.began_allreduce = False
        self.future = None

    def allow_allreduce(self):
        self._allow.set()

    def cancel(self):
        self._cancelled.set()
        self._allow.set()

    @property
    def cancelled(self) -> bool:
        return self._cancelled.is_set()

    def done(self) -> bool:
        return self.future is not None and self.future.done()

    def result(self, timeout: Optional[float] = None):
        return self.future.result(timeout)


class GroupInfo:
    def __init__(self, group_id: bytes, peer_ids: List[bytes], endpoints: List[Optional[Tuple[str, int]]],
                 modes: List[int], gathered: List[Any], bulk_ports: Optional[List[Optional[int]]] = None):
        self.group_id, self.peer_ids, self.endpoints, self.modes, self.gathered = group_id, peer_ids, endpoints, modes, gathered
        self.bulk_ports = list(bulk_ports) if bulk_ports is not None else [None] * len(peer_ids)

    def to_msg(self):
        return {"gid": self.group_id, "pids": self.peer_ids, "eps": [list(e) if e else None for e in self.endpoints],
                "modes": self.modes, "gathered": self.gathered, "bulk": self.bulk_ports}

    @classmethod
    def from_msg(cls, m):
        return cls(m["gid"], list(m["pids"]), [tuple(e) if e else None for e in m["eps"]], list(m["modes"]),
                   list(m["gathered"]), m.get("bulk"))

    def native_data_plane(self) -> bool:
        """Every member that owns a span (everyone but clients) listens on the C++ bulk transport and every member
        speaks it (clients declare 0): the round can run in b200h_allreduce_round."""
        return all(b is not None and (b > 0 or m == AveragingMode.CLIENT) for b, m in zip(self.bulk_ports, self.modes))

    @property
    def size(self):
        return len(self.peer_ids)


# ------------------------------------------------------------------------------------------------
# partitioning of the flat buffer
# ------------------------------------------------------------------------------------------------
class Partition:
    """Parts [(start, end, tensor_numel_for_codec_choice)] over the flat buffer + owner spans."""

    def __init__(self, tensor_sizes: Sequence[int], compression: C.CompressionBase, part_elems: int):
        self.total = int(sum(tensor_sizes))
        parts: List[Tuple[int, int, int]] = []
        # consecutive tensors that pick the same codec form a region; regions are cut into parts
        off, region_start, region_codec, region_numel = 0, 0, None, 0
        def codec_of(n):
            return compression.choose(n).name if isinstance(compression, C.SizeAdaptiveCompression) else compression.name
        def flush(end):
            nonlocal parts
            s = region_start
            while s < end:
                e = min(end, s + part_elems)
                parts.append((s, e, region_numel))
                s = e
        for n in tensor_sizes:
            c = codec_of(n)
            if region_codec is None:
                region_codec, region_numel = c, n
            elif c != region_codec:
                flush(off)
                region_start, region_codec, region_numel = off, c, n
            else:
                region_numel = max(region_numel, n)
            off += n
        if off > region_start:
            flush(off)
        self.parts = parts

    def assign(self, weights: Sequence[float]) -> List[Tuple[int, int]]:
        """Contiguous runs of part indices per member, proportional to ``weights`` (0 = owns nothing)."""
        n = len(self.parts)
        sizes = [e - s for s, e, _ in self.parts]
        total = float(sum(sizes))
        owners = [i for i, w in enumerate(weights) if w > 0]
        tw = float(sum(weights[i] for i in owners))
        spans: List[Tuple[int, int]] = [(0, 0)] * len(weights)
        p, cum, target = 0, 0.0, 0.0
        for idx, i in enumerate(owners):
            target += weights[i] / tw * total
            q = p
            if idx == len(owners) - 1:
                q = n
            else:
                while q < n and cum + sizes[q] <= target + 1e-6:
                    cum += sizes[q]
                    q += 1
            spans[i] = (p, q)
            p = q
        return spans


class _PartReducer:
    """Owner-side accumulation of one span: sum_i w_i x_i / sum_i w_i per part, tolerant to failing senders."""

    def __init__(self, part_shapes: List[int], senders: List[int]):
        self.acc = [torch.zeros(n, dtype=torch.float32) for n in part_shapes]
        self.denom = [0.0] * len(part_shapes)
        self.got = [set() for _ in part_shapes]
        self.active = set(senders)
        self.futures = [asyncio.get_event_loop().create_future() for _ in part_shapes]

    def _check(self, i):
        if not self.futures[i].done() and self.active <= self.got[i]:
            d = self.denom[i]
            self.futures[i].set_result(self.acc[i] / d if d > 0 else self.acc[i])

    def add(self, i: int, sender: int, x: Optional[torch.Tensor], w: float):
        if sender in self.got[i]:
            return
        if x is not None and w > 0:
            self.acc[i].add_(x, alpha=w)
            self.denom[i] += w
        self.got[i].add(sender)
        self._check(i)

    def fail(self, sender: int):
        self.active.discard(sender)
        for i in range(len(self.futures)):
            self._check(i)

    async def result(self, i, timeout):
        return await asyncio.wait_for(asyncio.shield(self.futures[i]), timeout)


class _Round:
    def __init__(self, group: GroupInfo, my_index: int, weight: float):
        self.group, self.my_index, self.weight = group, my_index, weight
        self.reducer: Optional[_PartReducer] = None
        self.ready = asyncio.get_event_loop().create_future()


class DecentralizedAverager:
    def __init__(self, tensor_shapes_or_tensors, dht: DHT, *, prefix: str, start: bool = True,
                 target_group_size: Optional[int] = None, min_group_size: int = 2, min_matchmaking_time: float = 5.0,
                 request_timeout: float = 3.0, allreduce_timeout: Optional[float] = None,
                 sender_timeout: Optional[float] = None, compression: C.CompressionBase = C.NoCompression(),
                 state_compression: C.CompressionBase = C.NoCompression(), part_size_bytes: int = 2 ** 19,
                 client_mode: Optional[bool] = None, auxiliary: bool = False, bandwidth: Optional[float] = None,
                 declare_state_period: float = 30.0, allow_state_sharing: bool = True, pin_memory: bool = False):
        self.dht, self.prefix = dht, prefix
        self.client_mode = dht.client_mode if client_mode is None else client_mode
        assert not (self.client_mode and auxiliary), "auxiliary peers must accept incoming connections"
        self.mode = AveragingMode.AUX if auxiliary else (AveragingMode.CLIENT if self.client_mode else AveragingMode.NODE)
        self.target_group_size, self.min_group_size = target_group_size, min_group_size
        self.min_matchmaking_time, self.request_timeout = min_matchmaking_time, request_timeout
        self.allreduce_timeout = allreduce_timeout
        self.sender_timeout = sender_timeout if sender_timeout is not None else (allreduce_timeout or 30.0)
        self.compression, self.state_compression = compression, state_compression
        self.part_elems = max(1, part_size_bytes // 4)
        self.bandwidth = bandwidth if bandwidth is not None else (0.0 if self.client_mode else 1.0)
        self.declare_state_period, self.allow_state_sharing = declare_state_period, allow_state_sharing
        self.state_sharing_priority = 0.0

        shapes = [tuple(t.shape) if isinstance(t, torch.Tensor) else tuple(t) for t in tensor_shapes_or_tensors]
        self._sizes = [int(np.prod(s)) if len(s) else 1 for s in shapes]
        total = sum(self._sizes)
        self._flat = torch.zeros(total, dtype=torch.float32)
        if pin_memory and torch.cuda.is_available():
            self._flat = self._flat.pin_memory()
        self._tensors, off = [], 0
        for s, n in zip(shapes, self._sizes):
            self._tensors.append(self._flat[off:off + n].view(s))
            off += n
        for src, dst in zip(tensor_shapes_or_tensors, self._tensors):
            if isinstance(src, torch.Tensor):
                dst.copy_(src.detach().cpu())
        self._partition = Partition(self._sizes, compression, self.part_elems)
        # what two peers must agree on to average the same thing: tensor sizes in order and the part boundaries
        # (which follow from the sizes, the codec choice per tensor and the part size).  Declared with the
        # matchmaking record and checked in the join handshake; a peer with another schema is never grouped.
        self.schema = hashlib.sha1(repr((self._sizes, [(a, b) for a, b, _ in self._partition.parts])).encode()).hexdigest()[:16]
        self._schema_warned: set = set()
        self.lock_averaged_tensors = threading.Lock()
        self._rounds: Dict[bytes, _Round] = {}
        self._round_waiters: Dict[bytes, asyncio.Future] = {}
        self._looking = False
        self._leading = False
        self._followers: Dict[bytes, dict] = {}
        self._assembled: Optional[asyncio.Future] = None
        self._mm_lock: Optional[asyncio.Lock] = None
        self._declare_task = None
        self.last_group: Optional[GroupInfo] = None
        # C++ data plane of the all-reduce (native/allreduce_pipe.cpp): one persistent listener per averager; its
        # port travels with the matchmaking record.  DALLE_B200_NATIVE_ALLREDUCE=0 keeps the asyncio part loop.
        self._bulk_listener = None
        self._bulk_port: Optional[int] = None
        if _bulk_available() and os.environ.get("DALLE_B200_NATIVE_ALLREDUCE", "1") != "0":
            if self.client_mode:
                self._bulk_port = 0
            else:
                from ..transport.bulk import BulkListener

                try:
                    self._bulk_listener = BulkListener(dht.p2p.endpoint[0] if dht.p2p.endpoint else "127.0.0.1", 0)
                    self._bulk_port = self._bulk_listener.port
                except OSError:
                    self._bulk_listener = None
        self.native_rounds = 0
        self._round_key: Optional[int] = None
        p2p = dht.p2p
        p2p.add_stream_handler(f"{prefix}.join", self._rpc_join)
        p2p.add_stream_handler(f"{prefix}.allreduce", self._rpc_allreduce)
        p2p.add_stream_handler(f"{prefix}.download", self._rpc_download)
        if start and allow_state_sharing and not self.client_mode:
            self._declare_task = asyncio.run_coroutine_threadsafe(self._declare_loop(), dht.loop)

    # ---- tensors --------------------------------------------------------------------------------
    @contextlib.contextmanager
    def get_tensors(self):
        with self.lock_averaged_tensors:
            yield self._tensors

    @property
    def total_size(self) -> int:
        return self._flat.numel()

    def shutdown(self):
        if self._declare_task is not None:
            self._declare_task.cancel()
        if self._bulk_listener is not None:
            self._bulk_listener.close()
            self._bulk_listener = None

    # ---- step -----------------------------------------------------------------------------------
    def step(self, gather: Any = None, weight: Optional[float] = None, timeout: Optional[float] = None,
             allow_retries: bool = True, wait: bool = True, target_group_size: Optional[int] = None,
             round_key: Optional[int] = None):
        """Find a group and run one all-reduce.  Returns {peer_id_bytes: gathered} on success, None on failure
        (the local tensors are then unchanged or partially averaged, never corrupted).  ``round_key`` (for instance
        the epoch): peers that declare different keys are never grouped; None matches anything (auxiliary peers)."""
        if weight is None:
            weight = float(self.mode != AveragingMode.AUX)
        fut = asyncio.run_coroutine_threadsafe(
            self._step(gather, float(weight), timeout, allow_retries, target_group_size, None, round_key), self.dht.loop)
        return fut.result() if wait else fut

    def schedule_step(self, scheduled_time: Optional[float] = None, gather: Any = None, weight: Optional[float] = None,
                      timeout: Optional[float] = None, allow_retries: bool = False,
                      target_group_size: Optional[int] = None, round_key: Optional[int] = None) -> StepControl:
        """Start looking for a group NOW, aiming to have it assembled at ``scheduled_time`` (DHT time), and hold the
        all-reduce until ``control.allow_allreduce()`` (reference flag --matchmaking_time, arguments.py:66-68: peers
        begin matchmaking that many seconds before the tracker's ETA, so the wait overlaps the last micro-batches)."""
        if weight is None:
            weight = float(self.mode != AveragingMode.AUX)
        control = StepControl(weight, scheduled_time)
        control.future = asyncio.run_coroutine_threadsafe(
            self._step(gather, float(weight), timeout, allow_retries, target_group_size, control, round_key),
            self.dht.loop)
        return control

    async def _step(self, gather, weight, timeout, allow_retries, target_group_size,
                    control: Optional[StepControl] = None, round_key: Optional[int] = None):
        if self._mm_lock is None:
            self._mm_lock = asyncio.Lock()
        deadline = time.monotonic() + timeout if timeout is not None else None
        async with self._mm_lock:
            attempt = 0
            while True:
                attempt += 1
                mm_time = self.min_matchmaking_time
                if control is not None and control.scheduled_time is not None and attempt == 1:
                    mm_time = max(0.5, control.scheduled_time - get_dht_time())
                if deadline is not None:
                    mm_time = min(mm_time, max(0.1, deadline - time.monotonic()))
                try:
                    if control is not None and control.cancelled:
                        return None
                    self._round_key = round_key
                    t_mm = time.monotonic()
                    group = await self._look_for_group(gather, weight, mm_time, target_group_size or self.target_group_size)
                    logger.debug(f"{self.prefix}: matchmaking (key {round_key}, window {mm_time:.2f}s, scheduled "
                                 f"{control is not None}) -> {'group of %d' % group.size if group else 'nobody'} after "
                                 f"{time.monotonic() - t_mm:.2f}s")
                    if group is None:
                        raise MatchmakingException("no group found")
                    if control is not None:
                        # the group exists; the data may not: wait for the owner of this round to release it
                        while not control._allow.is_set():
                            if deadline is not None and time.monotonic() >= deadline:
                                raise asyncio.TimeoutError()
                            await asyncio.sleep(0.002)
                        if control.cancelled:
                            return None
                        weight = control.weight
                        control.began_allreduce = True
                    remaining = None if deadline is None else max(0.1, deadline - time.monotonic())
                    ar_timeout = self.allreduce_timeout if remaining is None else (
                        min(self.allreduce_timeout, remaining) if self.allreduce_timeout else remaining)
                    await asyncio.wait_for(self._run_allreduce(group, weight), ar_timeout)
                    self.last_group = group
                    return {pid: g for pid, g in zip(group.peer_ids, group.gathered)}
                except (MatchmakingException, AllreduceException, asyncio.TimeoutError, P2PError, ConnectionError,
                        OSError) as e:
                    logger.debug(f"{self.prefix}: averaging attempt {attempt} failed: {e!r}")
                    if not allow_retries or (deadline is not None and time.monotonic() >= deadline - 0.05):
                        return None
                    await asyncio.sleep(0.05)

    # ---- matchmaking ----------------------------------------------------------------------------
    def _my_info(self, gather, deadline):
        ep = self.dht.p2p.endpoint
        return {"pid": self.dht.peer_id.raw, "ep": list(ep) if ep else None, "mode": int(self.mode),
                "deadline": deadline, "gather": gather, "bw": self.bandwidth, "schema": self.schema,
                "bulk": self._bulk_port, "rk": self._round_key}

    def _same_round(self, theirs) -> bool:
        mine = self._round_key
        return mine is None or theirs is None or int(mine) == int(theirs)

    def _warn_schema(self, who: bytes, theirs):
        if who not in self._schema_warned:
            self._schema_warned.add(who)
            logger.warning(f"{self.prefix}: peer {who.hex()[:12] if not who.startswith(b'leader@') else who.decode()} "
                           f"declares tensor schema {theirs}, this peer has {self.schema} (different model, codec or "
                           f"part size) - not grouping with it")

    async def _declared(self):
        entry = await self.dht.node.get(f"{self.prefix}.matchmaking", latest=True)
        now = get_dht_time()
        out = []
        if entry is not None and isinstance(entry.value, dict):
            for pid, v in entry.value.items():
                d = v.value
                if d and d.get("ep") and v.expiration_time > now:
                    if d.get("schema", self.schema) != self.schema:
                        self._warn_schema(bytes(pid), d.get("schema"))
                        continue
                    if not self._same_round(d.get("rk")):
                        continue       # that peer is looking for a group for another epoch
                    out.append((d["deadline"], bytes(pid), tuple(d["ep"])))
        return sorted(out)

    async def _look_for_group(self, gather, weight, mm_time, target_group_size) -> Optional[GroupInfo]:
        me = self.dht.peer_id.raw
        deadline = get_dht_time() + mm_time
        info = self._my_info(gather, deadline)
        self._looking, self._leading = True, False
        self._followers = {}
        self._assembled = asyncio.get_event_loop().create_future()
        try:
            if self.mode != AveragingMode.CLIENT:
                await self.dht.node.store(f"{self.prefix}.matchmaking",
                                          {"ep": info["ep"], "deadline": deadline, "schema": self.schema,
                                           "rk": self._round_key},
                                          deadline + 1.0, subkey=me)
            tried_at: Dict[bytes, float] = {}
            while True:
                now = get_dht_time()
                if self._assembled.done():  # a leader (possibly me, via a follower filling the group) closed it
                    return self._assembled.result()
                cands = [c for c in await self._declared() if c[0] > now - 0.5]
                better = [c for c in cands if c[1] != me and (self.mode == AveragingMode.CLIENT or (c[0], c[1]) < (deadline, me))]
                better = [c for c in better if time.monotonic() - tried_at.get(c[1], -1e9) > 0.5]
                if better and not self._followers:
                    self._leading = False
                    ld, lpid, lep = better[0]
                    tried_at[lpid] = time.monotonic()
                    group = await self._join_leader(lep, info, max(ld, deadline) - now + self.request_timeout)
                    if group is not None:
                        return group
                    continue
                # lead
                self._leading = self.mode != AveragingMode.CLIENT
                if self._leading and target_group_size and len(self._followers) + 1 >= target_group_size:
                    return self._assemble(info)
                if now >= deadline:
                    if self._leading and len(self._followers) + 1 >= self.min_group_size:
                        return self._assemble(info)
                    self._disband()
                    return None
                await asyncio.sleep(min(0.1, max(0.01, deadline - now)))
        finally:
            self._looking = False
            self._leading = False

    def _assemble(self, my_info) -> GroupInfo:
        members = [my_info] + [f["info"] for f in self._followers.values()]
        members.sort(key=lambda m: bytes(m["pid"]))
        group = GroupInfo(os.urandom(16), [bytes(m["pid"]) for m in members],
                          [tuple(m["ep"]) if m["ep"] else None for m in members], [m["mode"] for m in members],
                          [m["gather"] for m in members], [m.get("bulk") for m in members])
        if not self._assembled.done():
            self._assembled.set_result(group)
        self._looking = False
        return group

    def _disband(self):
        if self._assembled is not None and not self._assembled.done():
            self._assembled.set_result(None)

    async def _rpc_join(self, stream: Stream, payload):
        try:
            if not (self._looking and self._leading) or self._assembled is None or self._assembled.done():
                await stream.send({"status": "rejected"})
                return
            pid = bytes(payload["pid"])
            if payload.get("schema") != self.schema:
                self._warn_schema(pid, payload.get("schema"))
                await stream.send({"status": "schema_mismatch", "schema": self.schema})
                return
            if not self._same_round(payload.get("rk")):
                await stream.send({"status": "rejected"})
                return
            self._followers[pid] = {"info": payload}
            await stream.send({"status": "accepted"})
            assembled = self._assembled
            try:
                group = await asyncio.wait_for(asyncio.shield(assembled), payload.get("wait", 30.0))
            except asyncio.TimeoutError:
                group = None
            if group is None or pid not in group.peer_ids:
                await stream.send({"status": "disband"})
            else:
                await stream.send({"status": "begin", "group": group.to_msg()})
        except (ConnectionError, asyncio.IncompleteReadError):
            self._followers.pop(bytes(payload.get("pid", b"")), None)
        finally:
            await stream.close()

    async def _join_leader(self, leader_ep, info, wait) -> Optional[GroupInfo]:
        try:
            stream = await self.dht.p2p.open_stream(leader_ep, f"{self.prefix}.join", dict(info, wait=wait),
                                                    timeout=self.request_timeout)
        except Exception:
            return None
        try:
            msg = await stream.recv(self.request_timeout)
            if msg.get("status") == "schema_mismatch":
                self._warn_schema(b"leader@" + repr(tuple(leader_ep)).encode(), msg.get("schema"))
            if msg.get("status") != "accepted":
                return None
            msg = await stream.recv(wait + self.request_timeout)
            if msg.get("status") == "begin":
                return GroupInfo.from_msg(msg["group"])
            return None
        except Exception:
            return None
        finally:
            await stream.close()

    # ---- all-reduce -----------------------------------------------------------------------------
    def _spans(self, group: GroupInfo):
        owner_w = [0.0 if (m == AveragingMode.CLIENT or ep is None) else 1.0 for m, ep in zip(group.modes, group.endpoints)]
        return self._partition.assign(owner_w)

    async def _run_allreduce(self, group: GroupInfo, weight: float):
        if group.native_data_plane() and (self._bulk_listener is not None or self.client_mode):
            return await self._run_allreduce_native(group, weight)
        me = group.peer_ids.index(self.dht.peer_id.raw)
        spans = self._spans(group)
        parts = self._partition.parts
        senders = [i for i, m in enumerate(group.modes) if m != AveragingMode.AUX]
        rnd = _Round(group, me, weight)
        p0, p1 = spans[me]
        if p1 > p0:
            rnd.reducer = _PartReducer([parts[i][1] - parts[i][0] for i in range(p0, p1)], senders)
        self._rounds[group.group_id] = rnd
        waiter = self._round_waiters.pop(group.group_id, None)
        if waiter is not None and not waiter.done():
            waiter.set_result(rnd)
        try:
            tasks = []
            for k, (q0, q1) in enumerate(spans):
                if q1 <= q0:
                    continue
                if k == me:
                    tasks.append(self._reduce_own_span(rnd, p0, p1))
                elif self.mode != AveragingMode.AUX:
                    tasks.append(self._send_span(rnd, k, q0, q1))
            results = await asyncio.gather(*tasks, return_exceptions=True)
            failures = [r for r in results if isinstance(r, Exception)]
            if failures and len(failures) == len(results):
                raise AllreduceException(f"all spans failed: {failures[0]!r}")
        finally:
            # keep serving late senders for a moment, then forget the round
            asyncio.get_event_loop().call_later(5.0, self._rounds.pop, group.group_id, None)

    _CODEC_IDS = {"none": 0, "fp16": 1, "uniform8": 2, "u8": 2}

    async def _run_allreduce_native(self, group: GroupInfo, weight: float):
        """The whole round in C++ threads (native/allreduce_pipe.cpp): this coroutine only builds the descriptor and
        waits for the call in a worker thread, with the GIL released."""
        import ctypes

        from ..native import host

        me = group.peer_ids.index(self.dht.peer_id.raw)
        n, parts, spans = group.size, self._partition.parts, self._spans(group)
        comp = self.compression

        def codec_id(tensor_numel):
            c = comp.choose(tensor_numel) if isinstance(comp, C.SizeAdaptiveCompression) else comp
            return self._CODEC_IDS[c.name]

        I, LL = ctypes.c_int, ctypes.c_longlong
        modes = (I * n)(*[int(m) for m in group.modes])
        hosts = (ctypes.c_char_p * n)(*[(ep[0].encode() if ep else None) for ep in group.endpoints])
        ports = (I * n)(*[int(b or 0) for b in group.bulk_ports])
        sb, se = (I * n)(*[s[0] for s in spans]), (I * n)(*[s[1] for s in spans])
        ps, pe = (LL * len(parts))(*[p[0] for p in parts]), (LL * len(parts))(*[p[1] for p in parts])
        pc = (I * len(parts))(*[codec_id(p[2]) for p in parts])
        failed = (I * n)()
        gid = (ctypes.c_ubyte * 16)(*group.group_id[:16])
        desc = host.AllreduceDesc(n, me, modes, hosts, ports, sb, se, len(parts), ps, pe, pc, self._flat.data_ptr(),
                                  float(weight), gid, self._bulk_listener._l if self._bulk_listener else None,
                                  int(1000 * self.sender_timeout), failed)
        def run():
            with self.lock_averaged_tensors:      # the round reads and updates the buffer in place
                return host.allreduce_round(desc)

        rc = await asyncio.get_event_loop().run_in_executor(None, run)
        self.native_rounds += 1
        if rc < 0:
            raise AllreduceException(f"native all-reduce setup failed ({rc})")
        n_exchanges = sum(1 for k in range(n) if k != me and (spans[k][1] > spans[k][0] or group.modes[k] != AveragingMode.AUX))
        if rc > 0 and sum(failed) >= max(1, n_exchanges):
            raise AllreduceException(f"every exchange of the round failed ({list(failed)})")

    async def _reduce_own_span(self, rnd: _Round, p0: int, p1: int):
        parts = self._partition.parts
        if self.mode != AveragingMode.AUX:
            for j, i in enumerate(range(p0, p1)):
                s, e, _ = parts[i]
                rnd.reducer.add(j, rnd.my_index, self._flat[s:e].clone(), rnd.weight)
        else:
            pass  # contributes nothing; not in `senders`
        if self.mode == AveragingMode.AUX:
            for j in range(p1 - p0):  # wait so that the round is served until every part is reduced
                await rnd.reducer.result(j, self.sender_timeout * 4)
            return
        for j, i in enumerate(range(p0, p1)):
            try:
                avg = await rnd.reducer.result(j, self.sender_timeout)
            except asyncio.TimeoutError:
                # stragglers: drop every sender that has not delivered this part and finish with the rest
                for sdr in list(rnd.reducer.active - rnd.reducer.got[j]):
                    rnd.reducer.fail(sdr)
                avg = await rnd.reducer.result(j, 1.0)
            s, e, _ = parts[i]
            with self.lock_averaged_tensors:
                self._flat[s:e].copy_(avg)

    def _compress_part(self, x: torch.Tensor, tensor_numel: int):
        if isinstance(self.compression, C.SizeAdaptiveCompression):
            return self.compression.compress(x, tensor_numel)
        return self.compression.compress(x)

    async def _send_span(self, rnd: _Round, owner: int, q0: int, q1: int):
        group, parts = rnd.group, self._partition.parts
        stream = await self.dht.p2p.open_stream(group.endpoints[owner], f"{self.prefix}.allreduce",
                                                {"gid": group.group_id, "sender": rnd.my_index, "w": rnd.weight},
                                                timeout=self.request_timeout)
        loop = asyncio.get_event_loop()
        sent: Dict[int, torch.Tensor] = {}
        try:
            async def pump():
                for i in range(q0, q1):
                    s, e, tn = parts[i]
                    x = self._flat[s:e].clone()
                    msg = await loop.run_in_executor(None, self._compress_part, x, tn)
                    sent[i] = C.extract(msg)  # the owner sees the LOSSY version; the delta is relative to it
                    await stream.send({"i": i, "m": msg}, bulk=True)
                await stream.send({"done": 1})

            pump_task = asyncio.ensure_future(pump())
            received = 0
            while received < q1 - q0:
                msg = await stream.recv(self.sender_timeout * 2)
                if "err" in msg:
                    raise AllreduceException(msg["err"])
                i = msg["i"]
                delta = await loop.run_in_executor(None, C.extract, msg["m"])
                s, e, _ = parts[i]
                with self.lock_averaged_tensors:
                    # x <- x^ + delta would drop the local rounding error; hivemind applies x += delta
                    self._flat[s:e].add_(delta)
                received += 1
            await pump_task
        except Exception:
            raise
        finally:
            await stream.close()

    async def _rpc_allreduce(self, stream: Stream, payload):
        gid, sender, w = bytes(payload["gid"]), int(payload["sender"]), float(payload["w"])
        rnd = self._rounds.get(gid)
        if rnd is None:
            fut = self._round_waiters.setdefault(gid, asyncio.get_event_loop().create_future())
            try:
                rnd = await asyncio.wait_for(asyncio.shield(fut), self.sender_timeout)
            except asyncio.TimeoutError:
                await stream.send({"err": "unknown group"})
                await stream.close()
                return
        reducer, parts = rnd.reducer, self._partition.parts
        p0, p1 = self._spans(rnd.group)[rnd.my_index]
        loop = asyncio.get_event_loop()
        lossy: Dict[int, torch.Tensor] = {}
        try:
            async def replies():
                for j, i in enumerate(range(p0, p1)):
                    avg = await reducer.result(j, self.sender_timeout * 4)
                    while i not in lossy:
                        await asyncio.sleep(0.001)
                        if recv_task.done() and i not in lossy:
                            return
                    delta = avg - lossy.pop(i)
                    msg = await loop.run_in_executor(None, self._compress_part, delta, parts[i][2])
                    await stream.send({"i": i, "m": msg}, bulk=True)

            async def receive():
                while True:
                    msg = await stream.recv(self.sender_timeout)
                    if "done" in msg:
                        return
                    i = msg["i"]
                    x = await loop.run_in_executor(None, C.extract, msg["m"])
                    if not (p0 <= i < p1) or x.numel() != parts[i][1] - parts[i][0] or not torch.isfinite(x).all():
                        raise AllreduceException("malformed part")
                    lossy[i] = x
                    reducer.add(i - p0, sender, x, w)

            recv_task = asyncio.ensure_future(receive())
            reply_task = asyncio.ensure_future(replies())
            done, pending = await asyncio.wait({recv_task, reply_task}, return_when=asyncio.FIRST_EXCEPTION)
            for t in done:
                if t.exception() is not None:
                    raise t.exception()
            await asyncio.gather(*pending)
        except Exception:
            reducer.fail(sender)   # this sender no longer counts for the parts it did not deliver
            for t in (locals().get("recv_task"), locals().get("reply_task")):
                if t is not None:
                    t.cancel()
        finally:
            await stream.close()

    # ---- state sharing --------------------------------------------------------------------------
    def get_current_state(self):
        """-> (metadata, list of tensors). Override in subclasses (TrainingStateAverager)."""
        with self.get_tensors() as tensors:
            return {}, [t.clone() for t in tensors]

    async def _declare_loop(self):
        try:
            while True:
                if self.allow_state_sharing:
                    ep = self.dht.p2p.endpoint
                    await self.dht.node.store(f"{self.prefix}.all_averagers",
                                              {"ep": list(ep), "prio": float(self.state_sharing_priority)},
                                              get_dht_time() + 2 * self.declare_state_period, subkey=self.dht.peer_id.raw)
                await asyncio.sleep(self.declare_state_period)
        except asyncio.CancelledError:
            pass

    def declare_state_now(self):
        if self.allow_state_sharing and not self.client_mode:
            ep = self.dht.p2p.endpoint
            self.dht.store(f"{self.prefix}.all_averagers", {"ep": list(ep), "prio": float(self.state_sharing_priority)},
                           get_dht_time() + 2 * self.declare_state_period, subkey=self.dht.peer_id.raw)

    async def _rpc_download(self, stream: Stream, payload):
        try:
            if not self.allow_state_sharing:
                await stream.send({"err": "state sharing disabled"})
                return
            loop = asyncio.get_event_loop()
            meta, tensors = await loop.run_in_executor(None, self.get_current_state)
            use_bulk = bool(payload and payload.get("bulk")) and _bulk_available() and \
                isinstance(self.state_compression, C.NoCompression)
            if use_bulk:
                # tensors travel over the C++ bulk transport (raw frames, GIL released); control stays here
                from ..transport.bulk import BulkListener

                lst = BulkListener(self.dht.p2p.endpoint[0] if self.dht.p2p.endpoint else "127.0.0.1", 0)
                await stream.send({"meta": msgpack.packb(meta, use_bin_type=True), "n": len(tensors),
                                   "bulk_port": lst.port, "shapes": [list(t.shape) for t in tensors],
                                   "dtypes": [str(t.dtype) for t in tensors]})

                def serve():
                    conn = lst.accept(self.request_timeout * 3)
                    try:
                        if conn is not None:
                            for i, t in enumerate(tensors):
                                conn.send_tensor(t, tag=i)
                    finally:
                        if conn is not None:
                            conn.close()
                        lst.close()

                await loop.run_in_executor(None, serve)
                await stream.send({"done": 1})
                return
            await stream.send({"meta": msgpack.packb(meta, use_bin_type=True), "n": len(tensors)})
            for i, t in enumerate(tensors):
                flat = t.reshape(-1)
                for s in range(0, max(flat.numel(), 1), 1 << 22):
                    chunk = flat[s:s + (1 << 22)]
                    msg = self.state_compression.compress(chunk) if chunk.dtype.is_floating_point else \
                        {"c": "raw", "dt": str(chunk.dtype), "d": chunk.contiguous().numpy().tobytes()}
                    await stream.send({"i": i, "shape": list(t.shape), "m": msg}, bulk=True)
            await stream.send({"done": 1})
        except (ConnectionError, asyncio.IncompleteReadError):
            pass
        finally:
            await stream.close()

    def load_state_from_peers(self, wait: bool = True, timeout: Optional[float] = None):
        fut = asyncio.run_coroutine_threadsafe(self._load_state_from_peers(timeout), self.dht.loop)
        return fut.result() if wait else fut

    async def _load_state_from_peers(self, timeout):
        entry = await self.dht.node.get(f"{self.prefix}.all_averagers", latest=True)
        if entry is None or not isinstance(entry.value, dict):
            return None
        me = self.dht.peer_id.raw
        donors = sorted(((v.value["prio"], bytes(pid), tuple(v.value["ep"])) for pid, v in entry.value.items()
                         if bytes(pid) != me and v.value), reverse=True)
        for prio, pid, ep in donors:
            try:
                stream = await self.dht.p2p.open_stream(ep, f"{self.prefix}.download", {"bulk": _bulk_available()},
                                                        timeout=self.request_timeout)
                try:
                    head = await stream.recv(timeout or 30.0)
                    if "err" in head:
                        continue
                    meta = msgpack.unpackb(head["meta"], raw=False, strict_map_key=False)
                    if "bulk_port" in head:
                        from ..transport.bulk import BulkConnection

                        def fetch():
                            conn = BulkConnection.connect((ep[0], head["bulk_port"]), self.request_timeout)
                            try:
                                out = []
                                for shape, dt in zip(head["shapes"], head["dtypes"]):
                                    t = torch.empty(shape, dtype=getattr(torch, dt.replace("torch.", "")))
                                    conn.recv_into(t, timeout or 60.0)
                                    out.append(t)
                                return out
                            finally:
                                conn.close()

                        tensors = await asyncio.get_event_loop().run_in_executor(None, fetch)
                        return meta, tensors
                    chunks: Dict[int, list] = {}
                    shapes: Dict[int, list] = {}
                    while True:
                        msg = await stream.recv(timeout or 60.0)
                        if "done" in msg:
                            break
                        m = msg["m"]
                        if m["c"] == "raw":
                            dt = getattr(torch, m["dt"].replace("torch.", ""))
                            t = torch.frombuffer(bytearray(m["d"]), dtype=dt) if len(m["d"]) else torch.zeros(0, dtype=dt)
                        else:
                            t = C.extract(m)
                        chunks.setdefault(msg["i"], []).append(t)
                        shapes[msg["i"]] = msg["shape"]
                    tensors = [torch.cat(chunks[i]).view(shapes[i]) if i in chunks else torch.zeros(shapes.get(i, [0]))
                               for i in range(head["n"])]
                    return meta, tensors
                finally:
                    await stream.close()
            except Exception:
                continue
        return None
