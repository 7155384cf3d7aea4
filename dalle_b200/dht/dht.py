"""``DHT``: thread-hosted DHT node with the hivemind.DHT surface the reference touches:
constructor keywords of reference task.py:104-114, ``peer_id``, ``get_visible_maddrs()``,
``store(key, subkey, value, expiration_time, return_future)``, ``get(key, latest)`` -> object with ``.value``
(dict sub-key -> object with ``.value``), ``shutdown()``.

The node runs on a private asyncio loop in a daemon thread (hivemind forks a process); the same loop and the same
``P2P`` endpoint are shared by the averagers, so one TCP port serves DHT RPCs, matchmaking and tensor streams.
"""
from __future__ import annotations

import asyncio
import concurrent.futures
import os
import threading
from typing import Any, Iterable, List, Optional, Sequence

from ..transport.p2p import P2P, Multiaddr, PeerID
from .node import DHTNode
from .storage import ValueWithExpiration, get_dht_time
from .validation import CompositeValidator, RecordValidatorBase


class DHT:
    def __init__(self, initial_peers: Sequence[str] = (), *, start: bool = False, client_mode: bool = False,
                 host_maddrs: Sequence[str] = ("/ip4/0.0.0.0/tcp/0",), announce_maddrs: Sequence[str] = (),
                 use_ipfs: bool = False, record_validators: Iterable[RecordValidatorBase] = (),
                 identity_path: Optional[str] = None, authorizer=None, num_replicas: int = 5, **kwargs):
        if use_ipfs:
            raise ValueError("--use_ipfs needs the public IPFS bootstrap network; this transport is plain TCP without "
                             "internet access - pass explicit /ip4/.../tcp/... multiaddrs in --initial_peers instead")
        self.initial_peers = [str(p) for p in initial_peers]
        self.client_mode = client_mode
        self._validator = CompositeValidator(record_validators)
        self.authorizer = authorizer
        host, port = "0.0.0.0", 0
        if host_maddrs:
            m = Multiaddr(host_maddrs[0])
            host, port = m.host or host, m.port or 0
        announce = Multiaddr(announce_maddrs[0]).host if announce_maddrs else None
        identity = None
        if identity_path is not None:
            if os.path.exists(identity_path):
                identity = open(identity_path, "rb").read()
            else:
                identity = os.urandom(32)
                with open(identity_path, "wb") as f:
                    f.write(identity)
        self._p2p = P2P(identity=identity, host=host, port=port, announce_host=announce, client_mode=client_mode)
        self._num_replicas = num_replicas
        self._loop: Optional[asyncio.AbstractEventLoop] = None
        self._thread: Optional[threading.Thread] = None
        self._node: Optional[DHTNode] = None
        self._ready = threading.Event()
        self._error: Optional[BaseException] = None
        self.is_alive = False
        if start:
            self.run_in_background(await_ready=True)

    # ---- lifecycle ------------------------------------------------------------------------------
    def run_in_background(self, await_ready: bool = True, timeout: float = 30.0):
        self._thread = threading.Thread(target=self._run, name="DHT", daemon=True)
        self._thread.start()
        if await_ready:
            if not self._ready.wait(timeout):
                raise TimeoutError("DHT did not start in time")
            if self._error is not None:
                raise self._error

    def _run(self):
        self._loop = asyncio.new_event_loop()
        asyncio.set_event_loop(self._loop)

        async def boot():
            await self._p2p.start()
            self._node = DHTNode(self._p2p, self._validator, num_replicas=self._num_replicas,
                                 authorizer=self.authorizer)
            if self.initial_peers:
                await self._node.bootstrap(self.initial_peers)

        try:
            self._loop.run_until_complete(boot())
            self.is_alive = True
        except BaseException as e:  # surfaced to the constructor
            self._error = e
            self._ready.set()
            return
        self._ready.set()
        try:
            self._loop.run_forever()
        finally:
            self.is_alive = False

    def shutdown(self):
        if self._loop is not None and self._loop.is_running():
            fut = asyncio.run_coroutine_threadsafe(self._p2p.shutdown(), self._loop)
            try:
                fut.result(2.0)
            except Exception:
                pass

            def _cancel_all():
                for t in asyncio.all_tasks(self._loop):
                    t.cancel()
                self._loop.call_later(0.05, self._loop.stop)

            self._loop.call_soon_threadsafe(_cancel_all)
        if self._thread is not None:
            self._thread.join(2.0)
        self.is_alive = False

    # ---- plumbing shared with the averagers --------------------------------------------------------
    @property
    def p2p(self) -> P2P:
        return self._p2p

    @property
    def loop(self) -> asyncio.AbstractEventLoop:
        return self._loop

    @property
    def node(self) -> DHTNode:
        return self._node

    def run_coroutine(self, coro, return_future: bool = False):
        fut = asyncio.run_coroutine_threadsafe(coro, self._loop)
        return fut if return_future else fut.result()

    # ---- reference surface ----------------------------------------------------------------------
    @property
    def peer_id(self) -> PeerID:
        return self._p2p.peer_id

    def get_visible_maddrs(self, latest: bool = False) -> List[Multiaddr]:
        return self._p2p.visible_maddrs()

    def add_validators(self, validators: Iterable[RecordValidatorBase]):
        self._validator.extend(validators)

    def store(self, key, value, expiration_time: float, subkey=None, return_future: bool = False):
        return self.run_coroutine(self._node.store(key, value, expiration_time, subkey), return_future)

    def get(self, key, latest: bool = False, return_future: bool = False):
        return self.run_coroutine(self._node.get(key, latest), return_future)
