"""Kademlia-style DHT node over the TCP P2P layer: 160-bit ids, XOR metric, iterative nearest-peer
search, replicated stores, expirations, sub-keyed values, record validation on every store and every read."""
from __future__ import annotations

import asyncio
import hashlib
import time
from typing import Any, Dict, List, Optional, Tuple

from ..transport.p2p import P2P, PeerID, parse_endpoint
from .storage import DHTLocalStorage, DictionaryDHTValue, ValueWithExpiration, get_dht_time
from .validation import IS_REGULAR_VALUE, CompositeValidator, DHTRecord, pack, unpack


def key_id(raw_key: bytes) -> bytes:
    return hashlib.sha1(raw_key).digest()


def xor_distance(a: bytes, b: bytes) -> int:
    return int.from_bytes(a, "big") ^ int.from_bytes(b, "big")


class DHTNode:
    def __init__(self, p2p: P2P, validator: CompositeValidator, *, num_replicas: int = 5, beam: int = 3,
                 rpc_timeout: float = 5.0, blacklist_time: float = 5.0, authorizer=None):
        self.p2p, self.validator = p2p, validator
        self.node_id = key_id(p2p.peer_id.raw)
        self.storage = DHTLocalStorage()
        self.raw_keys: Dict[bytes, bytes] = {}
        self.peers: Dict[PeerID, Tuple[Tuple[str, int], bytes]] = {}   # peer -> (endpoint, node id)
        self.blacklist: Dict[PeerID, float] = {}
        self.num_replicas, self.beam, self.rpc_timeout, self.blacklist_time = num_replicas, beam, rpc_timeout, blacklist_time
        self.authorizer = authorizer
        p2p.add_unary_handler("dht.ping", self._rpc_ping)
        p2p.add_unary_handler("dht.store", self._rpc_store)
        p2p.add_unary_handler("dht.find", self._rpc_find)

    # ---- membership -----------------------------------------------------------------------------
    def _me(self):
        ep = self.p2p.endpoint
        return {"pid": self.p2p.peer_id.raw, "ep": list(ep) if ep else None}

    def _learn(self, info: Optional[dict]):
        if not info or not info.get("ep"):
            return  # client-mode peers are never routed to
        pid = PeerID(info["pid"])
        if pid != self.p2p.peer_id:
            self.peers[pid] = (tuple(info["ep"]), key_id(pid.raw))

    async def _call(self, pid: Optional[PeerID], ep, name, payload):
        try:
            payload = dict(payload, frm=self._me())
            if self.authorizer is not None:
                payload["auth"] = await self.authorizer.sign_request_fields()
            res = await self.p2p.call(ep, name, payload, timeout=self.rpc_timeout)
            self._learn(res.get("frm"))
            return res
        except Exception:
            if pid is not None:
                self.peers.pop(pid, None)
                self.blacklist[pid] = time.monotonic() + self.blacklist_time
            return None

    async def bootstrap(self, initial_peers: List[str]):
        eps = [parse_endpoint(m)[0] for m in initial_peers]
        results = await asyncio.gather(*(self._call(None, ep, "dht.ping", {}) for ep in eps))
        if eps and not any(results):
            raise RuntimeError(f"could not reach any of the initial peers {initial_peers}")
        await self.find_nearest(self.node_id)

    async def _check_auth(self, payload):
        if self.authorizer is not None and not await self.authorizer.validate_request_fields(payload.get("auth")):
            raise PermissionError("request failed authorization")

    async def _rpc_ping(self, payload, sender):
        await self._check_auth(payload)
        self._learn(payload.get("frm"))
        return {"frm": self._me()}

    def _nearest_known(self, kid: bytes, k: int, exclude=()) -> List[list]:
        now = time.monotonic()
        cands = [(xor_distance(nid, kid), pid, ep) for pid, (ep, nid) in self.peers.items()
                 if pid not in exclude and self.blacklist.get(pid, 0) < now]
        cands.sort(key=lambda t: t[0])
        return [[pid.raw, ep[0], ep[1]] for _, pid, ep in cands[:k]]

    async def find_nearest(self, kid: bytes, k: Optional[int] = None) -> List[Tuple[PeerID, Tuple[str, int]]]:
        """Iterative lookup: repeatedly ask the closest not-yet-queried peers for their closest peers."""
        k = k or self.num_replicas
        queried = set()
        while True:
            cands = [(PeerID(raw), (h, p)) for raw, h, p in self._nearest_known(kid, k)]
            todo = [(pid, ep) for pid, ep in cands if pid not in queried][: self.beam]
            if not todo:
                return cands
            queried.update(pid for pid, _ in todo)
            res = await asyncio.gather(*(self._call(pid, ep, "dht.find", {"kid": kid, "peers_only": True})
                                         for pid, ep in todo))
            for r in res:
                for raw, h, p in (r or {}).get("nearest", []):
                    self._learn({"pid": raw, "ep": [h, p]})

    # ---- store ----------------------------------------------------------------------------------
    def _store_local(self, kid, raw_key, subkey_b, value_b, exp) -> bool:
        record = DHTRecord(raw_key, subkey_b, value_b, exp)
        if not self.validator.validate(record):
            return False
        self.raw_keys[kid] = raw_key
        sk = None if subkey_b == IS_REGULAR_VALUE else subkey_b
        return self.storage.store(kid, value_b, exp, subkey=sk)

    async def _rpc_store(self, payload, sender):
        await self._check_auth(payload)
        self._learn(payload.get("frm"))
        ok = self._store_local(payload["kid"], payload["raw"], payload["sk"], payload["v"], payload["exp"])
        return {"ok": ok, "frm": self._me()}

    async def store(self, key: Any, value: Any, expiration_time: float, subkey: Any = None) -> bool:
        raw_key = key.encode() if isinstance(key, str) else bytes(key) if isinstance(key, (bytes, bytearray)) else pack(key)
        kid = key_id(raw_key)
        subkey_b = IS_REGULAR_VALUE if subkey is None else pack(subkey)
        record = DHTRecord(raw_key, subkey_b, pack(value), expiration_time)
        value_b = self.validator.sign_value(record)
        nearest = await self.find_nearest(kid)
        # we are one of the replicas when we are closer than the furthest chosen peer (or the swarm is small)
        my_d = xor_distance(self.node_id, kid)
        store_self = not self.p2p.client_mode and (
            len(nearest) < self.num_replicas or my_d < xor_distance(key_id(nearest[-1][0].raw), kid))
        ok_local = self._store_local(kid, raw_key, subkey_b, value_b, expiration_time) if store_self else False
        msg = {"kid": kid, "raw": raw_key, "sk": subkey_b, "v": value_b, "exp": expiration_time}
        targets = nearest[: self.num_replicas - (1 if store_self else 0)] if store_self else nearest
        res = await asyncio.gather(*(self._call(pid, ep, "dht.store", msg) for pid, ep in targets))
        return ok_local or any(r and r.get("ok") for r in res)

    # ---- get ------------------------------------------------------------------------------------
    def _serialize_local(self, kid):
        rec = self.storage.get(kid)
        if rec is None:
            return None
        if isinstance(rec.value, DictionaryDHTValue):
            return {"raw": self.raw_keys.get(kid, b""), "d": [[sk, v.value, v.expiration_time] for sk, v in rec.value.items()]}
        return {"raw": self.raw_keys.get(kid, b""), "r": [rec.value, rec.expiration_time]}

    async def _rpc_find(self, payload, sender):
        await self._check_auth(payload)
        self._learn(payload.get("frm"))
        kid = payload["kid"]
        out = {"nearest": self._nearest_known(kid, self.num_replicas, exclude=(sender,)), "frm": self._me()}
        if not payload.get("peers_only"):
            out["value"] = self._serialize_local(kid)
        return out

    async def get(self, key: Any, latest: bool = False) -> Optional[ValueWithExpiration]:
        raw_key = key.encode() if isinstance(key, str) else bytes(key) if isinstance(key, (bytes, bytearray)) else pack(key)
        kid = key_id(raw_key)
        found = [self._serialize_local(kid)]
        nearest = await self.find_nearest(kid)
        res = await asyncio.gather(*(self._call(pid, ep, "dht.find", {"kid": kid}) for pid, ep in nearest))
        found += [r.get("value") for r in res if r]
        regular: Optional[Tuple[bytes, float]] = None
        merged: Dict[bytes, Tuple[bytes, float]] = {}
        now = get_dht_time()
        for f in found:
            if not f:
                continue
            if "r" in f:
                v, exp = f["r"]
                if exp > now and self.validator.validate(DHTRecord(raw_key, IS_REGULAR_VALUE, v, exp)):
                    if regular is None or exp > regular[1]:
                        regular = (v, exp)
            for sk, v, exp in f.get("d", []):
                if exp > now and self.validator.validate(DHTRecord(raw_key, sk, v, exp)):
                    if sk not in merged or exp > merged[sk][1]:
                        merged[sk] = (v, exp)
        if merged:
            out = {}
            for sk, (v, exp) in merged.items():
                val = unpack(self.validator.strip_value(DHTRecord(raw_key, sk, v, exp)))
                out[unpack(sk)] = ValueWithExpiration(val, exp)
                self._store_local(kid, raw_key, sk, v, exp)  # cache
            return ValueWithExpiration(out, max(e for _, e in merged.values()))
        if regular is not None:
            v, exp = regular
            return ValueWithExpiration(unpack(self.validator.strip_value(DHTRecord(raw_key, IS_REGULAR_VALUE, v, exp))), exp)
        return None
