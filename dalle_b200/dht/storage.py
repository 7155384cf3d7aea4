"""Local DHT storage with expirations and sub-keyed dictionary values.

Semantics kept from hivemind: every record carries an absolute expiration time; a store only succeeds if its
expiration is LATER than the stored one; dictionary values merge per sub-key with the same rule; expired records
vanish on access.
"""
from __future__ import annotations

import heapq
import time
from typing import Any, Dict, NamedTuple, Optional


def get_dht_time() -> float:
    return time.time()


class ValueWithExpiration(NamedTuple):
    value: Any
    expiration_time: float


class DictionaryDHTValue:
    """A value that is a dict sub-key -> ValueWithExpiration; the record expires when its newest sub-key does."""

    def __init__(self):
        self.data: Dict[Any, ValueWithExpiration] = {}

    def store(self, subkey, value, expiration_time: float) -> bool:
        old = self.data.get(subkey)
        if old is not None and old.expiration_time >= expiration_time:
            return False
        self.data[subkey] = ValueWithExpiration(value, expiration_time)
        return True

    def prune(self, now: float):
        for k in [k for k, v in self.data.items() if v.expiration_time <= now]:
            del self.data[k]

    @property
    def latest_expiration_time(self) -> float:
        return max((v.expiration_time for v in self.data.values()), default=-float("inf"))

    def items(self):
        return self.data.items()

    def __len__(self):
        return len(self.data)


class DHTLocalStorage:
    def __init__(self, maxsize: Optional[int] = None):
        self.data: Dict[bytes, ValueWithExpiration] = {}
        self.maxsize = maxsize

    def _prune(self, key: bytes, now: float):
        rec = self.data.get(key)
        if rec is None:
            return
        if isinstance(rec.value, DictionaryDHTValue):
            rec.value.prune(now)
            if len(rec.value) == 0:
                del self.data[key]
        elif rec.expiration_time <= now:
            del self.data[key]

    def store(self, key: bytes, value, expiration_time: float, subkey=None) -> bool:
        now = get_dht_time()
        if expiration_time <= now:
            return False
        self._prune(key, now)
        rec = self.data.get(key)
        if subkey is not None:
            if rec is None or not isinstance(rec.value, DictionaryDHTValue):
                if rec is not None and rec.expiration_time >= expiration_time:
                    return False
                d = DictionaryDHTValue()
                d.store(subkey, value, expiration_time)
                self.data[key] = ValueWithExpiration(d, expiration_time)
                return True
            ok = rec.value.store(subkey, value, expiration_time)
            self.data[key] = ValueWithExpiration(rec.value, rec.value.latest_expiration_time)
            return ok
        if rec is not None and rec.expiration_time >= expiration_time:
            return False
        self.data[key] = ValueWithExpiration(value, expiration_time)
        if self.maxsize is not None and len(self.data) > self.maxsize:
            oldest = min(self.data, key=lambda k: self.data[k].expiration_time)
            del self.data[oldest]
        return True

    def get(self, key: bytes) -> Optional[ValueWithExpiration]:
        self._prune(key, get_dht_time())
        return self.data.get(key)

    def __len__(self):
        return len(self.data)
