"""T0/T2: DHT store/get/expiry/subkeys/signatures/schema over localhost TCP (hivemind-style: several peers in
one process)."""
import time

import pytest
from pydantic import BaseModel, StrictFloat, conint

from dalle_b200.dht import (DHT, BytesWithPublicKey, RSASignatureValidator, SchemaValidator, get_dht_time)
from dalle_b200.dht.storage import DHTLocalStorage


def _swarm(n, **kw):
    first = DHT(start=True, host_maddrs=["/ip4/127.0.0.1/tcp/0"], **kw)
    peers = [str(m) for m in first.get_visible_maddrs()]
    rest = [DHT(initial_peers=peers, start=True, host_maddrs=["/ip4/127.0.0.1/tcp/0"], **kw) for _ in range(n - 1)]
    return [first] + rest


def test_local_storage_expiration_and_subkeys():
    s = DHTLocalStorage()
    now = get_dht_time()
    assert s.store(b"k", b"v1", now + 10)
    assert not s.store(b"k", b"v0", now + 5)      # older expiration loses
    assert s.store(b"k", b"v2", now + 20)
    assert s.get(b"k").value == b"v2"
    assert not s.store(b"e", b"x", now - 1)
    assert s.store(b"d", b"a", now + 10, subkey=b"s1") and s.store(b"d", b"b", now + 0.2, subkey=b"s2")
    assert len(s.get(b"d").value) == 2
    time.sleep(0.25)
    assert len(s.get(b"d").value) == 1


def test_store_get_across_peers_and_client_mode():
    dhts = _swarm(4)
    try:
        assert dhts[1].store("key1", {"a": 1}, get_dht_time() + 30)
        for d in dhts:
            assert d.get("key1").value == {"a": 1}
        assert dhts[2].store("dict", 1.5, get_dht_time() + 30, subkey="p2")
        assert dhts[3].store("dict", 2.5, get_dht_time() + 30, subkey="p3")
        got = dhts[0].get("dict", latest=True)
        assert {k: v.value for k, v in got.value.items()} == {"p2": 1.5, "p3": 2.5}
        client = DHT(initial_peers=[str(m) for m in dhts[0].get_visible_maddrs()], start=True, client_mode=True)
        assert client.get_visible_maddrs() == []
        assert client.get("key1").value == {"a": 1}
        assert client.store("from_client", 7, get_dht_time() + 30)
        assert dhts[3].get("from_client").value == 7
        client.shutdown()
        assert dhts[0].get("missing") is None
        assert dhts[1].store("short", 1, get_dht_time() + 0.3)
        time.sleep(0.4)
        assert dhts[2].get("short") is None
    finally:
        for d in dhts:
            d.shutdown()


class _Metrics(BaseModel):
    step: conint(ge=0, strict=True)
    loss: StrictFloat


class _Schema(BaseModel):
    metrics: dict[BytesWithPublicKey, _Metrics]


def test_signature_and_schema_validators():
    from dalle_b200.dht.crypto import RSAPrivateKey

    va, vb = RSASignatureValidator(RSAPrivateKey()), RSASignatureValidator(RSAPrivateKey())
    a = DHT(start=True, host_maddrs=["/ip4/127.0.0.1/tcp/0"], record_validators=[SchemaValidator(_Schema, prefix="exp"), va])
    b = DHT(initial_peers=[str(m) for m in a.get_visible_maddrs()], start=True, host_maddrs=["/ip4/127.0.0.1/tcp/0"],
            record_validators=[SchemaValidator(_Schema, prefix="exp"), vb])
    try:
        exp = get_dht_time() + 30
        assert a.store("exp_metrics", {"step": 1, "loss": 0.5}, exp, subkey=va.local_public_key)
        assert b.store("exp_metrics", {"step": 2, "loss": 0.25}, exp, subkey=vb.local_public_key)
        got = b.get("exp_metrics", latest=True).value
        assert {v.value["step"] for v in got.values()} == {1, 2}
        assert set(got.keys()) == {va.local_public_key, vb.local_public_key}
        # wrong schema (loss must be a strict float) is rejected everywhere
        assert not a.store("exp_metrics", {"step": 1, "loss": "bad"}, exp + 1, subkey=va.local_public_key)
        # b cannot write under a's public key: the signature will not verify
        assert not b.store("exp_metrics", {"step": 9, "loss": 0.1}, exp + 2, subkey=va.local_public_key)
        assert b.get("exp_metrics", latest=True).value[va.local_public_key].value["step"] == 1
    finally:
        a.shutdown(); b.shutdown()


def test_use_ipfs_is_rejected_with_a_clear_message():
    with pytest.raises(ValueError, match="plain TCP"):
        DHT(use_ipfs=True)
