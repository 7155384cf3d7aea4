"""R12 / R14-R16 / H12: token authorization, alternative optimizers, fault injection."""
import asyncio
import threading
from datetime import datetime, timedelta

import torch

from dalle_b200.collab.auth import AccessToken, TokenAuthorizerBase
from dalle_b200.dht import DHT, get_dht_time
from dalle_b200.dht.crypto import RSAPrivateKey
from dalle_b200.optim.extras import LambWithGradientClipping, OffloadOptimizer


class _Authority:
    def __init__(self):
        self.key = RSAPrivateKey()

    def issue(self, username, public_key: bytes, minutes=5):
        t = AccessToken(username=username, public_key=public_key,
                        expiration_time=str(datetime.utcnow() + timedelta(minutes=minutes)))
        t.signature = self.key.sign(f"{t.username} {t.public_key} {t.expiration_time}".encode())
        return t


class _Authorizer(TokenAuthorizerBase):
    def __init__(self, authority: _Authority, username: str, minutes=5):
        super().__init__(RSAPrivateKey())
        self.authority, self.username, self.minutes = authority, username, minutes

    async def get_token(self):
        return self.authority.issue(self.username, self.local_public_key.to_bytes(), self.minutes)

    def is_token_valid(self, t: AccessToken) -> bool:
        ok = self.authority.key.get_public_key().verify(f"{t.username} {t.public_key} {t.expiration_time}".encode(), t.signature)
        return ok and datetime.fromisoformat(t.expiration_time) >= datetime.utcnow()

    def does_token_need_refreshing(self, t: AccessToken) -> bool:
        return datetime.fromisoformat(t.expiration_time) < datetime.utcnow() + timedelta(minutes=1)


def test_token_authorizer_signs_and_validates_requests():
    auth = _Authority()
    a, b = _Authorizer(auth, "alice"), _Authorizer(auth, "bob")
    fields = asyncio.run(a.sign_request_fields())
    assert asyncio.run(b.validate_request_fields(fields))
    assert not asyncio.run(b.validate_request_fields(fields))          # replayed nonce
    forged = asyncio.run(a.sign_request_fields())
    forged["token"]["username"] = "mallory"                             # token no longer matches its signature
    assert not asyncio.run(b.validate_request_fields(forged))
    other = _Authorizer(_Authority(), "eve")                            # token from a different authority
    assert not asyncio.run(b.validate_request_fields(asyncio.run(other.sign_request_fields())))
    assert not asyncio.run(b.validate_request_fields(None))


def test_dht_rejects_unauthorized_peers():
    auth = _Authority()
    first = DHT(start=True, host_maddrs=["/ip4/127.0.0.1/tcp/0"], authorizer=_Authorizer(auth, "a"))
    peers = [str(m) for m in first.get_visible_maddrs()]
    try:
        good = DHT(initial_peers=peers, start=True, host_maddrs=["/ip4/127.0.0.1/tcp/0"], authorizer=_Authorizer(auth, "b"))
        assert good.store("k", 1, get_dht_time() + 10) and first.get("k").value == 1
        good.shutdown()
        try:
            DHT(initial_peers=peers, start=True, host_maddrs=["/ip4/127.0.0.1/tcp/0"])   # no credentials
            raise AssertionError("an unauthorized peer joined")
        except RuntimeError:
            pass
    finally:
        first.shutdown()


def test_clipped_lamb_and_offload_optimizer():
    torch.manual_seed(0)
    w = torch.nn.Parameter(torch.randn(10, 10))
    w2 = torch.nn.Parameter(w.detach().clone())
    direct = LambWithGradientClipping([w], lr=0.1, max_grad_norm=1.0, weight_decay=0.01)
    off = OffloadOptimizer([w2], LambWithGradientClipping, lr=0.1, max_grad_norm=1.0, weight_decay=0.01)
    for _ in range(3):
        g = torch.randn(10, 10)
        w.grad, w2.grad = g.clone(), g.clone()
        direct.step(); off.step()
    assert torch.allclose(w, w2, atol=1e-6)
    assert off.state_dict()["state"][0]["step"] == 3


def test_allreduce_survives_a_peer_that_dies_midway(monkeypatch):
    """Fault injection: one sender's stream is cut; the survivors still finish the round."""
    from dalle_b200.collab.averager import DecentralizedAverager
    import dalle_b200.transport.p2p as p2p

    d0 = DHT(start=True, host_maddrs=["/ip4/127.0.0.1/tcp/0"])
    peers = [str(m) for m in d0.get_visible_maddrs()]
    d1 = DHT(initial_peers=peers, start=True, host_maddrs=["/ip4/127.0.0.1/tcp/0"])
    d2 = DHT(initial_peers=peers, start=True, host_maddrs=["/ip4/127.0.0.1/tcp/0"])
    try:
        data = [[torch.full((200_000,), float(i + 1))] for i in range(3)]
        avs = [DecentralizedAverager(data[i], d, prefix="fault", min_matchmaking_time=2.0, target_group_size=3,
                                     allreduce_timeout=20, sender_timeout=2.0, part_size_bytes=2 ** 16)
               for i, d in enumerate((d0, d1, d2))]
        victim_pid = d2.peer_id.raw
        orig = p2p.write_frame
        sent = {"n": 0}

        async def flaky(writer, obj, bulk=False):
            if bulk and isinstance(obj, dict) and "m" in obj and flaky.active:
                sent["n"] += 1
                if sent["n"] > 3:
                    raise ConnectionResetError("injected failure")
            return await orig(writer, obj, bulk)

        flaky.active = False
        results = [None] * 3

        def work(i):
            results[i] = avs[i].step(weight=1.0, timeout=25, allow_retries=False)

        ts = [threading.Thread(target=work, args=(i,)) for i in range(3)]
        # the victim's outgoing bulk frames start failing after a few parts
        monkeypatch.setattr(p2p.Stream, "send", lambda self, obj, bulk=False: (
            flaky(self.writer, obj, bulk) if getattr(self, "_victim", False) else orig(self.writer, obj, bulk)))
        orig_open = avs[2].dht.p2p.open_stream

        async def tagging_open(*a, **k):
            s = await orig_open(*a, **k)
            s._victim = True
            return s

        avs[2].dht.p2p.open_stream = tagging_open
        flaky.active = True
        [t.start() for t in ts]
        [t.join(60) for t in ts]
        assert results[0] is not None and results[1] is not None      # survivors completed the round
        for i in (0, 1):
            with avs[i].get_tensors() as tt:
                assert torch.isfinite(tt[0]).all()
                assert 1.0 <= tt[0].min().item() and tt[0].max().item() <= 3.0
    finally:
        for d in (d0, d1, d2):
            d.shutdown()


def test_phase_trace_is_inert_without_cuda_and_times_with_it():
    """phase tracing (NVTX + CUDA events) must be free when off and never change results."""
    import torch
    from dalle_b200.utils.phases import PhaseTrace

    t = PhaseTrace(enabled=True)
    with t.phase("forward"):
        x = torch.ones(4).sum()
    assert x.item() == 4.0
    s = t.summary()
    if torch.cuda.is_available():
        assert "forward" in s and s["forward"] >= 0.0
    else:
        assert s == {}
