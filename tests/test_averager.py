"""T2: matchmaking + butterfly all-reduce between several peers on localhost (threads, like hivemind's tests)."""
import threading
import time

import pytest
import torch

from dalle_b200.collab import compression as C
from dalle_b200.collab.averager import DecentralizedAverager, Partition
from dalle_b200.dht import DHT


def _dhts(n, client_last=False):
    first = DHT(start=True, host_maddrs=["/ip4/127.0.0.1/tcp/0"])
    peers = [str(m) for m in first.get_visible_maddrs()]
    out = [first]
    for i in range(1, n):
        cm = client_last and i == n - 1
        out.append(DHT(initial_peers=peers, start=True, host_maddrs=["/ip4/127.0.0.1/tcp/0"], client_mode=cm))
    return out


def _run(averagers, weights, **kw):
    results = [None] * len(averagers)

    def work(i):
        results[i] = averagers[i].step(weight=weights[i], gather={"i": i}, timeout=20, **kw)

    ts = [threading.Thread(target=work, args=(i,)) for i in range(len(averagers))]
    [t.start() for t in ts]
    [t.join(40) for t in ts]
    return results


def test_partition_regions_and_spans():
    comp = C.SizeAdaptiveCompression(2 ** 16 + 1, C.Float16Compression(), C.Uniform8BitQuantization())
    p = Partition([1000, 70000, 200000, 10], comp, part_elems=65536)
    assert p.parts[0] == (0, 1000, 1000)                 # small tensor: own fp16 region
    assert all(e - s <= 65536 for s, e, _ in p.parts) and p.parts[-1][1] == p.total
    spans = p.assign([1.0, 0.0, 1.0])
    assert spans[1][0] == spans[1][1] and spans[0][1] == spans[2][0] and spans[2][1] == len(p.parts)


@pytest.mark.parametrize("plane", ["native", "asyncio"])
@pytest.mark.parametrize("compression", ["none", "adaptive"])
def test_weighted_allreduce_three_peers(compression, plane, monkeypatch):
    """Both data planes: the C++ thread pipeline (native/allreduce_pipe.cpp) and the asyncio part loop."""
    monkeypatch.setenv("DALLE_B200_NATIVE_ALLREDUCE", "1" if plane == "native" else "0")
    comp = C.NoCompression() if compression == "none" else C.SizeAdaptiveCompression(
        2 ** 16 + 1, C.Float16Compression(), C.Uniform8BitQuantization())
    dhts = _dhts(3)
    try:
        g = torch.Generator().manual_seed(0)
        shapes = [(300,), (70, 1000), (17,)]
        data = [[torch.randn(*s, generator=g) for s in shapes] for _ in range(3)]
        avs = [DecentralizedAverager(data[i], dhts[i], prefix="t1", min_matchmaking_time=4.0, target_group_size=3,
                                     compression=comp, part_size_bytes=2 ** 16, allreduce_timeout=15, start=True)
               for i in range(3)]
        weights = [1.0, 2.0, 5.0]
        res = _run(avs, weights)
        assert all(r is not None and len(r) == 3 for r in res), res
        assert {v["i"] for v in res[0].values()} == {0, 1, 2}
        assert all(a.native_rounds == (1 if plane == "native" else 0) for a in avs), [a.native_rounds for a in avs]
        for k in range(len(shapes)):
            ref = sum(w * d[k] for w, d in zip(weights, data)) / sum(weights)
            for a in avs:
                with a.get_tensors() as ts:
                    if compression == "none":
                        assert (ts[k] - ref).abs().max().item() < 1e-5
                    else:  # 8-bit buckets clip the tails (values beyond 3 sigma share one bucket): judge by norm
                        assert ((ts[k] - ref).norm() / ref.norm()).item() < 0.05
    finally:
        for d in dhts:
            d.shutdown()


def test_client_mode_and_aux_peer():
    dhts = _dhts(3, client_last=True)
    try:
        data = [[torch.full((5000,), float(i + 1))] for i in range(3)]
        avs = [DecentralizedAverager(data[0], dhts[0], prefix="t2", min_matchmaking_time=2.0, target_group_size=3,
                                     allreduce_timeout=15),
               DecentralizedAverager(data[1], dhts[1], prefix="t2", min_matchmaking_time=2.0, target_group_size=3,
                                     allreduce_timeout=15, auxiliary=True),
               DecentralizedAverager(data[2], dhts[2], prefix="t2", min_matchmaking_time=2.0, target_group_size=3,
                                     allreduce_timeout=15)]
        res = _run(avs, [1.0, 0.0, 3.0])
        assert all(r is not None for r in res)
        assert all(a.native_rounds == 1 for a in avs)   # aux owner and client sender both speak the C++ data plane
        ref = (1.0 * 1 + 3.0 * 3) / 4.0   # the auxiliary peer contributes nothing
        for i in (0, 2):
            with avs[i].get_tensors() as ts:
                assert torch.allclose(ts[0], torch.full((5000,), ref), atol=1e-5)
    finally:
        for d in dhts:
            d.shutdown()


def test_native_round_drops_a_sender_with_non_finite_values_and_averages_the_rest():
    """hivemind's reducer bans a sender whose part is malformed and finishes with the others: the C++ owner refuses a
    part with Inf in it, the two healthy peers still converge to their own mean on that span."""
    dhts = _dhts(3)
    try:
        data = [[torch.full((4000,), 1.0)], [torch.full((4000,), 3.0)], [torch.full((4000,), float("inf"))]]
        avs = [DecentralizedAverager(data[i], dhts[i], prefix="t7", min_matchmaking_time=2.0, target_group_size=3,
                                     allreduce_timeout=15, sender_timeout=3.0, part_size_bytes=2 ** 12) for i in range(3)]
        res = _run(avs, [1.0, 1.0, 1.0])
        assert res[0] is not None and res[1] is not None
        assert all(a.native_rounds == 1 for a in avs)
        own = [a._spans(a.last_group or avs[0].last_group) for a in avs[:1]][0]
        parts = avs[0]._partition.parts
        bad_owner = avs[0].last_group.peer_ids.index(dhts[2].peer_id.raw)
        for i in (0, 1):
            with avs[i].get_tensors() as ts:
                for k, (q0, q1) in enumerate(own):
                    if k == bad_owner or q1 <= q0:
                        continue   # the span reduced BY the bad peer carries its own Inf; the others must be clean
                    s, e = parts[q0][0], parts[q1 - 1][1]
                    assert torch.allclose(ts[0][s:e], torch.full((e - s,), 2.0), atol=1e-5), (i, k, ts[0][s:e][:4])
    finally:
        for d in dhts:
            d.shutdown()


def test_lonely_peer_gives_up_and_keeps_local_values():
    dhts = _dhts(1)
    try:
        x = torch.arange(10.0)
        av = DecentralizedAverager([x], dhts[0], prefix="t3", min_matchmaking_time=0.5)
        t0 = time.time()
        assert av.step(timeout=1.5) is None
        assert time.time() - t0 < 6
        with av.get_tensors() as ts:
            assert torch.equal(ts[0], x)
    finally:
        dhts[0].shutdown()


def test_peers_with_another_tensor_schema_are_never_grouped():
    """Two peers average the same model; a third one under the same prefix holds other tensor sizes (a padded layout,
    another model): it is refused in matchmaking and in the join handshake, and the two keep exact results."""
    dhts = _dhts(3)
    try:
        good = [DecentralizedAverager([torch.full((300,), float(i)), torch.full((7,), float(i))], dhts[i], prefix="t6",
                                      min_matchmaking_time=1.5, target_group_size=2) for i in range(2)]
        odd = DecentralizedAverager([torch.full((304,), 9.0), torch.full((3,), 9.0)], dhts[2], prefix="t6",
                                    min_matchmaking_time=1.5, target_group_size=2)
        assert good[0].schema == good[1].schema != odd.schema
        res = _run(good + [odd], [1.0, 1.0, 1.0])
        assert res[0] is not None and res[1] is not None and res[2] is None
        for av in good:
            with av.get_tensors() as ts:
                assert torch.allclose(ts[0], torch.full((300,), 0.5)) and torch.allclose(ts[1], torch.full((7,), 0.5))
        with odd.get_tensors() as ts:
            assert torch.equal(ts[0], torch.full((304,), 9.0))
    finally:
        [d.shutdown() for d in dhts]


def test_state_download_from_highest_priority_donor():
    dhts = _dhts(3)
    try:
        avs = [DecentralizedAverager([torch.full((100,), float(i))], dhts[i], prefix="t4", declare_state_period=1.0)
               for i in range(3)]
        avs[0].state_sharing_priority, avs[1].state_sharing_priority = 1, 5
        avs[0].declare_state_now(); avs[1].declare_state_now()
        time.sleep(0.3)
        meta, tensors = avs[2].load_state_from_peers(timeout=5)
        assert torch.equal(tensors[0], torch.full((100,), 1.0))
    finally:
        for d in dhts:
            d.shutdown()
