"""C++ host library vs the PyTorch oracles: block-wise quantiser (N1), uniform 8-bit codec (H8), weighted
reduction (H7) and the bulk TCP transport (N3)."""
import threading

import pytest
import torch

from dalle_b200.native import host
from dalle_b200.optim.lamb8bit import dequantize_blockwise_ref, prepared_code, quantize_blockwise_ref

pytestmark = pytest.mark.skipif(not host.available(), reason="host library not built")


def test_blockwise_quantiser_matches_oracle():
    g = torch.Generator().manual_seed(0)
    x = torch.randn(4096 * 5 + 77, generator=g) * 0.02
    x[4096:8192] = 0.0   # an all-zero block
    for signed in (True, False):
        code = prepared_code(signed)
        xs = x if signed else x.abs()
        q, a = host.quantize_blockwise(xs, code)
        qr, ar = quantize_blockwise_ref(xs, code)
        assert torch.equal(a, ar)
        assert (q != qr).float().mean().item() < 1e-3          # ties only
        assert torch.allclose(host.dequantize_blockwise(q, a, code), dequantize_blockwise_ref(q, a, code))


def test_uniform8_matches_oracle():
    from dalle_b200.collab.compression import uniform8_quantize_ref

    x = torch.randn(100_000, generator=torch.Generator().manual_seed(1)) * 0.3 + 0.1
    q, l = host.uniform8_quantize(x)
    qr, lr = uniform8_quantize_ref(x)
    assert (q != qr).float().mean().item() < 1e-3
    assert torch.allclose(l, lr, atol=1e-4)
    acc = torch.zeros(1000)
    host.weighted_add_(acc, torch.ones(1000), 2.5)
    assert torch.allclose(acc, torch.full((1000,), 2.5))


def test_bulk_transport_roundtrip():
    from dalle_b200.transport.bulk import BulkConnection, BulkListener

    lst = BulkListener()
    data = torch.randn(3_000_000)
    got = torch.empty_like(data)
    out = {}

    def server():
        c = lst.accept(5.0)
        out["tag"] = c.recv_into(got, 10.0)
        c.send_tensor(got[:10] * 2, tag=7)
        c.close()

    th = threading.Thread(target=server)
    th.start()
    c = BulkConnection.connect(("127.0.0.1", lst.port))
    c.send_tensor(data, tag=42)
    back = torch.empty(10)
    assert c.recv_into(back, 10.0) == 7
    th.join(10)
    c.close(); lst.close()
    assert out["tag"] == 42 and torch.equal(got, data) and torch.allclose(back, data[:10] * 2)
