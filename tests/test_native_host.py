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


import pytest  # noqa: E402


@pytest.mark.parametrize("sanitizer", ["thread", "address,undefined"])
def test_cpp_data_plane_under_sanitizers(tmp_path, sanitizer):
    """the threaded C++ all-reduce (native/allreduce_pipe.cpp + transport.cpp) built with ThreadSanitizer
    and with Address/UB sanitizers, three and four members (node, node, auxiliary, client) over loopback, all codecs;
    a sanitizer report or a wrong average fails the run.  (OpenMP is left out of this build: libgomp is not
    instrumented and would drown the report in false positives; the pragmas are ignored.)"""
    import os
    import shutil
    import subprocess

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    if shutil.which("g++") is None:
        pytest.skip("no g++")
    exe = str(tmp_path / "ar_san")
    srcs = [os.path.join(root, "tests", "native", "allreduce_tsan.cpp")] + [
        os.path.join(root, "dalle_b200", "native", f) for f in ("allreduce_pipe.cpp", "transport.cpp", "quant_host.cpp")]
    build = subprocess.run(["g++", "-O1", "-g", "-std=c++17", "-march=x86-64-v3", f"-fsanitize={sanitizer}",
                            "-fno-omit-frame-pointer", "-I", os.path.join(root, "include"), *srcs, "-lpthread", "-o", exe],
                           capture_output=True, text=True)
    if build.returncode != 0 and "sanitizer" in build.stderr.lower() or "cannot find -l" in build.stderr:
        pytest.skip("sanitizer runtime not installed")
    assert build.returncode == 0, build.stderr[-2000:]
    env = dict(os.environ, TSAN_OPTIONS="halt_on_error=1", ASAN_OPTIONS="detect_leaks=1", UBSAN_OPTIONS="halt_on_error=1")
    run = subprocess.run([exe], capture_output=True, text=True, env=env, timeout=300)
    assert run.returncode == 0, (run.stdout + run.stderr)[-4000:]
    assert "0 + 0 problems" in run.stdout and "WARNING: ThreadSanitizer" not in run.stderr and "ERROR: AddressSanitizer" not in run.stderr
