"""world = 1: ``reduce_gradients`` of the fused kernel returns decode(encode(g)) - compare its hivemind codecs with the
oracle in collab/compression.py chunk by chunk.  Prints CODECS_OK."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import torch

from dalle_b200.collab.compression import uniform8_quantize_ref
from dalle_b200.parallel.group import FusedGroupAverager
from check_group import init_single_rank, make


def main():
    init_single_rank()
    dev = torch.device("cuda", 0)
    for payload in ("fp16", "u8h"):
        ps, opt = make(dev)
        group = FusedGroupAverager(opt, payload=payload)
        g = torch.Generator(device=dev).manual_seed(5)
        L = opt._fused.layout
        for p in ps:
            p.grad.copy_(torch.randn(p.shape, generator=g, device=dev) * 0.05 + 0.01)
        src = L.flat_grad.clone()
        group.reduce_gradients(grad_scale=1.0)
        torch.cuda.synchronize()
        got = L.flat_grad.cpu()
        src = src.cpu()
        if payload == "fp16":
            want = src.clamp(-65504, 65504).half().float()
            assert torch.equal(got, want), (got - want).abs().max()
        else:
            chunk = 32 * 4096
            worst = 0.0
            for c0 in range(0, src.numel(), chunk):
                x = src[c0:c0 + chunk]
                q, lookup = uniform8_quantize_ref(x)
                want = lookup[q.long()]
                d = (got[c0:c0 + chunk] - want).abs()
                bin_w = 6.0 * (x - x.mean()).norm().item() / (x.numel() - 1) ** 0.5 / 256
                # identical up to: a value on a bin edge that fp32 rounding of mu / sigma puts into the neighbouring bin
                # (that element moves by about one bin width, and the MEANS of the two bins it left / joined - hence
                # every element of those bins - by about a 500th of it)
                frac_diff = (d > 0.05 * bin_w).float().mean().item()
                assert frac_diff < 1e-3, (c0, frac_diff)
                assert d.max().item() <= 1.5 * bin_w + 1e-6, (c0, d.max().item(), bin_w)
                worst = max(worst, frac_diff)
            print(f"u8h: worst fraction of elements in a neighbouring bin {worst:.2e}")
    print("CODECS_OK")
    torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
