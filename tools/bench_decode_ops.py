"""Per-kernel device times of the second-generation decode step at TTT dimensions, batch 32: each kernel kind is
launched once per layer (64 different caches / 5 weight blocks in turn) inside a CUDA graph between two CUDA events;
argv[2] = 0 | 1 switches programmatic dependent launch.  Last line: the graph of all 64 real layers."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from dalle_b200.config import ttt_config
from dalle_b200.models.dalle import DALLE
from dalle_b200.models.decode_engine import DecodeEngine
from dalle_b200.models.masks import ATTN_TYPE_CODES
from dalle_b200.ops import lib

B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
pdl = int(sys.argv[2]) if len(sys.argv) > 2 else 0
cfg = ttt_config()
model = DALLE(cfg, seed=0).cuda()
eng = DecodeEngine(model, B)
L = lib.cuda_lib()
L.b200_decode2_set_pdl(pdl)
m = model
D, I, hid = cfg.dim, cfg.inner_dim, cfg.dim * cfg.ff_mult
eng.pos.fill_(cfg.text_len + 600)


def qkv(i):
    a = m.attn_index[i]
    eng._gemm(0, eng._w(m.attn_qkv_w[a]), 3 * I, D, x=eng.x2, site=2 * i, ln=(m.norm_f_w[i], m.norm_f_b[i]),
              ring=eng.ring[i][0], out=eng.qkv, ldo=3 * I, rot=m.rotary_pairs, q_cols=I, q_scale=0.125)


def attn(i):
    lib.check(L.b200_decode2_attn(eng.qkv.data_ptr(), eng.kc[i].data_ptr(), eng.vc[i].data_ptr(), eng.ao.data_ptr(),
                                  eng.pos.data_ptr(), B, cfg.heads, cfg.text_len, cfg.image_fmap_size,
                                  ATTN_TYPE_CODES[cfg.attn_types[i]] + 1, cfg.conv_kernel_size, eng.slots[i],
                                  torch.cuda.current_stream().cuda_stream), "attn")


def outp(i):
    a = m.attn_index[i]
    eng._gemm(1, eng._w(m.attn_out_w[a]), D, I, xin=eng.ao, bias=m.attn_out_b[a], scale=m.scale_f[i].view(-1),
              xout=eng.x1, site_out=2 * i + 1)


def ff1(i):
    f = m.ff_index[i]
    eng._gemm(2, eng._w(m.ff_w1[f]), hid, D, x=eng.x1, site=2 * i + 1, ln=(m.norm_g_w[i], m.norm_g_b[i]),
              ring=eng.ring[i][1], bias=m.ff_b1[f], out=eng.u, ldo=hid)


def ff2(i):
    f = m.ff_index[i]
    eng._gemm(1, eng._w(m.ff_w2[f]), D, hid, xin=eng.u, bias=m.ff_b2[f], scale=m.scale_g[i].view(-1), xout=eng.x2,
              site_out=2 * i + 2)


def timed(fn, name):
    for i in range(cfg.depth):
        fn(i)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):            # one launch per layer inside a graph: no host time in the measurement
        for i in range(cfg.depth):
            fn(i)
    g.replay()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(3):
        s, e = torch.cuda.Event(True), torch.cuda.Event(True)
        s.record()
        g.replay()
        e.record()
        torch.cuda.synchronize()
        best = min(best, s.elapsed_time(e) / cfg.depth * 1e3)
    print(f"{name:10s} {best:7.2f} us per launch (graph of {cfg.depth} launches of this kind, pdl={pdl})")
    return best


tot = sum(timed(f, n) for f, n in ((qkv, "qkv"), (attn, "attn"), (outp, "out-proj"), (ff1, "ff1"), (ff2, "ff2")))
print(f"sum per layer {tot:.1f} us -> {tot * cfg.depth / 1e3:.2f} ms per step for the layers")
g = torch.cuda.CUDAGraph()
eng.pos.fill_(cfg.text_len + 600)
with torch.cuda.graph(g):
    eng._layers()
g.replay()
torch.cuda.synchronize()
s, e = torch.cuda.Event(True), torch.cuda.Event(True)
s.record()
for _ in range(5):
    g.replay()
e.record()
torch.cuda.synchronize()
print(f"graph of the {cfg.depth} layers: {s.elapsed_time(e) / 5:.3f} ms per replay (pdl={pdl})")
