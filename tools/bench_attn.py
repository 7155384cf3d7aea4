"""Attention forward/backward device times at TTT shapes (CUDA events after warm-up; inputs 250 MB+ per call, larger
than L2).  usage: bench_attn.py [B] [kinds]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from dalle_b200.config import ttt_config
from dalle_b200.ops import attention as A
from dalle_b200.models.rotary import rotary_pair_table
from tools.bench_ops import timeit

B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
kinds = sys.argv[2].split(",") if len(sys.argv) > 2 else ["axial_row", "axial_col", "conv_like"]
cfg = ttt_config()
M, I = B * cfg.seq_len, cfg.inner_dim
dev = "cuda"
rot = rotary_pair_table(cfg).cuda()
qkv = torch.zeros(M + 128, 3 * I, device=dev, dtype=torch.bfloat16); qkv[:M] = (torch.randn(M, 3 * I, device=dev) * 0.5).bfloat16()
out = torch.zeros(M, I, device=dev, dtype=torch.bfloat16); lse = torch.zeros(B, cfg.heads, cfg.seq_len, device=dev)
dout = torch.zeros(M + 128, I, device=dev, dtype=torch.bfloat16); dout[:M] = torch.randn(M, I, device=dev).bfloat16()
delta = torch.zeros_like(lse); dqkv = torch.zeros(M, 3 * I, device=dev, dtype=torch.bfloat16); kv_acc = torch.zeros(M, 2 * I, device=dev)
for t in kinds:
    f = timeit(lambda: A.attn_fwd(qkv, out, lse, cfg, t))
    b = timeit(lambda: A.attn_bwd(qkv, out, dout, lse, delta, dqkv, kv_acc, rot, cfg, t))
    print(f"B={B} attn {t}: fwd {f:.0f} us  bwd {b:.0f} us", flush=True)
