"""Per-kernel device times of attention forward/backward via torch.profiler (CUPTI): which kernels ran, how long.
usage: kernel_times.py [B] [kind]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import profile, ProfilerActivity
from dalle_b200.config import ttt_config
from dalle_b200.ops import attention as A
from dalle_b200.models.rotary import rotary_pair_table

B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
kind = sys.argv[2] if len(sys.argv) > 2 else "axial_row"
cfg = ttt_config()
M, I = B * cfg.seq_len, cfg.inner_dim
dev = "cuda"
rot = rotary_pair_table(cfg).cuda()
qkv = torch.zeros(M + 128, 3 * I, device=dev, dtype=torch.bfloat16); qkv[:M] = (torch.randn(M, 3 * I, device=dev) * 0.5).bfloat16()
out = torch.zeros(M, I, device=dev, dtype=torch.bfloat16); lse = torch.zeros(B, cfg.heads, cfg.seq_len, device=dev)
dout = torch.zeros(M + 128, I, device=dev, dtype=torch.bfloat16); dout[:M] = torch.randn(M, I, device=dev).bfloat16()
delta = torch.zeros_like(lse); dqkv = torch.zeros(M, 3 * I, device=dev, dtype=torch.bfloat16); kv_acc = torch.zeros(M, 2 * I, device=dev)
def step():
    A.attn_fwd(qkv, out, lse, cfg, kind)
    A.attn_bwd(qkv, out, dout, lse, delta, dqkv, kv_acc, rot, cfg, kind)
for _ in range(3): step()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(5): step()
    torch.cuda.synchronize()
for e in sorted(prof.key_averages(), key=lambda e: -e.device_time_total):
    if e.device_time_total > 0:
        print(f"{e.device_time_total / e.count:9.1f} us x{e.count:3d}  {e.key[:110]}")
