"""Short program for `ncu --set full -k regex:attn`: attention forward + backward only (TTT shapes, B=16)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from dalle_b200.config import ttt_config
from dalle_b200.ops import attention as A
from dalle_b200.models.rotary import rotary_pair_table

B = int(sys.argv[1]) if len(sys.argv) > 1 else 16
kinds = sys.argv[2].split(",") if len(sys.argv) > 2 else ["axial_row"]
cfg = ttt_config()
M, I = B * cfg.seq_len, cfg.inner_dim
dev = "cuda"
rot = rotary_pair_table(cfg).cuda()
qkv = torch.zeros(M + 128, 3 * I, device=dev, dtype=torch.bfloat16); qkv[:M] = (torch.randn(M, 3 * I, device=dev) * 0.5).bfloat16()
out = torch.zeros(M, I, device=dev, dtype=torch.bfloat16); lse = torch.zeros(B, cfg.heads, cfg.seq_len, device=dev)
dout = torch.zeros(M + 128, I, device=dev, dtype=torch.bfloat16); dout[:M] = torch.randn(M, I, device=dev).bfloat16()
delta = torch.zeros_like(lse); dqkv = torch.zeros(M, 3 * I, device=dev, dtype=torch.bfloat16); kv_acc = torch.zeros(M, 2 * I, device=dev)
for _ in range(2):
    for kind in kinds:
        A.attn_fwd(qkv, out, lse, cfg, kind)
        A.attn_bwd(qkv, out, dout, lse, delta, dqkv, kv_acc, rot, cfg, kind)
torch.cuda.synchronize()
print("ok")
