"""Short program for `ncu --set full`: one launch of each hot kernel at TTT shapes (B=16)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from dalle_b200.config import ttt_config
from dalle_b200.ops import attention as A, elementwise as E, gemm as G
from dalle_b200.models.rotary import rotary_pair_table

B = 16
cfg = ttt_config()
M, D, I = B * cfg.seq_len, cfg.dim, cfg.inner_dim
dev = "cuda"
h = (torch.randn(M, D, device=dev)).bfloat16()
W1 = (torch.randn(8192, D, device=dev) * 0.03).bfloat16(); b1 = torch.zeros(8192, device=dev)
u = torch.zeros(M, 4096, device=dev, dtype=torch.bfloat16); pre = torch.zeros(M, 8192, device=dev, dtype=torch.bfloat16)
x = torch.randn(M, D, device=dev); sc = torch.ones(D, device=dev); bb = torch.zeros(D, device=dev)
W2 = (torch.randn(D, 4096, device=dev) * 0.03).bfloat16()
rot = rotary_pair_table(cfg).cuda()
qkv = torch.zeros(M + 128, 3 * I, device=dev, dtype=torch.bfloat16); qkv[:M] = (torch.randn(M, 3 * I, device=dev) * 0.5).bfloat16()
out = torch.zeros(M, I, device=dev, dtype=torch.bfloat16); lse = torch.zeros(B, cfg.heads, cfg.seq_len, device=dev)
dout = torch.zeros(M + 128, I, device=dev, dtype=torch.bfloat16); dout[:M] = torch.randn(M, I, device=dev).bfloat16()
delta = torch.zeros_like(lse); dqkv = torch.zeros(M, 3 * I, device=dev, dtype=torch.bfloat16); kv_acc = torch.zeros(M, 2 * I, device=dev)
mean = torch.zeros(M, device=dev); rstd = torch.ones(M, device=dev); dx = torch.zeros(M, D, device=dev)
dw = torch.zeros(D, device=dev); db = torch.zeros(D, device=dev); gW = torch.zeros(8192, D, device=dev)
cs8k = torch.zeros(8192, device=dev)
for _ in range(2):
    G.gemm(G.NT, G.EPI_BF16, h, W1, pre)                                        # plain NT
    G.gemm(G.NT, G.EPI_GEGLU, h, W1, u, out2=pre, bias=b1, n_half=4096)         # GEGLU + pre
    G.gemm(G.NT, G.EPI_RESID, u, W2, x, bias=bb, scale=sc, res=x, alpha=1.0)    # residual epilogue
    G.gemm(G.NN, G.EPI_GEGLU_BWD, h, W2, pre, aux=pre, n_half=4096)             # GEGLU backward dgrad
    G.gemm(G.TN, G.EPI_ATOMIC_F32, pre, h, gW)                                  # wgrad split-K
    A.attn_fwd(qkv, out, lse, cfg, "axial_row")
    A.attn_bwd(qkv, out, dout, lse, delta, dqkv, kv_acc, rot, cfg, "axial_row")
    E.ln_shift_fwd(x, None, sc, bb, h, mean, rstd, cfg, True)
    E.ln_shift_bwd(h, x, None, mean, rstd, sc, dx, None, dw, db, cfg, True, 0)
    E.colsum(pre, cs8k, 1.0)
    E.resid_bwd_prep(dx, h, sc, out, dw, db)
torch.cuda.synchronize()
print("ok")
