"""LayerNorm+shift backward (plain and with the fused residual prologue of the next block) at benchmark shapes,
timed inside a CUDA graph of 10 launches.  DALLE_B200_LNB_RING=n sets the rows in flight per row group."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from dalle_b200.config import ttt_config
from dalle_b200.ops import elementwise as E
from tools.roofline import timeit_graph

cfg = ttt_config(); B = 32; M, D = B * cfg.seq_len, cfg.dim; dev = "cuda"
x = torch.randn(M, D, device=dev); w = torch.ones(D, device=dev)
h = torch.randn(M, D, device=dev).bfloat16(); mean = torch.zeros(M, device=dev); rstd = torch.ones(M, device=dev)
dx = torch.zeros(M, D, device=dev); dw = torch.zeros(D, device=dev); db = torch.zeros(D, device=dev)
o_n = torch.randn(M, D, device=dev).bfloat16(); d_on = torch.zeros(M, D, device=dev, dtype=torch.bfloat16)
dsn = torch.zeros(D, device=dev); dbn = torch.zeros(D, device=dev)
t0 = timeit_graph(lambda: E.ln_shift_bwd(h, x, None, mean, rstd, w, dx, None, dw, db, cfg, True, 0))
t1 = timeit_graph(lambda: E.ln_shift_bwd(h, x, None, mean, rstd, w, dx, None, dw, db, cfg, True, 0, next_prep=(o_n, w, d_on, dsn, dbn)))
print(f"ring={os.environ.get('DALLE_B200_LNB_RING', 'default')}: ln_shift_bwd {t0:.1f} us, with next-block prologue {t1:.1f} us "
      f"({M * D * 14 / t0 / 1e3:.0f} / {M * D * 18 / t1 / 1e3:.0f} GB/s)")
