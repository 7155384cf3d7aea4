import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from dalle_b200.ops import gemm as G
def timeit(fn, iters=20):
    for _ in range(5): fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(True), torch.cuda.Event(True)
    s.record()
    for _ in range(iters): fn()
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / iters * 1e3
M, D = 40960, 1024
dev = "cuda"
h = torch.randn(M, D, device=dev).bfloat16()
W1 = (torch.randn(8192, D, device=dev) * 0.03).bfloat16(); b1 = torch.zeros(8192, device=dev)
u = torch.zeros(M, 4096, device=dev, dtype=torch.bfloat16); pre = torch.zeros(M, 8192, device=dev, dtype=torch.bfloat16)
W2 = (torch.randn(D, 4096, device=dev) * 0.03).bfloat16()
d_o = torch.randn(M, D, device=dev).bfloat16(); dpre = torch.zeros(M, 8192, device=dev, dtype=torch.bfloat16)
for rep in range(2):
    a = timeit(lambda: G.gemm(G.NT, G.EPI_GEGLU, h, W1, u, bias=b1, n_half=4096))
    b = timeit(lambda: G.gemm(G.NT, G.EPI_GEGLU, h, W1, u, out2=pre, bias=b1, n_half=4096))
    c = timeit(lambda: G.gemm(G.NN, G.EPI_GEGLU_BWD, d_o, W2, dpre, aux=pre, n_half=4096))
    d = timeit(lambda: G.gemm(G.NT, G.EPI_BF16, h, W1, pre))
    print(f"GEGLU {a:.0f}  GEGLU+pre {b:.0f}  GEGLU_BWD {c:.0f}  plainNT(8192) {d:.0f} us", flush=True)
