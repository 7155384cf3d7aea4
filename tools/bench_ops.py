"""GPU micro-benchmarks of individual ops at TTT shapes (CUDA events, warm)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from dalle_b200.config import ttt_config
from dalle_b200.ops import attention as A, elementwise as E, gemm as G
from dalle_b200.models.rotary import rotary_pair_table


def timeit(fn, iters=30):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(True), torch.cuda.Event(True)
    s.record()
    for _ in range(iters): fn()
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / iters * 1e3  # us


def timeit_graph(fn, iters=20):
    """Device time per call: the calls are captured into one CUDA graph (no host launch cost between kernels)."""
    for _ in range(3): fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(iters): fn()
    g.replay(); torch.cuda.synchronize()
    s, e = torch.cuda.Event(True), torch.cuda.Event(True)
    s.record(); g.replay(); e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / iters * 1e3


def box_speed():
    """cuBLAS bf16 8192^3 matmul TF/s: a yardstick for the box's current power/clock state."""
    a = torch.randn(8192, 8192, device='cuda', dtype=torch.bfloat16); b = torch.randn(8192, 8192, device='cuda', dtype=torch.bfloat16)
    t = timeit(lambda: torch.matmul(a, b), iters=20)
    return 2 * 8192 ** 3 / t / 1e6


def main():
    print(f"box speed (cuBLAS 8192^3): {box_speed():.0f} TF/s", flush=True)
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
    cfg = ttt_config()
    M, D, I = B * cfg.seq_len, cfg.dim, cfg.inner_dim
    dev = "cuda"
    qkv = torch.zeros(M + 128, 3 * I, device=dev, dtype=torch.bfloat16)
    qkv[:M] = (torch.randn(M, 3 * I, device=dev) * 0.5).bfloat16()
    out = torch.zeros(M, I, device=dev, dtype=torch.bfloat16)
    lse = torch.zeros(B, cfg.heads, cfg.seq_len, device=dev)
    dout = torch.zeros(M + 128, I, device=dev, dtype=torch.bfloat16)
    dout[:M] = torch.randn(M, I, device=dev).bfloat16()
    delta = torch.zeros_like(lse)
    dqkv = torch.zeros(M, 3 * I, device=dev, dtype=torch.bfloat16)
    kv_acc = torch.zeros(M, 2 * I, device=dev)
    rot = rotary_pair_table(cfg).cuda()
    for t in ("axial_row", "axial_col", "conv_like"):
        f = timeit(lambda: A.attn_fwd(qkv, out, lse, cfg, t))
        b = timeit(lambda: A.attn_bwd(qkv, out, dout, lse, delta, dqkv, kv_acc, rot, cfg, t))
        print(f"attn {t}: fwd {f:.0f} us  bwd {b:.0f} us", flush=True)
    x = torch.randn(M, D, device=dev); w = torch.ones(D, device=dev); bb = torch.zeros(D, device=dev)
    h = torch.zeros(M, D, device=dev, dtype=torch.bfloat16); mean = torch.zeros(M, device=dev); rstd = torch.zeros(M, device=dev)
    print(f"ln_shift_fwd {timeit(lambda: E.ln_shift_fwd(x, None, w, bb, h, mean, rstd, cfg, True)):.0f} us  (graph: {timeit_graph(lambda: E.ln_shift_fwd(x, None, w, bb, h, mean, rstd, cfg, True)):.0f} us)")
    dx = torch.zeros(M, D, device=dev); dw = torch.zeros(D, device=dev); db = torch.zeros(D, device=dev)
    print(f"  (yardstick: torch fp32->bf16 cast of the same tensor, same bytes: {timeit_graph(lambda: h.copy_(x)):.0f} us; "
          f"fp32 copy read+write 336 MB: {timeit_graph(lambda: dx.copy_(x)):.0f} us)")
    print(f"ln_shift_bwd {timeit(lambda: E.ln_shift_bwd(h, x, None, mean, rstd, w, dx, None, dw, db, cfg, True, 0)):.0f} us")
    o_n = torch.randn(M, D, device=dev).bfloat16(); d_on = torch.zeros(M, D, device=dev, dtype=torch.bfloat16)
    dsn = torch.zeros(D, device=dev); dbn = torch.zeros(D, device=dev)
    print(f"ln_shift_bwd+prep {timeit(lambda: E.ln_shift_bwd(h, x, None, mean, rstd, w, dx, None, dw, db, cfg, True, 0, next_prep=(o_n, w, d_on, dsn, dbn))):.0f} us")
    # GEMM epilogue variants
    W1 = (torch.randn(8192, D, device=dev) * 0.03).bfloat16(); b1 = torch.zeros(8192, device=dev)
    u = torch.zeros(M, 4096, device=dev, dtype=torch.bfloat16); pre = torch.zeros(M, 8192, device=dev, dtype=torch.bfloat16)
    t_ = timeit(lambda: G.gemm(G.NT, G.EPI_GEGLU, h, W1, u, out2=pre, bias=b1, n_half=4096))
    print(f"FF1 GEGLU (+pre) {t_:.0f} us  {2*M*8192*D/t_/1e6:.0f} TF/s")
    t_ = timeit(lambda: G.gemm(G.NT, G.EPI_GEGLU, h, W1, u, bias=b1, n_half=4096))
    print(f"FF1 GEGLU        {t_:.0f} us  {2*M*8192*D/t_/1e6:.0f} TF/s")
    W2 = (torch.randn(D, 4096, device=dev) * 0.03).bfloat16(); sc = torch.ones(D, device=dev)
    t_ = timeit(lambda: G.gemm(G.NT, G.EPI_RESID, u, W2, x, out2=h, bias=bb, scale=sc, res=x, alpha=1.0))
    print(f"FF2 RESID        {t_:.0f} us  {2*M*4096*D/t_/1e6:.0f} TF/s")
    Wq = (torch.randn(3 * I, D, device=dev) * 0.03).bfloat16()
    t_ = timeit(lambda: G.gemm(G.NT, G.EPI_ROTARY, h, Wq, qkv[:M], rot=rot, seq_len=cfg.seq_len, q_cols=I, q_scale=0.125))
    print(f"QKV ROTARY       {t_:.0f} us  {2*M*3*I*D/t_/1e6:.0f} TF/s")
    d_o = torch.randn(M, D, device=dev).bfloat16(); dpre = torch.zeros(M, 8192, device=dev, dtype=torch.bfloat16)
    t_ = timeit(lambda: G.gemm(G.NN, G.EPI_GEGLU_BWD, d_o, W2, dpre, aux=pre, n_half=4096))
    print(f"GEGLU_BWD dgrad  {t_:.0f} us  {2*M*4096*D/t_/1e6:.0f} TF/s")
    dh = torch.zeros(M, D, device=dev, dtype=torch.bfloat16)
    t_ = timeit(lambda: G.gemm(G.NN, G.EPI_BF16, dpre, W1, dh))
    print(f"dgrad FF1        {t_:.0f} us  {2*M*8192*D/t_/1e6:.0f} TF/s")
    gW1 = torch.zeros(8192, D, device=dev)
    t_ = timeit(lambda: G.gemm(G.TN, G.EPI_ATOMIC_F32, dpre, h, gW1))
    print(f"wgrad FF1        {t_:.0f} us  {2*M*8192*D/t_/1e6:.0f} TF/s")
    db1 = torch.zeros(8192, device=dev)
    t_ = timeit(lambda: G.gemm(G.TN, G.EPI_ATOMIC_F32, dpre, h, gW1, a_colsum=db1))
    print(f"wgrad FF1 + db   {t_:.0f} us  {2*M*8192*D/t_/1e6:.0f} TF/s")
    gWo = torch.zeros(D, D, device=dev)
    t_ = timeit(lambda: G.gemm(G.TN, G.EPI_ATOMIC_F32, d_o, h, gWo))
    print(f"wgrad 1024x1024  {t_:.0f} us  {2*M*D*D/t_/1e6:.0f} TF/s")
    # the remaining per-layer shapes
    Wo = (torch.randn(D, I, device=dev) * 0.03).bfloat16(); ao = torch.randn(M, I, device=dev).bfloat16()
    t_ = timeit(lambda: G.gemm(G.NT, G.EPI_RESID, ao, Wo, x, out2=h, bias=bb, scale=sc, res=x, alpha=1.0))
    print(f"out-proj RESID   {t_:.0f} us  {2*M*I*D/t_/1e6:.0f} TF/s")
    dao = torch.zeros(M, I, device=dev, dtype=torch.bfloat16)
    t_ = timeit(lambda: G.gemm(G.NN, G.EPI_BF16, d_o, Wo, dao))
    print(f"dgrad out-proj   {t_:.0f} us  {2*M*I*D/t_/1e6:.0f} TF/s")
    dqkv2 = torch.randn(M, 3 * I, device=dev).bfloat16()
    t_ = timeit(lambda: G.gemm(G.NN, G.EPI_BF16, dqkv2, Wq, dh))
    print(f"dgrad QKV        {t_:.0f} us  {2*M*3*I*D/t_/1e6:.0f} TF/s")
    gWq = torch.zeros(3 * I, D, device=dev)
    t_ = timeit(lambda: G.gemm(G.TN, G.EPI_ATOMIC_F32, dqkv2, h, gWq))
    print(f"wgrad QKV        {t_:.0f} us  {2*M*3*I*D/t_/1e6:.0f} TF/s")
    gW2 = torch.zeros(D, 4096, device=dev)
    t_ = timeit(lambda: G.gemm(G.TN, G.EPI_ATOMIC_F32, d_o, u, gW2))
    print(f"wgrad FF2        {t_:.0f} us  {2*M*4096*D/t_/1e6:.0f} TF/s")
    dy = torch.randn(M, D, device=dev); dsc = torch.zeros(D, device=dev); dbi = torch.zeros(D, device=dev)
    t_ = timeit(lambda: E.resid_bwd_prep(dy, ao, sc, dao, dsc, dbi))
    print(f"resid_bwd_prep   {t_:.0f} us")
    cs = torch.zeros(8192, device=dev)
    t_ = timeit(lambda: E.colsum(dpre, cs, 1.0))
    print(f"colsum [M,8192]  {t_:.0f} us")


    print(f"box speed (cuBLAS 8192^3): {box_speed():.0f} TF/s", flush=True)


if __name__ == "__main__":
    main()