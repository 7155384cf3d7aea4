"""Per-kernel roofline table at the shapes of the headline benchmark (TTT model, micro-batch 32 -> M = 40960 rows).

Every op of a training step is timed alone (CUDA events around a CUDA graph of `iters` launches, so no host time), and
set against the MEASURED peaks of this box (MEASURED_PEAKS.json: cuBLAS bf16 TFLOP/s, copy GB/s): the time the op would
take at those peaks given its FLOPs and its minimum HBM traffic, which of the two binds, and the fraction reached.

    python tools/roofline.py [--batch 32] [--out profiles/roofline_r2.txt]
    python tools/roofline.py --once        # two launches of each op, nothing else: the program to run under ncu
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from dalle_b200.config import ttt_config
from dalle_b200.models.rotary import rotary_pair_table
from dalle_b200.ops import attention as A, elementwise as E, gemm as G

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def peaks():
    try:
        p = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(p["bf16_tflops"]), float(p["hbm_gbs"])
    except Exception:
        return 1600.0, 6500.0       # the profiling recipe's fallback


def timeit_graph(fn, iters=10):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(iters):
            fn()
    g.replay()
    torch.cuda.synchronize()
    best = 1e30
    for _ in range(3):
        s, e = torch.cuda.Event(True), torch.cuda.Event(True)
        s.record()
        g.replay()
        e.record()
        torch.cuda.synchronize()
        best = min(best, s.elapsed_time(e) / iters * 1e3)
    return best   # us


def attn_pairs(cfg, kind):
    """(query, key) pairs one (sample, head) evaluates under the layer's sparsity pattern."""
    T, F = cfg.text_len, cfg.image_fmap_size
    text = T * (T + 1) // 2
    n_img = F * F - 1
    if kind == "conv_like":
        h = cfg.conv_kernel_size // 2
        local = 0
        for p in range(n_img):
            r, c = divmod(p, F)
            for rr in range(max(0, r - h), r + 1):
                for cc in range(max(0, c - h), min(F - 1, c + h) + 1):
                    local += rr * F + cc <= p
    else:
        local = sum((p % F if kind == "axial_row" else p // F) + 1 for p in range(n_img))
    return text + n_img * T + local


def build_ops(B):
    cfg = ttt_config()
    M, D, I = B * cfg.seq_len, cfg.dim, cfg.inner_dim
    Hd = D * cfg.ff_mult
    dev = "cuda"
    bf = torch.bfloat16
    rn = lambda *s: torch.randn(*s, device=dev)
    qkv = torch.zeros(M + 128, 3 * I, device=dev, dtype=bf)
    qkv[:M] = (rn(M, 3 * I) * 0.5).to(bf)
    out = torch.zeros(M, I, device=dev, dtype=bf)
    lse = torch.zeros(B, cfg.heads, cfg.seq_len, device=dev)
    dout = torch.zeros(M + 128, I, device=dev, dtype=bf)
    dout[:M] = rn(M, I).to(bf)
    delta = torch.zeros_like(lse)
    dqkv = torch.zeros(M, 3 * I, device=dev, dtype=bf)
    kv_acc = torch.zeros(M, 2 * I, device=dev)
    rot = rotary_pair_table(cfg).cuda()
    x = rn(M, D)
    w, bb, sc = torch.ones(D, device=dev), torch.zeros(D, device=dev), torch.ones(D, device=dev)
    h = rn(M, D).to(bf)
    mean, rstd = torch.zeros(M, device=dev), torch.ones(M, device=dev)
    dx, dw, db = torch.zeros(M, D, device=dev), torch.zeros(D, device=dev), torch.zeros(D, device=dev)
    W1 = (rn(2 * Hd, D) * 0.03).to(bf)
    b1 = torch.zeros(2 * Hd, device=dev)
    u = torch.zeros(M, Hd, device=dev, dtype=bf)
    pre = (rn(M, 2 * Hd) * 0.5).to(bf)
    W2 = (rn(D, Hd) * 0.03).to(bf)
    Wq = (rn(3 * I, D) * 0.03).to(bf)
    Wo = (rn(D, I) * 0.03).to(bf)
    d_o = rn(M, D).to(bf)
    dpre = torch.zeros(M, 2 * Hd, device=dev, dtype=bf)
    dh = torch.zeros(M, D, device=dev, dtype=bf)
    gW1, db1 = torch.zeros(2 * Hd, D, device=dev), torch.zeros(2 * Hd, device=dev)
    gWo, gWq, gW2 = torch.zeros(D, D, device=dev), torch.zeros(3 * I, D, device=dev), torch.zeros(D, Hd, device=dev)
    ao = rn(M, I).to(bf)
    dao = torch.zeros(M, I, device=dev, dtype=bf)
    dqkv2 = rn(M, 3 * I).to(bf)
    dy, dsc, dbi = rn(M, D), torch.zeros(D, device=dev), torch.zeros(D, device=dev)
    # head: one image-segment chunk of the logits (rows x 8192) as the executor runs it
    rows_i, nv = 8192, cfg.num_image_tokens
    Wv = (rn(nv, D) * 0.03).to(bf)
    bv = torch.zeros(nv, device=dev)
    lg = torch.zeros(rows_i, nv, device=dev, dtype=bf)
    labels = torch.randint(0, nv, (rows_i,), device=dev)
    loss = torch.zeros(1, device=dev)
    gs = torch.ones(1, device=dev)

    def gemm(mode, epi, a, b_, o, **kw):
        return lambda: G.gemm(mode, epi, a, b_, o, **kw)

    ops = []   # (name, per_layer_count, fn, flops, min_bytes)
    add = lambda *t: ops.append(t)
    f = lambda n, k: 2.0 * M * n * k
    add("QKV GEMM + rotary (NT)", 1, gemm(G.NT, G.EPI_ROTARY, h, Wq, qkv[:M], rot=rot, seq_len=cfg.seq_len, q_cols=I, q_scale=0.125),
        f(3 * I, D), M * D * 2 + 3 * I * D * 2 + M * 3 * I * 2)
    for kind, cnt in (("axial_row", 47 / 64), ("axial_col", 16 / 64), ("conv_like", 1 / 64)):
        pairs = attn_pairs(cfg, kind) * B * cfg.heads
        add(f"attention fwd {kind}", cnt, (lambda k: lambda: A.attn_fwd(qkv, out, lse, cfg, k))(kind), 4.0 * 64 * pairs,
            M * 3 * I * 2 + M * I * 2)
        add(f"attention bwd {kind}", cnt, (lambda k: lambda: A.attn_bwd(qkv, out, dout, lse, delta, dqkv, kv_acc, rot, cfg, k))(kind),
            10.0 * 64 * pairs, M * 3 * I * 2 * 2 + M * I * 2 * 2)
    add("out-proj GEMM + residual (NT)", 1, gemm(G.NT, G.EPI_RESID, ao, Wo, x, out2=h, bias=bb, scale=sc, res=x, alpha=1.0),
        f(D, I), M * I * 2 + M * D * 4 * 2 + M * D * 2)
    add("LN + shift fwd", 2, lambda: E.ln_shift_fwd(x, None, w, bb, h, mean, rstd, cfg, True), 0.0, M * D * 4 + M * D * 2)
    add("FF1 GEMM + GEGLU (+pre) (NT)", 1, gemm(G.NT, G.EPI_GEGLU, h, W1, u, out2=pre, bias=b1, n_half=Hd),
        f(2 * Hd, D), M * D * 2 + M * Hd * 2 + M * 2 * Hd * 2)
    add("FF2 GEMM + residual (NT)", 1, gemm(G.NT, G.EPI_RESID, u, W2, x, out2=h, bias=bb, scale=sc, res=x, alpha=1.0),
        f(D, Hd), M * Hd * 2 + M * D * 4 * 2 + M * D * 2)
    add("FF2 dgrad + GEGLU backward (NN)", 1, gemm(G.NN, G.EPI_GEGLU_BWD, d_o, W2, dpre, aux=pre, n_half=Hd),
        f(Hd, D), M * D * 2 + M * 2 * Hd * 2 * 2)
    add("FF1 dgrad (NN)", 1, gemm(G.NN, G.EPI_BF16, dpre, W1, dh), f(2 * Hd, D), M * 2 * Hd * 2 + M * D * 2)
    add("FF1 wgrad + bias grad (TN)", 1, gemm(G.TN, G.EPI_ATOMIC_F32, dpre, h, gW1, a_colsum=db1), f(2 * Hd, D),
        M * 2 * Hd * 2 + M * D * 2)
    add("FF2 wgrad (TN)", 1, gemm(G.TN, G.EPI_ATOMIC_F32, d_o, u, gW2), f(Hd, D), M * D * 2 + M * Hd * 2)
    add("out-proj dgrad (NN)", 1, gemm(G.NN, G.EPI_BF16, d_o, Wo, dao), f(I, D), M * D * 2 + M * I * 2)
    add("out-proj wgrad (TN)", 1, gemm(G.TN, G.EPI_ATOMIC_F32, d_o, h, gWo), f(I, D), M * D * 2 * 2)
    add("QKV dgrad (NN)", 1, gemm(G.NN, G.EPI_BF16, dqkv2, Wq, dh), f(3 * I, D), M * 3 * I * 2 + M * D * 2)
    add("QKV wgrad (TN)", 1, gemm(G.TN, G.EPI_ATOMIC_F32, dqkv2, h, gWq), f(3 * I, D), M * 3 * I * 2 + M * D * 2)
    o_n, d_on = rn(M, D).to(bf), torch.zeros(M, D, device=dev, dtype=bf)
    add("LN + shift bwd + next block's residual prologue", 2,
        lambda: E.ln_shift_bwd(h, x, None, mean, rstd, w, dx, None, dw, db, cfg, True, 0, next_prep=(o_n, sc, d_on, dsc, dbi)),
        0.0, M * D * 2 + M * D * 4 * 2 + M * D * 4 + M * D * 2 * 2)
    add(f"head: logits GEMM ({rows_i} x {nv})", 0, gemm(G.NT, G.EPI_BF16, h[:rows_i], Wv, lg, bias=bv), 2.0 * rows_i * nv * D,
        rows_i * D * 2 + nv * D * 2 + rows_i * nv * 2)
    add(f"head: cross-entropy fwd+bwd ({rows_i} x {nv})", 0, lambda: E.ce_fwd_bwd(lg, labels, nv, 1e-4, gs, loss, write_grad=True),
        0.0, rows_i * nv * 2 * 2)
    return cfg, ops


def optimizer_op(cfg):
    """One fused LAMB-8bit step over the whole TTT parameter set (random gradients)."""
    from dalle_b200.models.dalle import DALLE
    from dalle_b200.optim.lamb8bit import LAMB8Bit
    from dalle_b200.trainer import param_groups

    model = DALLE(cfg, seed=0).cuda()
    opt = LAMB8Bit(param_groups(model, 0.045), lr=1e-3, betas=(0.9, 0.96), eps=1e-6, weight_decay=0.045,
                   max_grad_norm=4.0, clamp_value=10000.0, reuse_grad_buffers=True)
    opt._build_fused()
    flat = opt._fused.layout.flat_grad
    n = flat.numel()

    flat.normal_(0, 1e-3)

    def step():
        opt.step(zero_grad=False)

    # grad read twice (norm pass, moments pass) 8, param r/w 8, bf16 shadow 2, two 8-bit states r/w 4, update r/w in
    # the gradient buffer between the two passes 8
    return ("LAMB-8bit step (whole model, 125.9 M parameters)", 0, step, 0.0, n * 30.0), opt


def extra_ops_for_profiles(opt):
    """Kernels that are not per-layer training ops but must appear in the ncu capture: one decode step of a depth-2
    TTT-dimension model (second-generation decode kernels + sampler) and one fused group step at world 1."""
    import torch.distributed as dist

    from dalle_b200.config import get_config
    from dalle_b200.models.dalle import DALLE
    from dalle_b200.models.decode_engine import DecodeEngine

    cfg2 = get_config("ttt", depth=2, attn_types=("axial_row", "axial_col"), shared_attn_ids=(0, 1), shared_ff_ids=(0, 1))
    eng = DecodeEngine(DALLE(cfg2, seed=0).cuda(), 32)

    def decode_step():
        eng.pos.fill_(cfg2.text_len + 600)
        eng._step(True, 1.0, 0, 0.95)

    ops = [("decode step (depth 2, batch 32)", 0, decode_step, 0.0, 0.0)]
    if opt is not None:
        from dalle_b200.parallel.group import FusedGroupAverager

        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29677")
        if not dist.is_initialized():
            dist.init_process_group("nccl", rank=0, world_size=1, device_id=torch.device("cuda", 0))
        group = FusedGroupAverager(opt, payload="adaptive")
        ops.append(("fused group step (world 1, adaptive payload)", 0, lambda: group.step(1.0), 0.0, 0.0))
    return ops


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--once", action="store_true")
    ap.add_argument("--out", default=None)
    ap.add_argument("--no-optimizer", action="store_true")
    args = ap.parse_args()
    cfg, ops = build_ops(args.batch)
    opt = None
    if not args.no_optimizer:
        op, opt = optimizer_op(cfg)
        ops.append(op)
    if args.once:
        ops += extra_ops_for_profiles(opt)
        for name, _, fn, _, _ in ops:      # warm-up outside the profiled range
            fn()
        torch.cuda.synchronize()
        torch.cuda.profiler.start()        # ncu --profile-from-start off captures one launch of each op
        for name, _, fn, _, _ in ops:
            fn()
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()
        print("ok", len(ops), "ops")
        return
    tf_peak, gb_peak = peaks()
    a = torch.randn(8192, 8192, device="cuda", dtype=torch.bfloat16)
    t_box = timeit_graph(lambda: torch.matmul(a, a), iters=10)
    lines = [f"Roofline at the benchmark shapes: TTT model, micro-batch {args.batch} (M = {args.batch * cfg.seq_len} rows), one GPU.",
             f"Peaks (MEASURED_PEAKS.json): {tf_peak:.0f} TFLOP/s cuBLAS bf16 burst, {gb_peak:.0f} GB/s copy.  cuBLAS 8192^3 on this box "
             f"right now: {2 * 8192 ** 3 / t_box / 1e6:.0f} TFLOP/s.",
             "time: CUDA events around a CUDA graph of 10 launches (best of 3).  t_flop / t_mem: time at the measured peaks for the op's "
             "FLOPs / its minimum HBM traffic;", "bound: the larger of the two; reached = bound time / measured time.  x/layer: launches "
             "per layer of the 64-layer model (attention types by their share of the layers).", "",
             f"{'op':52s} {'us':>8s} {'TFLOP/s':>8s} {'GB/s':>7s} {'t_flop':>7s} {'t_mem':>7s} {'bound':>6s} {'reached':>8s} {'x/layer':>7s}"]
    total_layer, ideal_layer, rows = 0.0, 0.0, []
    for name, cnt, fn, flops, nbytes in ops:
        us = timeit_graph(fn)
        t_flop, t_mem = flops / tf_peak / 1e6, nbytes / gb_peak / 1e3
        bound = max(t_flop, t_mem)
        lines.append(f"{name:52s} {us:8.1f} {flops / us / 1e6:8.0f} {nbytes / us / 1e3:7.0f} {t_flop:7.1f} {t_mem:7.1f} "
                     f"{'flop' if t_flop >= t_mem else 'mem':>6s} {100 * bound / us:7.0f}% {cnt:7.2f}")
        total_layer += cnt * us
        ideal_layer += cnt * bound
        rows.append({"op": name, "us": us, "flops": flops, "bytes": nbytes, "bound_us": bound, "per_layer": cnt})
    lines += ["", f"sum over one layer: {total_layer:.0f} us measured, {ideal_layer:.0f} us at the measured peaks "
                  f"({100 * ideal_layer / total_layer:.0f}%); x 64 layers = {total_layer * 64 / 1e3:.1f} ms of the step"]
    text = "\n".join(lines)
    print(text)
    if args.out:
        with open(args.out, "w") as fo:
            fo.write(text + "\n")
        with open(os.path.splitext(args.out)[0] + ".json", "w") as fo:
            json.dump(rows, fo)


if __name__ == "__main__":
    main()
