"""tcgen05 GEMM family vs a plain PyTorch fp32 statement of the same op."""
import pytest
import torch

from dalle_b200.ops import gemm as G

pytestmark = pytest.mark.gpu


def _rand(*shape, scale=1.0):
    return (torch.randn(*shape, device="cuda") * scale).to(torch.bfloat16)


def _check(out, ref, tol=2e-2):
    err = (out.float() - ref.float()).abs().max().item()
    den = ref.float().abs().max().item() + 1e-6
    assert err / den < tol, f"rel err {err / den:.4g}"


@pytest.mark.parametrize("M,N,K", [(128, 256, 64), (256, 512, 256), (1280, 1024, 1024), (384, 1000, 192)])
def test_nt_bf16(M, N, K):
    A, B = _rand(M, K), _rand(N, K)
    bias = torch.randn(N, device="cuda")
    ldo = (N + 63) // 64 * 64
    out = torch.zeros(M, ldo, device="cuda", dtype=torch.bfloat16)[:, :N]
    G.gemm(G.NT, G.EPI_BF16, A, B, out, bias=bias)
    ref = torch.empty(M, N)
    G.gemm_ref(G.NT, G.EPI_BF16, A.cpu(), B.cpu(), ref, bias=bias.cpu())
    _check(out.cpu(), ref)


@pytest.mark.parametrize("M,N,K", [(128, 256, 64), (1280, 1024, 4096), (256, 1024, 1000)])
def test_nn_dgrad(M, N, K):
    ldk = (K + 7) // 8 * 8
    A = _rand(M, ldk)[:, :K]
    B = _rand(K, N)
    out = torch.zeros(M, N, device="cuda", dtype=torch.bfloat16)
    G.gemm(G.NN, G.EPI_BF16, A, B, out)
    _check(out.cpu(), A.float().cpu() @ B.float().cpu())


@pytest.mark.parametrize("M,N,K", [(128, 256, 64), (1024, 1024, 2560), (3072, 1024, 1280), (1000, 1024, 640)])
def test_tn_wgrad_accumulates(M, N, K):
    A = _rand(K, (M + 7) // 8 * 8)[:, :M]
    B = _rand(K, N)
    acc0 = torch.randn(M, N, device="cuda")
    out = acc0.clone()
    G.gemm(G.TN, G.EPI_ATOMIC_F32, A, B, out, alpha=0.5)
    ref = acc0.cpu() + 0.5 * (A.float().cpu().t() @ B.float().cpu())
    _check(out.cpu(), ref, tol=5e-3)


@pytest.mark.parametrize("M,N,K", [(512, 1024, 1024), (384, 1000, 2048), (200, 64, 4096)])   # K >= 2048: deep variant
def test_resid_epilogue(M, N, K):
    A, B = _rand(M, K), _rand(N, K, scale=0.05)
    bias, scale = torch.randn(N, device="cuda"), torch.rand(N, device="cuda")
    res = torch.randn(M, N, device="cuda")
    out = torch.empty(M, N, device="cuda")
    out2 = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    G.gemm(G.NT, G.EPI_RESID, A, B, out, out2=out2, bias=bias, scale=scale, res=res, alpha=-1.0)
    ref, ref2 = torch.empty(M, N), torch.empty(M, N)
    G.gemm_ref(G.NT, G.EPI_RESID, A.cpu(), B.cpu(), ref, out2=ref2, bias=bias.cpu(), scale=scale.cpu(),
               res=res.cpu(), alpha=-1.0)
    _check(out.cpu(), ref, tol=1e-3)
    _check(out2.cpu(), ref2)


def test_geglu_fwd_bwd_epilogues():
    M, K, H = 256, 1024, 512
    A, W = _rand(M, K), _rand(2 * H, K, scale=0.03)
    bias = torch.randn(2 * H, device="cuda") * 0.1
    out = torch.empty(M, H, device="cuda", dtype=torch.bfloat16)
    pre = torch.empty(M, 2 * H, device="cuda", dtype=torch.bfloat16)
    G.gemm(G.NT, G.EPI_GEGLU, A, W, out, out2=pre, bias=bias, n_half=H)
    ref, refpre = torch.empty(M, H), torch.empty(M, 2 * H)
    G.gemm_ref(G.NT, G.EPI_GEGLU, A.cpu(), W.cpu(), ref, out2=refpre, bias=bias.cpu(), n_half=H)
    _check(out.cpu(), ref)
    _check(pre.cpu(), refpre)
    # backward epilogue: d_out [M, H] = dY[M, N2] @ W2[N2, H]
    N2 = 256
    dY, W2 = _rand(M, N2), _rand(N2, H, scale=0.05)
    dpre = torch.empty(M, 2 * H, device="cuda", dtype=torch.bfloat16)
    G.gemm(G.NN, G.EPI_GEGLU_BWD, dY, W2, dpre, aux=pre, n_half=H)
    refd = torch.empty(M, 2 * H)
    G.gemm_ref(G.NN, G.EPI_GEGLU_BWD, dY.cpu(), W2.cpu(), refd, aux=pre.cpu(), n_half=H)
    _check(dpre.cpu(), refd)


def test_rotary_epilogue():
    seq, M, K, N = 160, 320, 256, 3 * 128
    A, W = _rand(M, K), _rand(N, K, scale=0.06)
    ang = torch.rand(seq, 32, device="cuda") * 6.28
    ang[:, 30:] = 0
    rot = torch.stack((ang.cos(), ang.sin()), dim=-1).contiguous()
    out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    G.gemm(G.NT, G.EPI_ROTARY, A, W, out, rot=rot, seq_len=seq, q_cols=128, q_scale=0.125)
    ref = torch.empty(M, N)
    G.gemm_ref(G.NT, G.EPI_ROTARY, A.cpu(), W.cpu(), ref, rot=rot.cpu(), seq_len=seq, q_cols=128, q_scale=0.125)
    _check(out.cpu(), ref)


def test_wgrad_with_fused_bias_gradient():
    """TN accumulate GEMM that also returns the column sums of A (dW and db from one pass over d_pre)."""
    K, M, N = 1000, 640, 512          # K not a multiple of 64, M spans several tiles
    A, B = _rand(K, M), _rand(K, N)
    out = torch.zeros(M, N, device="cuda")
    cs = torch.full((M,), 0.5, device="cuda")
    G.gemm(G.TN, G.EPI_ATOMIC_F32, A, B, out, a_colsum=cs)
    _check(out.cpu(), A.float().cpu().t() @ B.float().cpu(), tol=5e-3)
    ref = 0.5 + A.float().sum(0).cpu()
    assert (cs.cpu() - ref).abs().max().item() < 2e-3 * ref.abs().max().item() + 1e-3


def test_resid_few_rows_in_place_split_k():
    """Decode-shaped residual GEMM (M = batch rows, in place): takes the split-K accumulate path."""
    M, N, K = 32, 1024, 4096
    A, B = _rand(M, K), _rand(N, K, scale=0.03)
    bias, scale = torch.randn(N, device="cuda"), torch.rand(N, device="cuda")
    x = torch.randn(M, N, device="cuda")
    ref = torch.empty(M, N)
    G.gemm_ref(G.NT, G.EPI_RESID, A.cpu(), B.cpu(), ref, bias=bias.cpu(), scale=scale.cpu(), res=x.cpu().clone(), alpha=1.0)
    G.gemm(G.NT, G.EPI_RESID, A, B, x, bias=bias, scale=scale, res=x, alpha=1.0)
    _check(x.cpu(), ref, tol=1e-3)


# ---- production shapes: M = 32 x 1280 rows of the benchmark step, every epilogue (reference: the same PyTorch fp32
# statement, evaluated on the GPU; TF32 stays off) -------------------------------------------------------------------
M_PROD = 40960


def _close(out, ref, tol):
    err = (out.float() - ref.float()).abs().max().item()
    den = ref.float().abs().max().item() + 1e-6
    assert err / den < tol, f"rel err {err / den:.4g}"
    # and not only where the magnitudes are large: mean error relative to mean magnitude
    assert ((out.float() - ref.float()).abs().mean() / (ref.float().abs().mean() + 1e-9)).item() < tol


@pytest.mark.parametrize("case", ["qkv_rotary", "outproj_resid_k1024", "ff2_resid_k4096", "ff1_geglu_pre", "ff2_dgrad_geglu_bwd",
                                  "ff1_dgrad_nn", "ff1_wgrad_tn_colsum", "qkv_wgrad_tn", "head_logits"])
def test_production_shapes(case):
    assert not torch.backends.cuda.matmul.allow_tf32
    M, D, Hd = M_PROD, 1024, 4096
    dev = "cuda"
    if case == "qkv_rotary":
        seq = 1280
        A, W = _rand(M, D), _rand(3 * D, D, scale=0.03)
        ang = torch.rand(seq + 1, 32, device=dev) * 6.28
        rot = torch.stack((ang.cos(), ang.sin()), dim=-1).contiguous()
        out = torch.empty(M, 3 * D, device=dev, dtype=torch.bfloat16)
        G.gemm(G.NT, G.EPI_ROTARY, A, W, out, rot=rot, seq_len=seq, q_cols=D, q_scale=0.125)
        ref = torch.empty(M, 3 * D, device=dev)
        G.gemm_ref(G.NT, G.EPI_ROTARY, A, W, ref, rot=rot, seq_len=seq, q_cols=D, q_scale=0.125)
        _close(out, ref, 2e-2)
    elif case in ("outproj_resid_k1024", "ff2_resid_k4096"):
        K = 1024 if case == "outproj_resid_k1024" else Hd
        A, W = _rand(M, K), _rand(D, K, scale=0.03)
        bias, scale = torch.randn(D, device=dev), torch.rand(D, device=dev)
        x0 = torch.randn(M, D, device=dev)
        x, o2 = x0.clone(), torch.empty(M, D, device=dev, dtype=torch.bfloat16)
        G.gemm(G.NT, G.EPI_RESID, A, W, x, out2=o2, bias=bias, scale=scale, res=x, alpha=1.0)
        ref, ref2 = torch.empty(M, D, device=dev), torch.empty(M, D, device=dev)
        G.gemm_ref(G.NT, G.EPI_RESID, A, W, ref, out2=ref2, bias=bias, scale=scale, res=x0, alpha=1.0)
        _close(x, ref, 2e-3)
        _close(o2, ref2, 2e-2)
    elif case == "ff1_geglu_pre":
        A, W = _rand(M, D), _rand(2 * Hd, D, scale=0.03)
        bias = torch.randn(2 * Hd, device=dev) * 0.1
        u, pre = torch.empty(M, Hd, device=dev, dtype=torch.bfloat16), torch.empty(M, 2 * Hd, device=dev, dtype=torch.bfloat16)
        G.gemm(G.NT, G.EPI_GEGLU, A, W, u, out2=pre, bias=bias, n_half=Hd)
        ref, refp = torch.empty(M, Hd, device=dev), torch.empty(M, 2 * Hd, device=dev)
        G.gemm_ref(G.NT, G.EPI_GEGLU, A, W, ref, out2=refp, bias=bias, n_half=Hd)
        _close(u, ref, 2e-2)
        _close(pre, refp, 2e-2)
    elif case == "ff2_dgrad_geglu_bwd":
        dY, W2 = _rand(M, D), _rand(D, Hd, scale=0.05)
        pre = _rand(M, 2 * Hd, scale=0.7)
        dpre = torch.empty(M, 2 * Hd, device=dev, dtype=torch.bfloat16)
        G.gemm(G.NN, G.EPI_GEGLU_BWD, dY, W2, dpre, aux=pre, n_half=Hd)
        ref = torch.empty(M, 2 * Hd, device=dev)
        G.gemm_ref(G.NN, G.EPI_GEGLU_BWD, dY, W2, ref, aux=pre, n_half=Hd)
        _close(dpre, ref, 2e-2)
    elif case == "ff1_dgrad_nn":
        dpre, W1 = _rand(M, 2 * Hd), _rand(2 * Hd, D, scale=0.03)
        dh = torch.empty(M, D, device=dev, dtype=torch.bfloat16)
        G.gemm(G.NN, G.EPI_BF16, dpre, W1, dh)
        _close(dh, dpre.float() @ W1.float(), 2e-2)
    elif case in ("ff1_wgrad_tn_colsum", "qkv_wgrad_tn"):
        N_out = 2 * Hd if case == "ff1_wgrad_tn_colsum" else 3 * D
        dY, h = _rand(M, N_out, scale=0.1), _rand(M, D, scale=0.1)
        g0 = torch.randn(N_out, D, device=dev)
        gW, cs = g0.clone(), torch.zeros(N_out, device=dev)
        G.gemm(G.TN, G.EPI_ATOMIC_F32, dY, h, gW, a_colsum=cs if case == "ff1_wgrad_tn_colsum" else None)
        _close(gW, g0 + dY.float().t() @ h.float(), 5e-3)
        if case == "ff1_wgrad_tn_colsum":
            _close(cs, dY.float().sum(0), 5e-3)
    else:   # head_logits: one image-segment chunk of the tied vocabulary matrix
        rows, nv = 8192, 8192
        A, W = _rand(rows, D), _rand(nv, D, scale=0.03)
        bias = torch.randn(nv, device=dev)
        lg = torch.empty(rows, nv, device=dev, dtype=torch.bfloat16)
        G.gemm(G.NT, G.EPI_BF16, A, W, lg, bias=bias)
        _close(lg, A.float() @ W.float().t() + bias, 2e-2)
