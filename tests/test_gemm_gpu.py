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
