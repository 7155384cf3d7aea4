"""GPU diagnostic: numerics of the three GEMM modes + throughput vs torch.matmul (cuBLAS)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from dalle_b200.ops import gemm as G


def relerr(a, b):
    return ((a.float() - b.float()).abs().max() / (b.float().abs().max() + 1e-6)).item()


def main():
    torch.manual_seed(0)
    dev = "cuda"
    for (M, N, K) in [(128, 256, 64), (128, 256, 256), (256, 512, 128), (1280, 1024, 1024)]:
        A = torch.randn(M, K, device=dev).bfloat16(); B = torch.randn(N, K, device=dev).bfloat16()
        out = torch.zeros(M, N, device=dev, dtype=torch.bfloat16)
        G.gemm(G.NT, G.EPI_BF16, A, B, out); torch.cuda.synchronize()
        print("NT", M, N, K, "relerr", relerr(out, A.float() @ B.float().t()), flush=True)
        Bn = torch.randn(K, N, device=dev).bfloat16()
        G.gemm(G.NN, G.EPI_BF16, A, Bn, out); torch.cuda.synchronize()
        print("NN", M, N, K, "relerr", relerr(out, A.float() @ Bn.float()), flush=True)
        At = torch.randn(K, M, device=dev).bfloat16()
        o32 = torch.zeros(M, N, device=dev)
        G.gemm(G.TN, G.EPI_ATOMIC_F32, At, Bn, o32); torch.cuda.synchronize()
        print("TN", M, N, K, "relerr", relerr(o32, At.float().t() @ Bn.float()), flush=True)

    def bench(fn, flops, iters=20):
        for _ in range(3): fn()
        torch.cuda.synchronize()
        s, e = torch.cuda.Event(True), torch.cuda.Event(True)
        s.record()
        for _ in range(iters): fn()
        e.record(); torch.cuda.synchronize()
        ms = s.elapsed_time(e) / iters
        return ms, flops / ms / 1e9

    for (M, N, K) in [(40960, 1024, 1024), (40960, 3072, 1024), (40960, 1024, 4096), (40960, 8192, 1024), (8192, 8192, 8192)]:
        A = torch.randn(M, K, device=dev).bfloat16(); B = torch.randn(N, K, device=dev).bfloat16()
        out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
        fl = 2.0 * M * N * K
        ms, tf = bench(lambda: G.gemm(G.NT, G.EPI_BF16, A, B, out), fl)
        ms2, tf2 = bench(lambda: torch.matmul(A, B.t(), out=out), fl)
        print(f"NT {M}x{N}x{K}: ours {ms:.3f} ms {tf:.0f} TF/s | cublas {ms2:.3f} ms {tf2:.0f} TF/s", flush=True)
    # dgrad / wgrad shapes
    M, N, K = 40960, 1024, 4096
    dY = torch.randn(M, K, device=dev).bfloat16(); W = torch.randn(K, N, device=dev).bfloat16()
    out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    ms, tf = bench(lambda: G.gemm(G.NN, G.EPI_BF16, dY, W, out), 2.0 * M * N * K)
    print(f"NN {M}x{N}x{K}: ours {ms:.3f} ms {tf:.0f} TF/s", flush=True)
    X = torch.randn(M, N, device=dev).bfloat16()
    dW = torch.zeros(K, N, device=dev)
    ms, tf = bench(lambda: G.gemm(G.TN, G.EPI_ATOMIC_F32, dY, X, dW), 2.0 * M * N * K)
    print(f"TN wgrad {K}x{N} over {M}: ours {ms:.3f} ms {tf:.0f} TF/s", flush=True)
    dY2 = torch.randn(M, 1024, device=dev).bfloat16()
    dW2 = torch.zeros(1024, 1024, device=dev)
    ms, tf = bench(lambda: G.gemm(G.TN, G.EPI_ATOMIC_F32, dY2, X, dW2), 2.0 * M * 1024 * 1024)
    print(f"TN wgrad 1024x1024 over {M}: ours {ms:.3f} ms {tf:.0f} TF/s", flush=True)


if __name__ == "__main__":
    main()
