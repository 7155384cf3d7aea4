// Probe: one 256 x 256 x K tile computed by a CTA pair with tcgen05.mma.cta_group::2 (not part of the library).
//   nvcc -gencode arch=compute_100a,code=sm_100a -I dalle_b200/ops/csrc -I include tools/exp/cg2_probe.cu \
//        -L dalle_b200/ops -ldalle_b200 -lcuda -o tools/exp/cg2_probe
#include "sm100_cg2.cuh"
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_bf16.h>

namespace b200 {
int make_tmap_bf16_2d(CUtensorMap* m, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_cols,
                      uint32_t box_rows);
}
using namespace b200;

constexpr int KB = 4;   // k-blocks of 64 (each with its own smem stage: no ring reuse in the probe)

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(192, 1)
cg2_probe_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, float* C) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + (((raw + 1023u) & ~1023u) - raw);
  uint8_t* sA = smem;                    // KB x 16 KB : my 128 rows of A
  uint8_t* sB = smem + KB * 16384;       // KB x 16 KB : my 128 of the 256 B rows
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + KB * 32768);
  uint64_t* tfull = full + KB;
  uint32_t* tslot = reinterpret_cast<uint32_t*>(tfull + 1);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  if (warp == 0 && lane == 0) {
    for (int k = 0; k < KB; ++k) mbar_init(&full[k], 1);
    mbar_init(tfull, 1);
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc_cg2(tslot, 256);
    tmem_relinquish_cg2();
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();     // the peer's barriers exist before anything signals them
  tc_fence_after();
  const uint32_t tmem = *tslot;

  if (warp == 0 && lane == 0) {
    for (int k = 0; k < KB; ++k) {
      if (rank == 0) mbar_expect_tx(&full[k], 2 * 32768);            // both CTAs' A tile + B half
      const uint32_t bar = map_to_cta(&full[k], 0);                  // the leader's barrier
      tma_load_2d_cg2(sA + k * 16384, &tmA, bar, k * 64, rank * 128);
      tma_load_2d_cg2(sB + k * 16384, &tmB, bar, k * 64, rank * 128);
    }
  } else if (warp == 1 && lane == 0 && rank == 0) {
    constexpr uint32_t idesc = make_idesc_bf16(256, 256, 0, 0);
    for (int k = 0; k < KB; ++k) {
      mbar_wait(&full[k], 0);
      tc_fence_after();
      const uint32_t a = smem_u32(sA + k * 16384), b = smem_u32(sB + k * 16384);
#pragma unroll
      for (int kk = 0; kk < 4; ++kk)
        umma_bf16_ss_cg2(tmem, make_smem_desc_sw128(a + kk * 32, 0, 1024), make_smem_desc_sw128(b + kk * 32, 0, 1024),
                         idesc, (k > 0 || kk > 0) ? 1u : 0u);
    }
    umma_commit_cg2(tfull, 0x3);
  } else if (warp >= 2) {
    const int q = warp & 3;
    mbar_wait(tfull, 0);
    tc_fence_after();
    const int row = rank * 128 + q * 32 + lane;
    for (int c = 0; c < 8; ++c) {
      uint32_t r[32];
      tmem_ld32(tmem + (static_cast<uint32_t>(q * 32) << 16) + c * 32, r);
      tmem_ld_wait();
      for (int j = 0; j < 32; ++j) C[row * 256 + c * 32 + j] = __uint_as_float(r[j]);
    }
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc_cg2(tmem, 256);
  }
}

int main() {
  const int M = 256, N = 256, K = KB * 64;
  std::vector<__nv_bfloat16> hA(M * K), hB(N * K);
  std::vector<float> fA(M * K), fB(N * K);
  srand(1);
  for (int i = 0; i < M * K; ++i) { float v = (rand() % 2001 - 1000) / 1000.f; hA[i] = __float2bfloat16(v); fA[i] = __bfloat162float(hA[i]); }
  for (int i = 0; i < N * K; ++i) { float v = (rand() % 2001 - 1000) / 1000.f; hB[i] = __float2bfloat16(v); fB[i] = __bfloat162float(hB[i]); }
  __nv_bfloat16 *dA, *dB; float* dC;
  cudaMalloc(&dA, M * K * 2); cudaMalloc(&dB, N * K * 2); cudaMalloc(&dC, M * N * 4);
  cudaMemcpy(dA, hA.data(), M * K * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dB, hB.data(), N * K * 2, cudaMemcpyHostToDevice);
  cudaMemset(dC, 0xff, M * N * 4);
  CUtensorMap ta, tb;
  if (make_tmap_bf16_2d(&ta, dA, M, K, K, 64, 128) || make_tmap_bf16_2d(&tb, dB, N, K, K, 64, 128)) { printf("tmap failed\n"); return 2; }
  const int smem = KB * 32768 + 1024 + 256;
  cudaFuncSetAttribute(cg2_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  cg2_probe_kernel<<<2, 192, smem>>>(ta, tb, dC);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("kernel failed: %s\n", cudaGetErrorString(e)); return 3; }
  std::vector<float> hC(M * N);
  cudaMemcpy(hC.data(), dC, M * N * 4, cudaMemcpyDeviceToHost);
  double maxerr = 0;
  for (int i = 0; i < M; ++i)
    for (int j = 0; j < N; ++j) {
      double ref = 0;
      for (int k = 0; k < K; ++k) ref += (double)fA[i * K + k] * fB[j * K + k];
      double err = fabs(ref - hC[i * N + j]);
      if (!(err <= maxerr)) maxerr = err;
    }
  printf("cg2 probe: max abs error %.5f (%s)\n", maxerr, maxerr < 1e-2 ? "OK" : "MISMATCH");
  printf("C[0,0]=%f C[0,128]=%f C[128,0]=%f C[255,255]=%f\n", hC[0], hC[128], hC[128 * 256], hC[255 * 256 + 255]);
  return maxerr < 1e-2 ? 0 : 1;
}
