// Probe (not part of the library): D[128 x 64] = P[128 x 64] * V[64 x 64] with the A operand written to TMEM by
// thread-per-row tcgen05.st (two bf16 per column) and read by tcgen05.mma from TMEM; V is the MN-major B operand
// from a TMA-filled 128B-swizzled tile.  This is the P*V step of the attention kernels without the shared-memory
// round trip of P.
//   nvcc -gencode arch=compute_100a,code=sm_100a -I dalle_b200/ops/csrc -I include tools/exp/ts_probe.cu \
//        -L dalle_b200/ops -ldalle_b200 -lcuda -o tools/exp/ts_probe
#include "sm100_ptx.cuh"
#include <cstdio>
#include <cstdlib>
#include <vector>

namespace b200 {
int make_tmap_bf16_2d(CUtensorMap* m, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_cols,
                      uint32_t box_rows);
}
using namespace b200;

__global__ void __launch_bounds__(128, 1)
ts_probe_kernel(const __grid_constant__ CUtensorMap tmV, const __nv_bfloat16* P, float* D) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + (((raw + 1023u) & ~1023u) - raw);
  uint8_t* sV = smem;   // 64 keys x 128 B
  uint64_t* bar_v = reinterpret_cast<uint64_t*>(smem + 8192);
  uint64_t* bar_d = bar_v + 1;
  uint32_t* tslot = reinterpret_cast<uint32_t*>(bar_v + 2);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    mbar_init(bar_v, 1);
    mbar_init(bar_d, 1);
    fence_barrier_init();
  }
  if (warp == 0) {
    tmem_alloc(tslot, 128);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tslot;
  if (threadIdx.x == 0) {
    mbar_expect_tx(bar_v, 8192);
    tma_load_2d(sV, &tmV, bar_v, 0, 0);
  }
  // A: thread = row; 64 keys -> 32 packed columns at TMEM columns [64, 96)
  const int row = warp * 32 + lane;
  const uint32_t trow = tmem + (static_cast<uint32_t>(warp * 32) << 16);
  uint32_t pk[32];
  const uint32_t* prow = reinterpret_cast<const uint32_t*>(P + row * 64);
#pragma unroll
  for (int j = 0; j < 32; ++j) pk[j] = prow[j];
  {
    uint32_t lo[16], hi[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) { lo[j] = pk[j]; hi[j] = pk[16 + j]; }
    tmem_st16(trow + 64, lo);
    tmem_st16(trow + 64 + 16, hi);
  }
  tmem_st_wait();
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x == 0) {
    tc_fence_after();
    mbar_wait(bar_v, 0);
    constexpr uint32_t idesc = make_idesc_bf16(128, 64, 0, 1);
    const uint32_t v_addr = smem_u32(sV);
#pragma unroll
    for (int k16 = 0; k16 < 4; ++k16)
      umma_bf16_ts(tmem, tmem + 64 + k16 * 8, make_smem_desc_sw128(v_addr + k16 * 2048, 8192, 1024), idesc,
                   k16 > 0 ? 1u : 0u);
    umma_commit(bar_d);
  }
  mbar_wait(bar_d, 0);
  tc_fence_after();
  for (int c = 0; c < 2; ++c) {
    uint32_t r[32];
    tmem_ld32(trow + c * 32, r);
    tmem_ld_wait();
    for (int j = 0; j < 32; ++j) D[row * 64 + c * 32 + j] = __uint_as_float(r[j]);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem, 128);
  }
}

int main() {
  const int M = 128, K = 64, N = 64;
  std::vector<__nv_bfloat16> hP(M * K), hV(K * N);
  std::vector<float> fP(M * K), fV(K * N);
  srand(3);
  for (int i = 0; i < M * K; ++i) { float v = (rand() % 2001 - 1000) / 1000.f; hP[i] = __float2bfloat16(v); fP[i] = __bfloat162float(hP[i]); }
  for (int i = 0; i < K * N; ++i) { float v = (rand() % 2001 - 1000) / 1000.f; hV[i] = __float2bfloat16(v); fV[i] = __bfloat162float(hV[i]); }
  __nv_bfloat16 *dP, *dV; float* dD;
  cudaMalloc(&dP, M * K * 2); cudaMalloc(&dV, K * N * 2); cudaMalloc(&dD, M * N * 4);
  cudaMemcpy(dP, hP.data(), M * K * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dV, hV.data(), K * N * 2, cudaMemcpyHostToDevice);
  cudaMemset(dD, 0xff, M * N * 4);
  CUtensorMap tv;
  if (make_tmap_bf16_2d(&tv, dV, K, N, N, 64, 64)) { printf("tmap failed\n"); return 2; }
  const int smem = 8192 + 1024 + 256;
  cudaFuncSetAttribute(ts_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  ts_probe_kernel<<<1, 128, smem>>>(tv, dP, dD);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("kernel failed: %s\n", cudaGetErrorString(e)); return 3; }
  std::vector<float> hD(M * N);
  cudaMemcpy(hD.data(), dD, M * N * 4, cudaMemcpyDeviceToHost);
  double maxerr = 0;
  for (int i = 0; i < M; ++i)
    for (int j = 0; j < N; ++j) {
      double ref = 0;
      for (int k = 0; k < K; ++k) ref += (double)fP[i * K + k] * fV[k * N + j];
      double err = fabs(ref - hD[i * N + j]);
      if (!(err <= maxerr)) maxerr = err;
    }
  printf("ts probe: max abs error %.5f (%s)\n", maxerr, maxerr < 1e-2 ? "OK" : "MISMATCH");
  printf("D[0,0]=%f D[1,0]=%f D[64,5]=%f D[127,63]=%f\n", hD[0], hD[64], hD[64 * 64 + 5], hD[127 * 64 + 63]);
  return maxerr < 1e-2 ? 0 : 1;
}
