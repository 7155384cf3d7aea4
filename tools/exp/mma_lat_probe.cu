// Probe (not part of the library): issue cost, latency and throughput of small tcgen05.mma bursts as the attention
// kernels issue them (4 x [128x128x16] SS into one accumulator; 8 x [128x64x16] TS; long chains), one CTA, nothing
// else running.  Operand contents are irrelevant (shared memory is zero-filled).
//   nvcc -cudart shared -gencode arch=compute_100a,code=sm_100a -I dalle_b200/ops/csrc -I include \
//        tools/exp/mma_lat_probe.cu -o tools/exp/mma_lat_probe
#include "sm100_ptx.cuh"
#include <cstdio>
using namespace b200;

__global__ void __launch_bounds__(128, 1) probe(long long* out) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + (((raw + 1023u) & ~1023u) - raw);
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 65536);
  uint32_t* tslot = reinterpret_cast<uint32_t*>(bar + 2);
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < 65536 / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0u;
  if (threadIdx.x == 0) { mbar_init(bar, 1); fence_barrier_init(); }
  if (warp == 0) { tmem_alloc(tslot, 512); tmem_relinquish(); }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tslot;
  if (warp == 1 && elect_one()) {
    const uint64_t da = make_smem_desc_sw128(smem_u32(smem), 0, 1024);
    const uint64_t db = make_smem_desc_sw128(smem_u32(smem + 16384), 0, 1024);
    const uint64_t dv = make_smem_desc_sw128(smem_u32(smem + 32768), 8192, 1024);
    constexpr uint32_t id_s = make_idesc_bf16(128, 128, 0, 0);
    constexpr uint32_t id_pv = make_idesc_bf16(128, 64, 0, 1);
    uint32_t ph = 0;
    int o = 0;
    for (int rep = 0; rep < 3; ++rep) {
      // (a) 4 dependent SS MMAs, commit, wait
      long long t0 = clock64();
#pragma unroll
      for (int kk = 0; kk < 4; ++kk) umma_bf16_ss(tmem, da + 2 * kk, db + 2 * kk, id_s, kk > 0);
      long long t1 = clock64();
      umma_commit(bar);
      long long t2 = clock64();
      mbar_wait(bar, ph); ph ^= 1;
      long long t3 = clock64();
      out[o++] = t1 - t0; out[o++] = t2 - t0; out[o++] = t3 - t0;
      // (b) 8 dependent TS MMAs (A from TMEM cols 256..), commit, wait
      t0 = clock64();
#pragma unroll
      for (int k16 = 0; k16 < 8; ++k16) umma_bf16_ts(tmem + 128, tmem + 256 + k16 * 8, dv + 128 * k16, id_pv, k16 > 0);
      t1 = clock64();
      umma_commit(bar);
      t2 = clock64();
      mbar_wait(bar, ph); ph ^= 1;
      t3 = clock64();
      out[o++] = t1 - t0; out[o++] = t2 - t0; out[o++] = t3 - t0;
      // (c) 64 dependent SS MMAs
      t0 = clock64();
      for (int r = 0; r < 16; ++r)
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) umma_bf16_ss(tmem, da + 2 * kk, db + 2 * kk, id_s, 1u);
      t1 = clock64();
      umma_commit(bar);
      mbar_wait(bar, ph); ph ^= 1;
      t3 = clock64();
      out[o++] = t1 - t0; out[o++] = 0; out[o++] = t3 - t0;
      // (d) 64 SS MMAs alternating between two accumulators
      t0 = clock64();
      for (int r = 0; r < 16; ++r)
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) umma_bf16_ss(tmem + (kk & 1) * 128, da + 2 * kk, db + 2 * kk, id_s, 1u);
      t1 = clock64();
      umma_commit(bar);
      mbar_wait(bar, ph); ph ^= 1;
      t3 = clock64();
      out[o++] = t1 - t0; out[o++] = 0; out[o++] = t3 - t0;
      // (e) 64 dependent TS MMAs N=64
      t0 = clock64();
      for (int r = 0; r < 8; ++r)
#pragma unroll
        for (int k16 = 0; k16 < 8; ++k16) umma_bf16_ts(tmem + 128, tmem + 256 + k16 * 8, dv + 128 * k16, id_pv, 1u);
      t1 = clock64();
      umma_commit(bar);
      mbar_wait(bar, ph); ph ^= 1;
      t3 = clock64();
      out[o++] = t1 - t0; out[o++] = 0; out[o++] = t3 - t0;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) { tc_fence_after(); tmem_dealloc(tmem, 512); }
}

int main() {
  long long* d; cudaMalloc(&d, 64 * 8); cudaMemset(d, 0, 64 * 8);
  const int smem = 65536 + 1024 + 256;
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  probe<<<1, 128, smem>>>(d);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("failed: %s\n", cudaGetErrorString(e)); return 1; }
  long long h[64]; cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
  const char* names[5] = {"4 SS 128x128x16 (one acc)", "8 TS 128x64x16 (one acc)", "64 SS dependent", "64 SS two accs", "64 TS dependent"};
  for (int rep = 0; rep < 3; ++rep)
    for (int k = 0; k < 5; ++k)
      printf("rep %d  %-28s issue %6lld  +commit %6lld  done %6lld cycles\n", rep, names[k], h[(rep * 5 + k) * 3], h[(rep * 5 + k) * 3 + 1], h[(rep * 5 + k) * 3 + 2]);
  return 0;
}
