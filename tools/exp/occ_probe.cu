// Occupancy probe: does a kernel that uses tcgen05.alloc get more than one CTA per SM?
#include <cstdio>
#include <cuda_runtime.h>
#include <stdint.h>
__global__ void __launch_bounds__(128, 4) k_plain(float* out) {
  extern __shared__ uint8_t sm[];
  sm[threadIdx.x] = 1;
  __syncthreads();
  out[blockIdx.x * 128 + threadIdx.x] = sm[(threadIdx.x + 1) & 127];
}
template <int COLS>
__global__ void __launch_bounds__(128, 4) k_tmem(float* out, long long* clk) {
  extern __shared__ uint8_t sm[];
  __shared__ uint32_t slot;
  if (threadIdx.x < 32) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(&slot)), "r"(COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  __syncthreads();
  const long long t0 = clock64();
  while (clock64() - t0 < 2000000) {}   // ~1 ms: CTAs co-resident on an SM overlap in time
  if (threadIdx.x == 0) { clk[2 * blockIdx.x] = t0; unsigned s; asm("mov.u32 %0, %%smid;" : "=r"(s)); clk[2 * blockIdx.x + 1] = s; }
  out[blockIdx.x * 128 + threadIdx.x] = sm[threadIdx.x];
  __syncthreads();
  if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(slot), "r"(COLS) : "memory");
}
template <class K> void report(const char* name, K kern, int smem) {
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  int nb = 0;
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, kern, 128, smem);
  printf("%s smem %d: occupancy API says %d CTAs/SM\n", name, smem, nb);
}
int main() {
  float* out; long long* clk;
  cudaMalloc(&out, 1 << 22); cudaMalloc(&clk, 1 << 16);
  report("plain", k_plain, 100 * 1024);
  report("tmem256", k_tmem<256>, 100 * 1024);
  report("tmem256", k_tmem<256>, 32 * 1024);
  report("tmem64", k_tmem<64>, 32 * 1024);
  // launch 2 CTAs per SM worth of blocks and see how many share an SM concurrently (by wall time)
  for (int cols : {256, 64}) {
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    cudaEventRecord(a);
    if (cols == 256) k_tmem<256><<<296, 128, 32 * 1024>>>(out, clk); else k_tmem<64><<<296, 128, 32 * 1024>>>(out, clk);
    cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b);
    printf("tmem%d: 296 CTAs x ~1ms busy loop took %.2f ms (%s) err=%s\n", cols, ms, ms < 1.8 ? ">=2 CTAs/SM ran concurrently" : "serialized", cudaGetErrorString(cudaGetLastError()));
  }
  return 0;
}
