// Device helpers shared by the single-GPU LAMB kernels (lamb.cu) and the fused intra-box averaging kernel
// (group_avg.cu): block reductions, nearest-code search of the 8-bit dynamic maps, block/hyper-parameter records.
#pragma once
#include "sm100_ptx.cuh"

namespace b200 {

constexpr int QBLOCK = 4096;
constexpr int LAMB_THREADS = 256;
constexpr int EPT = QBLOCK / LAMB_THREADS;  // 16 elements per thread

__device__ __forceinline__ float block_sum(float v, float* sred) {
  v = warp_sum(v);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __syncthreads();
  if (lane == 0) sred[warp] = v;
  __syncthreads();
  float t = 0.f;
  if (warp == 0) {
    t = lane < (blockDim.x >> 5) ? sred[lane] : 0.f;
    t = warp_sum(t);
    if (lane == 0) sred[0] = t;
  }
  __syncthreads();
  return sred[0];
}
__device__ __forceinline__ float block_max(float v, float* sred) {
  v = warp_max(v);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __syncthreads();
  if (lane == 0) sred[warp] = v;
  __syncthreads();
  float t = 0.f;
  if (warp == 0) {
    t = lane < (blockDim.x >> 5) ? sred[lane] : -1e38f;
    t = warp_max(t);
    if (lane == 0) sred[0] = t;
  }
  __syncthreads();
  return sred[0];
}

// nearest entry of a sorted 256-entry code (bitsandbytes quantize_cpu: lower bound, then compare neighbours)
__device__ __forceinline__ int nearest_code(const float* code, float x) {
  int lo = 0, hi = 255;  // find largest idx with code[idx] <= x (clamped)
#pragma unroll
  for (int it = 0; it < 8; ++it) {
    const int mid = (lo + hi + 1) >> 1;
    if (code[mid] <= x) lo = mid; else hi = mid - 1;
  }
  int idx = lo;
  if (idx < 255) {
    const float dl = fabsf(x - code[idx]), dr = fabsf(x - code[idx + 1]);
    if (dr < dl) idx += 1;
  }
  return idx;
}

struct LambBlock {   // one entry per 4096-element block of the flat buffer
  int tensor;        // tensor index
  int valid;         // valid elements in this block
  long long state_off;  // offset into the uint8 state (quantised) or fp32 state (small tensors)
  int quantized;
  int absmax_idx;
};

struct LambHyper {
  float beta1, beta2, eps, max_grad_norm, clamp_value, lr, grad_scale;
  int use_clip;
};


}  // namespace b200
