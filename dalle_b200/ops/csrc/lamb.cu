// Fused LAMB with block-wise 8-bit moments over ONE flat parameter buffer (replaces the
// per-tensor CPU loop of reference lib/training/lamb_8bit.py:83-250 and the bitsandbytes quantize/dequantize calls
// it makes).  Math = SURVEY Appendix D.1:
//   clip   : g <- g * min(1, max_norm / (||g|| + 1e-6))                     (lamb_8bit.py:85-87)
//   moments: m <- b1 m + (1-b1) g ; v <- b2 v + (1-b2) g^2                   (dequantise -> update -> requantise,
//            4096-element blocks, 256-entry dynamic maps, absmax per block;   lamb_8bit.py:181-192)
//   update : u = m / (sqrt(v) + eps) + wd * p                                (lamb_8bit.py:195-199)
//   trust  : r = clamp(||p||, 0, clamp) / ||u||  (1 if either is 0) ; p <- p - lr * r * u   (lamb_8bit.py:149-158)
// Layout: every tensor starts on a 4096-element boundary of the flat buffer, so a CUDA block == a quantisation
// block == part of exactly one tensor.  Three launches: grad-norm (+finite flag) / moments+u / apply.
#include "b200_api.h"

#include <cstdio>
#include <mutex>

#include "lamb_device.cuh"

namespace b200 {

// stats[0] += sum g^2 ; stats[1] = 1 if any non-finite value was seen
__global__ void __launch_bounds__(LAMB_THREADS)
lamb_grad_norm_kernel(const float* __restrict__ g, long long n, float* __restrict__ stats) {
  __shared__ float sred[8];
  float s = 0.f;
  bool bad = false;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x * 4;
  for (long long i = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) * 4; i < n; i += stride) {
    const float4 v = *reinterpret_cast<const float4*>(g + i);
    s += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
    bad |= !(isfinite(v.x) && isfinite(v.y) && isfinite(v.z) && isfinite(v.w));
  }
  s = block_sum(s, sred);
  if (threadIdx.x == 0) atomicAdd(&stats[0], s);
  if (bad) stats[1] = 1.f;
}

__global__ void __launch_bounds__(LAMB_THREADS)
lamb_moments_kernel(float* __restrict__ g, const float* __restrict__ p, uint8_t* __restrict__ q1,
                    uint8_t* __restrict__ q2, float* __restrict__ f1, float* __restrict__ f2,
                    float* __restrict__ absmax1, float* __restrict__ absmax2, const float* __restrict__ code1,
                    const float* __restrict__ code2, const LambBlock* __restrict__ blocks,
                    const float* __restrict__ wd_per_tensor, float* __restrict__ tensor_sums,
                    const float* __restrict__ stats, LambHyper hp) {
  __shared__ float sc1[256], sc2[256], sred[8];
  if (stats[1] != 0.f) return;  // non-finite gradients: the whole step is skipped (finite-grad gate)
  sc1[threadIdx.x] = code1[threadIdx.x];
  sc2[threadIdx.x] = code2[threadIdx.x];
  const LambBlock blk = blocks[blockIdx.x];
  const long long base = static_cast<long long>(blockIdx.x) * QBLOCK;
  const int e0 = threadIdx.x * EPT;
  float clip = hp.grad_scale;
  if (hp.use_clip) {
    const float norm = sqrtf(stats[0]) * hp.grad_scale;
    clip *= fminf(1.0f, hp.max_grad_norm / (norm + 1e-6f));
  }
  const float wd = wd_per_tensor[blk.tensor];
  __syncthreads();

  float gv[EPT], pv[EPT], m[EPT], v[EPT];
#pragma unroll
  for (int j = 0; j < EPT; j += 4) {
    const float4 a = *reinterpret_cast<const float4*>(g + base + e0 + j);
    const float4 b = *reinterpret_cast<const float4*>(p + base + e0 + j);
    gv[j] = a.x * clip; gv[j + 1] = a.y * clip; gv[j + 2] = a.z * clip; gv[j + 3] = a.w * clip;
    pv[j] = b.x; pv[j + 1] = b.y; pv[j + 2] = b.z; pv[j + 3] = b.w;
  }
  if (blk.quantized) {
    const float a1 = absmax1[blk.absmax_idx], a2 = absmax2[blk.absmax_idx];
    const uint4 u1 = *reinterpret_cast<const uint4*>(q1 + blk.state_off + e0);
    const uint4 u2 = *reinterpret_cast<const uint4*>(q2 + blk.state_off + e0);
    const uint8_t* b1 = reinterpret_cast<const uint8_t*>(&u1);
    const uint8_t* b2 = reinterpret_cast<const uint8_t*>(&u2);
#pragma unroll
    for (int j = 0; j < EPT; ++j) {
      m[j] = sc1[b1[j]] * a1;
      v[j] = sc2[b2[j]] * a2;
    }
  } else {
#pragma unroll
    for (int j = 0; j < EPT; j += 4) {
      const float4 a = *reinterpret_cast<const float4*>(f1 + blk.state_off + e0 + j);
      const float4 b = *reinterpret_cast<const float4*>(f2 + blk.state_off + e0 + j);
      m[j] = a.x; m[j + 1] = a.y; m[j + 2] = a.z; m[j + 3] = a.w;
      v[j] = b.x; v[j + 1] = b.y; v[j + 2] = b.z; v[j + 3] = b.w;
    }
  }
  float mx1 = 0.f, mx2 = 0.f, su = 0.f, sp = 0.f;
  float u[EPT];
#pragma unroll
  for (int j = 0; j < EPT; ++j) {
    const bool ok = e0 + j < blk.valid;
    m[j] = ok ? hp.beta1 * m[j] + (1.f - hp.beta1) * gv[j] : 0.f;
    v[j] = ok ? hp.beta2 * v[j] + (1.f - hp.beta2) * gv[j] * gv[j] : 0.f;
    mx1 = fmaxf(mx1, fabsf(m[j]));
    mx2 = fmaxf(mx2, fabsf(v[j]));
    u[j] = ok ? m[j] / (sqrtf(v[j]) + hp.eps) + wd * pv[j] : 0.f;
    su += u[j] * u[j];
    sp += ok ? pv[j] * pv[j] : 0.f;
  }
  if (blk.quantized) {
    mx1 = block_max(mx1, sred);
    mx2 = block_max(mx2, sred);
    const float i1 = mx1 > 0.f ? 1.f / mx1 : 0.f, i2 = mx2 > 0.f ? 1.f / mx2 : 0.f;
    uint4 o1, o2;
    uint8_t* b1 = reinterpret_cast<uint8_t*>(&o1);
    uint8_t* b2 = reinterpret_cast<uint8_t*>(&o2);
#pragma unroll
    for (int j = 0; j < EPT; ++j) {
      b1[j] = static_cast<uint8_t>(nearest_code(sc1, m[j] * i1));
      b2[j] = static_cast<uint8_t>(nearest_code(sc2, v[j] * i2));
    }
    *reinterpret_cast<uint4*>(q1 + blk.state_off + e0) = o1;
    *reinterpret_cast<uint4*>(q2 + blk.state_off + e0) = o2;
    if (threadIdx.x == 0) {
      absmax1[blk.absmax_idx] = mx1;
      absmax2[blk.absmax_idx] = mx2;
    }
  } else {
#pragma unroll
    for (int j = 0; j < EPT; j += 4) {
      *reinterpret_cast<float4*>(f1 + blk.state_off + e0 + j) = make_float4(m[j], m[j + 1], m[j + 2], m[j + 3]);
      *reinterpret_cast<float4*>(f2 + blk.state_off + e0 + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
    }
  }
#pragma unroll
  for (int j = 0; j < EPT; j += 4)
    *reinterpret_cast<float4*>(g + base + e0 + j) = make_float4(u[j], u[j + 1], u[j + 2], u[j + 3]);
  su = block_sum(su, sred);
  sp = block_sum(sp, sred);
  if (threadIdx.x == 0) {
    atomicAdd(&tensor_sums[2 * blk.tensor], su);
    atomicAdd(&tensor_sums[2 * blk.tensor + 1], sp);
  }
}

__global__ void __launch_bounds__(LAMB_THREADS)
lamb_apply_kernel(float* __restrict__ g, float* __restrict__ p, __nv_bfloat16* __restrict__ p_bf16,
                  const LambBlock* __restrict__ blocks, const float* __restrict__ tensor_sums,
                  float* __restrict__ trust_out, const float* __restrict__ stats, LambHyper hp, int zero_grad) {
  const LambBlock blk = blocks[blockIdx.x];
  const long long base = static_cast<long long>(blockIdx.x) * QBLOCK;
  const int e0 = threadIdx.x * EPT;
  const bool skip = stats[1] != 0.f;
  float step = 0.f;
  if (!skip) {
    const float un = sqrtf(tensor_sums[2 * blk.tensor]);
    const float wn = fminf(sqrtf(tensor_sums[2 * blk.tensor + 1]), hp.clamp_value);
    const float trust = (wn != 0.f && un != 0.f) ? wn / un : 1.0f;
    step = hp.lr * trust;
    if (threadIdx.x == 0 && trust_out) {
      trust_out[3 * blk.tensor] = wn;
      trust_out[3 * blk.tensor + 1] = un;
      trust_out[3 * blk.tensor + 2] = trust;
    }
  }
#pragma unroll
  for (int j = 0; j < EPT; j += 4) {
    const long long off = base + e0 + j;
    float4 w = *reinterpret_cast<float4*>(p + off);
    if (!skip) {
      const float4 u = *reinterpret_cast<const float4*>(g + off);
      w.x -= step * u.x; w.y -= step * u.y; w.z -= step * u.z; w.w -= step * u.w;
      *reinterpret_cast<float4*>(p + off) = w;
    }
    if (p_bf16) {
      uint2 o;
      o.x = pack_bf16x2(w.x, w.y);
      o.y = pack_bf16x2(w.z, w.w);
      *reinterpret_cast<uint2*>(p_bf16 + off) = o;
    }
    if (zero_grad) *reinterpret_cast<float4*>(g + off) = make_float4(0.f, 0.f, 0.f, 0.f);
  }
}

}  // namespace b200

using namespace b200;

namespace b200 { int num_sms(); }  // gemm_sm100.cu
static int grad_norm_grid() { return b200::num_sms() * 8; }   // eight 256-thread blocks per SM, grid-stride

extern "C" int b200_lamb_step(float* grad, float* param, void* param_bf16, long long n_padded, void* q1, void* q2,
                              float* f1, float* f2, float* absmax1, float* absmax2, const float* code1,
                              const float* code2, const void* blocks, int n_blocks, const float* wd_per_tensor,
                              float* tensor_sums, int n_tensors, float* trust_out, float* stats, float beta1,
                              float beta2, float eps, float max_grad_norm, float clamp_value, float lr,
                              float grad_scale, int zero_grad, void* stream) {
  if (n_padded != static_cast<long long>(n_blocks) * QBLOCK) return -50;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  LambHyper hp;
  hp.beta1 = beta1; hp.beta2 = beta2; hp.eps = eps; hp.max_grad_norm = max_grad_norm; hp.clamp_value = clamp_value;
  hp.lr = lr; hp.grad_scale = grad_scale; hp.use_clip = max_grad_norm > 0.f;
  cudaMemsetAsync(stats, 0, 2 * sizeof(float), st);
  cudaMemsetAsync(tensor_sums, 0, 2 * n_tensors * sizeof(float), st);
  lamb_grad_norm_kernel<<<grad_norm_grid(), LAMB_THREADS, 0, st>>>(grad, n_padded, stats);
  lamb_moments_kernel<<<n_blocks, LAMB_THREADS, 0, st>>>(
      grad, param, reinterpret_cast<uint8_t*>(q1), reinterpret_cast<uint8_t*>(q2), f1, f2, absmax1, absmax2, code1,
      code2, reinterpret_cast<const LambBlock*>(blocks), wd_per_tensor, tensor_sums, stats, hp);
  lamb_apply_kernel<<<n_blocks, LAMB_THREADS, 0, st>>>(grad, param, reinterpret_cast<__nv_bfloat16*>(param_bf16),
                                                       reinterpret_cast<const LambBlock*>(blocks), tensor_sums,
                                                       trust_out, stats, hp, zero_grad);
  return cudaGetLastError() == cudaSuccess ? 0 : -51;
}

// -------------------------------------------------------------------------------------------------
// b200_lamb_step_noalloc: the same optimizer step for callers that have no way to provide scratch memory.
//
// Who needs this: host programs that gave the whole device-memory budget to their own arena at start-up (a serving
// or training runtime with a fixed pool, or an embedding application that is simply not allowed to call cudaMalloc
// or go through torch's caching allocator after initialisation), and control loops - such as the averaging thread of
// the collaborative optimizer - where an allocation could fail late, or stall the device, in the middle of a step.
// b200_lamb_step asks its caller for three small work arrays (stats, tensor_sums, trust_out); here the first two
// live in the fixed array below instead, so the caller passes inputs and outputs only.
//
// The array is part of the library image: every GPU that loads libdalle_b200.so carries these
// B200_LAMB_NOALLOC_SCRATCH_FLOATS * 4 bytes (about 33 KB) until the process exits.  That bounds the problem size:
// at most B200_LAMB_NOALLOC_MAX_TENSORS parameter tensors per call; larger requests are refused with
// B200_ERR_NOALLOC_TOO_MANY_TENSORS and a message on stderr, nothing is launched.
//
// All calls on one device share the array and find in it whatever the previous call left there.  The entry point
// clears, on the caller's stream, exactly the part that has to start from zero (the two statistics and
// 2*n_tensors partial sums) before the kernels run.  Two calls whose kernels can overlap on the GPU (different
// streams without an ordering between them) would race on the array - serialise them.
//
// The first call on each device resolves the device address of the array (cudaGetSymbolAddress, with that device
// current) and caches it; make that first call before starting any stream capture.  No cudaMalloc, cudaFree or
// allocating library routine is called, here or in the kernels.
// -------------------------------------------------------------------------------------------------
namespace b200 {

constexpr int kNoAllocStatsFloats = 64;  // stats[2], padded so the sums start on a 256-byte boundary
static __device__ __align__(256) float g_lamb_noalloc_scratch[B200_LAMB_NOALLOC_SCRATCH_FLOATS];

constexpr int kMaxDevices = 64;
static float* g_lamb_noalloc_ptr[kMaxDevices] = {nullptr};
static std::mutex g_lamb_noalloc_mutex;

// In host code `g_lamb_noalloc_scratch` is only a handle for the symbol, not a device address; the address a
// kernel can use has to be asked for, once per device, with that device current.
static int lamb_noalloc_scratch(float** out) {
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) return B200_ERR_NOALLOC_SETUP;
  std::lock_guard<std::mutex> lock(g_lamb_noalloc_mutex);
  if (g_lamb_noalloc_ptr[dev] == nullptr) {
    float* ptr = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void**>(&ptr), g_lamb_noalloc_scratch);
    if (e != cudaSuccess || ptr == nullptr) {
      fprintf(stderr, "b200_lamb_step_noalloc: cudaGetSymbolAddress failed on device %d: %s\n", dev,
              cudaGetErrorString(e));
      return B200_ERR_NOALLOC_SETUP;
    }
    g_lamb_noalloc_ptr[dev] = ptr;
  }
  *out = g_lamb_noalloc_ptr[dev];
  return 0;
}

}  // namespace b200

extern "C" int b200_lamb_step_noalloc(float* grad, float* param, void* param_bf16, long long n_padded, void* q1,
                                      void* q2, float* f1, float* f2, float* absmax1, float* absmax2,
                                      const float* code1, const float* code2, const void* blocks, int n_blocks,
                                      const float* wd_per_tensor, int n_tensors, float* trust_out, float* stats_out,
                                      float beta1, float beta2, float eps, float max_grad_norm, float clamp_value,
                                      float lr, float grad_scale, int zero_grad, void* stream) {
  if (n_tensors < 0 || n_tensors > B200_LAMB_NOALLOC_MAX_TENSORS) {
    fprintf(stderr,
            "b200_lamb_step_noalloc: %d parameter tensors exceed the built-in scratch (max %d); use b200_lamb_step "
            "with caller-provided scratch instead\n", n_tensors, B200_LAMB_NOALLOC_MAX_TENSORS);
    return B200_ERR_NOALLOC_TOO_MANY_TENSORS;
  }
  if (n_padded != static_cast<long long>(n_blocks) * QBLOCK) return -50;
  float* scratch = nullptr;
  const int rc = lamb_noalloc_scratch(&scratch);
  if (rc != 0) return rc;
  float* stats = scratch;
  float* tensor_sums = scratch + kNoAllocStatsFloats;

  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  LambHyper hp;
  hp.beta1 = beta1; hp.beta2 = beta2; hp.eps = eps; hp.max_grad_norm = max_grad_norm; hp.clamp_value = clamp_value;
  hp.lr = lr; hp.grad_scale = grad_scale; hp.use_clip = max_grad_norm > 0.f;
  // the previous call's statistics and partial sums are still in the array: clear what accumulates
  if (cudaMemsetAsync(stats, 0, 2 * sizeof(float), st) != cudaSuccess) return B200_ERR_NOALLOC_SETUP;
  if (n_tensors > 0 &&
      cudaMemsetAsync(tensor_sums, 0, 2 * static_cast<size_t>(n_tensors) * sizeof(float), st) != cudaSuccess)
    return B200_ERR_NOALLOC_SETUP;
  lamb_grad_norm_kernel<<<grad_norm_grid(), LAMB_THREADS, 0, st>>>(grad, n_padded, stats);
  lamb_moments_kernel<<<n_blocks, LAMB_THREADS, 0, st>>>(
      grad, param, reinterpret_cast<uint8_t*>(q1), reinterpret_cast<uint8_t*>(q2), f1, f2, absmax1, absmax2, code1,
      code2, reinterpret_cast<const LambBlock*>(blocks), wd_per_tensor, tensor_sums, stats, hp);
  lamb_apply_kernel<<<n_blocks, LAMB_THREADS, 0, st>>>(grad, param, reinterpret_cast<__nv_bfloat16*>(param_bf16),
                                                       reinterpret_cast<const LambBlock*>(blocks), tensor_sums,
                                                       trust_out, stats, hp, zero_grad);
  if (cudaGetLastError() != cudaSuccess) return -51;
  if (stats_out != nullptr &&  // {sum g^2, non-finite flag} for callers that want to look at them
      cudaMemcpyAsync(stats_out, stats, 2 * sizeof(float), cudaMemcpyDeviceToDevice, st) != cudaSuccess)
    return -51;
  return 0;
}
