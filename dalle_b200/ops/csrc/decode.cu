// Single-token decode kernels for KV-cached sampling (replaces the ~2000 eager launches
// per token of dalle-pytorch's cached forward + top-k/gumbel sampling used by reference inference/run_inference.py).
// Every kernel reads the current position from DEVICE memory so that one whole decode step (64 layers + head +
// sampler) is captured once in a CUDA graph and replayed 1281 times.
//
//   decode_embed      x1 = x2 = W[ids[b, pos]]
//   decode_ln_shift   LN of one token per sample + token shift from the per-block cache of earlier normalised tokens
//   decode_attn       append (k, v) of the token to the layer's cache, then attention over the keys the sparsity
//                     pattern allows at this position (all text keys + row prefix / column prefix / conv window)
//   decode_sample     temperature / top-k / top-p / multinomial over the image-vocabulary logits
//   decode_advance    pos += 1
#include "sm100_ptx.cuh"
#include "b200_api.h"

namespace b200 {

constexpr float kLog2eDec = 1.4426950408889634f;

struct DecGeom {
  int B, D, H, seq, T, F, conv_half;
};

__global__ void decode_embed_kernel(const long long* __restrict__ ids, int ids_stride, const float* __restrict__ W,
                                    float* __restrict__ x1, float* __restrict__ x2, const int* __restrict__ pos_ptr,
                                    int B, int D) {
  const int b = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (b >= B) return;
  const long long id = ids[static_cast<long long>(b) * ids_stride + *pos_ptr];
  for (int c = lane * 4; c < D; c += 128) {
    const float4 v = __ldg(reinterpret_cast<const float4*>(W + id * D + c));
    *reinterpret_cast<float4*>(x1 + static_cast<long long>(b) * D + c) = v;
    *reinterpret_cast<float4*>(x2 + static_cast<long long>(b) * D + c) = v;
  }
}

// cache: [B, seq, D/2] bf16 holding the first half of the channels of every earlier normalised token
__global__ void decode_ln_shift_kernel(const float* __restrict__ x, const float* __restrict__ x2,
                                       const float* __restrict__ w, const float* __restrict__ bias,
                                       __nv_bfloat16* __restrict__ cache, __nv_bfloat16* __restrict__ out,
                                       const int* __restrict__ pos_ptr, DecGeom g, float eps, int shift) {
  const int b = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (b >= g.B) return;
  const int pos = *pos_ptr, D = g.D;
  const int nv = D / 128;
  // which earlier positions feed the shifted channels
  int src_a = -1, src_b = -1;  // source position for channels [0, D/4) and [D/4, D/2)
  if (shift) {
    if (pos < g.T) {
      src_a = src_b = pos >= 1 ? pos - 1 : -1;
    } else {
      const int p = pos - g.T, r = p / g.F, c = p - r * g.F;
      src_a = r >= 1 ? pos - g.F : -1;
      src_b = c >= 1 ? pos - 1 : -1;
    }
  }
  // every load that does not depend on the statistics is issued up front (the kernel is one short latency chain)
  float4 v[8], ww[8], bb[8];
  uint2 cached[8];
  float s = 0.f;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    if (k < nv) {
      const int col = k * 128 + lane * 4;
      const long long off = static_cast<long long>(b) * D + col;
      v[k] = *reinterpret_cast<const float4*>(x + off);
      if (x2) {
        const float4 u = *reinterpret_cast<const float4*>(x2 + off);
        v[k] = make_float4(0.5f * (v[k].x + u.x), 0.5f * (v[k].y + u.y), 0.5f * (v[k].z + u.z), 0.5f * (v[k].w + u.w));
      }
      ww[k] = __ldg(reinterpret_cast<const float4*>(w + col));
      bb[k] = __ldg(reinterpret_cast<const float4*>(bias + col));
      cached[k] = make_uint2(0u, 0u);
      if (shift && col < D / 2) {
        const int src = col < D / 4 ? src_a : src_b;
        if (src >= 0)
          cached[k] = *reinterpret_cast<const uint2*>(cache + (static_cast<long long>(b) * g.seq + src) * (D / 2) + col);
      }
      s += v[k].x + v[k].y + v[k].z + v[k].w;
    }
  }
  const float mu = warp_sum(s) / D;
  float q = 0.f;
#pragma unroll
  for (int k = 0; k < 8; ++k)
    if (k < nv) {
      const float a = v[k].x - mu, c = v[k].y - mu, d = v[k].z - mu, e = v[k].w - mu;
      q += a * a + c * c + d * d + e * e;
    }
  const float rs = rsqrtf(warp_sum(q) / D + eps);
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    if (k < nv) {
      const int col = k * 128 + lane * 4;
      uint2 o;
      o.x = pack_bf16x2((v[k].x - mu) * rs * ww[k].x + bb[k].x, (v[k].y - mu) * rs * ww[k].y + bb[k].y);
      o.y = pack_bf16x2((v[k].z - mu) * rs * ww[k].z + bb[k].z, (v[k].w - mu) * rs * ww[k].w + bb[k].w);
      if (shift && col < D / 2) {
        __nv_bfloat16* crow = cache + (static_cast<long long>(b) * g.seq + pos) * (D / 2) + col;
        *reinterpret_cast<uint2*>(crow) = o;  // remember this token's own normalised half for later positions
        o = cached[k];
      }
      *reinterpret_cast<uint2*>(out + static_cast<long long>(b) * D + col) = o;
    }
  }
}

// one block of 4 warps per (sample, head); the key list is dealt to the warps in groups of 32.  Phase 1: lanes <-> keys
// (each lane dots q with whole K rows, 128 bytes in flight per key); phase 2: block softmax; phase 3: lanes <-> head
// dims, 16 V rows in flight per warp; the four partial outputs meet in shared memory.  (One warp per (sample, head)
// left 4 warps per SM chasing L2 round trips: 26 us per layer.)
constexpr int DEC_MAX_KEYS = 352;   // 257 text + 32 axial (+ window) rounded up
__global__ void __launch_bounds__(128)
decode_attn_kernel(const __nv_bfloat16* __restrict__ qkv, __nv_bfloat16* __restrict__ kcache,
                   __nv_bfloat16* __restrict__ vcache, __nv_bfloat16* __restrict__ out,
                   const int* __restrict__ pos_ptr, DecGeom g, int mode) {
  __shared__ float sq[64];
  __shared__ float sp[DEC_MAX_KEYS];
  __shared__ int spos[DEC_MAX_KEYS];
  __shared__ float sred[2][4];
  __shared__ float sacc[4][64];
  __shared__ int s_cnt;
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31, tid = threadIdx.x;
  const int wid = blockIdx.x;
  if (wid >= g.B * g.H) return;
  const int b = wid / g.H, h = wid - b * g.H, pos = *pos_ptr;
  const int HD = g.H * 64;
  const __nv_bfloat16* row = qkv + static_cast<long long>(b) * 3 * HD + h * 64;
  const long long cbase = (static_cast<long long>(b) * g.H + h) * g.seq * 64;
  if (w == 0) {
    const float2 q = unpack_bf16x2(*reinterpret_cast<const uint32_t*>(row + lane * 2));
    sq[lane * 2] = q.x;
    sq[lane * 2 + 1] = q.y;
  } else if (w == 1) {
    *reinterpret_cast<uint32_t*>(kcache + cbase + static_cast<long long>(pos) * 64 + lane * 2) =
        *reinterpret_cast<const uint32_t*>(row + HD + lane * 2);
  } else if (w == 2) {
    *reinterpret_cast<uint32_t*>(vcache + cbase + static_cast<long long>(pos) * 64 + lane * 2) =
        *reinterpret_cast<const uint32_t*>(row + 2 * HD + lane * 2);
  }

  // ---- the list of key positions this query may attend to ----
  int n_keys;
  if (pos < g.T) {
    n_keys = pos + 1;
    for (int i = tid; i < n_keys; i += 128) spos[i] = i;
  } else {
    for (int i = tid; i < g.T; i += 128) spos[i] = i;
    const int p = pos - g.T, r = p / g.F, c = p - r * g.F;
    if (mode == 1) {
      n_keys = g.T + c + 1;
      for (int i = tid; i <= c; i += 128) spos[g.T + i] = g.T + r * g.F + i;
    } else if (mode == 2) {
      n_keys = g.T + r + 1;
      for (int i = tid; i <= r; i += 128) spos[g.T + i] = g.T + i * g.F + c;
    } else if (mode == 3) {
      if (tid == 0) {
        int cnt = 0;
        for (int rr = max(0, r - g.conv_half); rr <= r; ++rr)
          for (int cc = max(0, c - g.conv_half); cc <= min(g.F - 1, c + g.conv_half); ++cc)
            if (rr * g.F + cc <= p) spos[g.T + cnt++] = g.T + rr * g.F + cc;
        s_cnt = cnt;
      }
      __syncthreads();
      n_keys = g.T + s_cnt;
    } else {
      n_keys = min(pos + 1, DEC_MAX_KEYS);
      for (int i = g.T + tid; i < n_keys; i += 128) spos[i] = i;
    }
  }
  __syncthreads();   // key list, q, and this position's K / V rows (global writes by this block) are in place

  // ---- phase 1: scores (thread tid takes keys tid, tid + 128, ...) ----
  float mx = -1e30f;
  for (int i = tid; i < n_keys; i += 128) {
    const uint4* kr = reinterpret_cast<const uint4*>(kcache + cbase + static_cast<long long>(spos[i]) * 64);
    uint4 kv[8];
#pragma unroll
    for (int c8 = 0; c8 < 8; ++c8) kv[c8] = kr[c8];   // the whole 128-byte row in flight at once
    float sc = 0.f;
#pragma unroll
    for (int c8 = 0; c8 < 8; ++c8) {
      const uint4 u = kv[c8];
      float2 f;
      const float* qq = &sq[c8 * 8];
      f = unpack_bf16x2(u.x); sc += f.x * qq[0] + f.y * qq[1];
      f = unpack_bf16x2(u.y); sc += f.x * qq[2] + f.y * qq[3];
      f = unpack_bf16x2(u.z); sc += f.x * qq[4] + f.y * qq[5];
      f = unpack_bf16x2(u.w); sc += f.x * qq[6] + f.y * qq[7];
    }
    sc *= kLog2eDec;
    sp[i] = sc;
    mx = fmaxf(mx, sc);
  }
  mx = warp_max(mx);
  if (lane == 0) sred[0][w] = mx;
  __syncthreads();
  mx = fmaxf(fmaxf(sred[0][0], sred[0][1]), fmaxf(sred[0][2], sred[0][3]));
  // ---- phase 2: probabilities ----
  float l = 0.f;
  for (int i = tid; i < n_keys; i += 128) {
    const float p = exp2f(sp[i] - mx);
    sp[i] = p;
    l += p;
  }
  l = warp_sum(l);
  if (lane == 0) sred[1][w] = l;
  __syncthreads();
  l = (sred[1][0] + sred[1][1]) + (sred[1][2] + sred[1][3]);
  // ---- phase 3: weighted sum of V rows; lane owns dims (2*lane, 2*lane+1); warp w takes keys w, w + 4, ... ----
  float a0 = 0.f, a1 = 0.f;
  int i = w;
  for (; i + 60 < n_keys; i += 64) {   // 16 rows of this warp in flight
    uint32_t vu[16];
#pragma unroll
    for (int j = 0; j < 16; ++j)
      vu[j] = *reinterpret_cast<const uint32_t*>(vcache + cbase + static_cast<long long>(spos[i + 4 * j]) * 64 + lane * 2);
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      const float2 vf = unpack_bf16x2(vu[j]);
      const float p = sp[i + 4 * j];
      a0 += p * vf.x;
      a1 += p * vf.y;
    }
  }
  for (; i < n_keys; i += 4) {
    const float2 vf = unpack_bf16x2(*reinterpret_cast<const uint32_t*>(vcache + cbase + static_cast<long long>(spos[i]) * 64 + lane * 2));
    a0 += sp[i] * vf.x;
    a1 += sp[i] * vf.y;
  }
  sacc[w][lane * 2] = a0;
  sacc[w][lane * 2 + 1] = a1;
  __syncthreads();
  if (w == 0) {
    const float inv = 1.0f / l;
    const float o0 = (sacc[0][lane * 2] + sacc[1][lane * 2]) + (sacc[2][lane * 2] + sacc[3][lane * 2]);
    const float o1 = (sacc[0][lane * 2 + 1] + sacc[1][lane * 2 + 1]) + (sacc[2][lane * 2 + 1] + sacc[3][lane * 2 + 1]);
    *reinterpret_cast<uint32_t*>(out + static_cast<long long>(b) * HD + h * 64 + lane * 2) = pack_bf16x2(o0 * inv, o1 * inv);
  }
}

__device__ __forceinline__ uint32_t hash_u32(uint32_t x) {
  x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16;
  return x;
}

// One block (1024 threads) per sample: sort the n (<= 8192) logits descending (bitonic, in shared memory),
// cut by top-k / top-p on the RAW logits (temperature 1), then draw one token from the kept set at the given
// temperature - the order of dalle-pytorch's generate_images and of models/generate.py:sample_from_logits
// (filter, then divide by the temperature).  Writes ids[b, pos + 1] = token + id_offset and codes[b, pos + 1 - T].
constexpr int SAMPLE_MAX = 8192;
template <int J>
__device__ __forceinline__ void bitonic_local(uint32_t (&key)[8], int i0, int k) {
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    if ((e & J) == 0) {
      const bool desc = ((i0 + e) & k) == 0;
      const uint32_t x = key[e], y = key[e | J];
      const bool sw = desc ? (x < y) : (x > y);
      key[e] = sw ? y : x;
      key[e | J] = sw ? x : y;
    }
  }
}
__global__ void __launch_bounds__(1024)
decode_sample_kernel(const __nv_bfloat16* __restrict__ logits, int ld, int n, float inv_temperature, int top_k,
                     float top_p, const unsigned long long* __restrict__ seed_ptr, long long* __restrict__ ids, int ids_stride,
                     long long id_offset, long long* __restrict__ codes, int codes_stride,
                     const int* __restrict__ pos_ptr, int T) {
  // 32-bit keys: the logits are bf16, so (ordered 16-bit pattern << 13) | (8191 - index) orders by value, then by index
  extern __shared__ uint32_t skeys[];
  __shared__ float sred[32];
  __shared__ float sbc[2];
  const int b = blockIdx.x, tid = threadIdx.x, pos = *pos_ptr;
  const __nv_bfloat16* row = logits + static_cast<long long>(b) * ld;
  // Bitonic sort of 8192 keys (padding keys are 0 and sort last), descending, 8 consecutive keys per thread held in
  // REGISTERS: exchange distances below 8 stay inside a thread, 8..128 are warp shuffles, only distances >= 256 (15 of
  // the 91 stages) go through shared memory.  (The first version did all 91 stages in shared memory with a block
  // barrier each: ~90 us of the 113 us kernel.)
  constexpr int npow = SAMPLE_MAX;
  constexpr int E = SAMPLE_MAX / 1024;
  uint32_t key[E];
#pragma unroll
  for (int e = 0; e < E; ++e) {
    const int i = tid * E + e;
    uint32_t kk = 0u;  // sorts last
    if (i < n) {
      uint32_t u = *reinterpret_cast<const unsigned short*>(row + i);   // the order is the same at any temperature > 0
      u = (u & 0x8000u) ? (~u & 0xffffu) : (u | 0x8000u);              // monotone map bf16 -> 16-bit uint
      kk = (u << 13) | (8191u - static_cast<uint32_t>(i));
    }
    key[e] = kk;
  }
  const int lane_ = tid & 31;
  for (int k = 2; k <= npow; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      if (j >= E * 32) {
        // partner in another warp: through shared memory
        const int m = j / E;
#pragma unroll
        for (int e = 0; e < E; ++e) skeys[tid * E + e] = key[e];
        __syncthreads();
        const bool lower = (tid & m) == 0;
        const bool desc = ((tid * E) & k) == 0;       // (k > j >= 256: the bit is the same for the thread's 8 keys)
#pragma unroll
        for (int e = 0; e < E; ++e) {
          const uint32_t o = skeys[(tid ^ m) * E + e];
          const bool take_max = desc == lower;
          key[e] = take_max ? (key[e] > o ? key[e] : o) : (key[e] < o ? key[e] : o);
        }
        __syncthreads();
      } else if (j >= E) {
        const int m = j / E;                           // lane distance 1..16
        const bool lower = (lane_ & m) == 0;
#pragma unroll
        for (int e = 0; e < E; ++e) {
          const uint32_t o = __shfl_xor_sync(0xffffffffu, key[e], m);
          const bool desc = ((tid * E + e) & k) == 0;
          const bool take_max = desc == lower;
          key[e] = take_max ? (key[e] > o ? key[e] : o) : (key[e] < o ? key[e] : o);
        }
      } else {
        // inside the thread (j = 4, 2, 1): compile-time distances keep the keys in registers
        if (j == 4) bitonic_local<4>(key, tid * E, k);
        else if (j == 2) bitonic_local<2>(key, tid * E, k);
        else bitonic_local<1>(key, tid * E, k);
      }
    }
  }
#pragma unroll
  for (int e = 0; e < E; ++e) skeys[tid * E + e] = key[e];
  __syncthreads();
  auto val = [&](int i) -> float {
    uint32_t u = skeys[i] >> 13;                       // ordered 16-bit pattern (0 for the padding keys)
    u = (u & 0x8000u) ? (u & 0x7fffu) : (~u & 0xffffu);
    return __uint_as_float(u << 16);
  };
  const float mx = val(0);
  int keep = n;
  if (top_k > 0 && top_k < keep) keep = top_k;
  // Block-parallel nucleus cut and draw over the sorted row (a single-thread walk costs > 1 ms when the
  // distribution is flat and the nucleus holds thousands of tokens): every thread owns `per` consecutive ranks.
  __shared__ float swarp[32];
  __shared__ int smin[2];
  const int per = (npow + blockDim.x - 1) / blockDim.x;   // <= 8
  const int i0 = tid * per;
  float e[8];    // unnormalised probabilities at temperature 1 (nucleus cut), then at the sampling temperature (draw)
  float local = 0.f;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const int i = i0 + k;
    e[k] = (k < per && i < keep) ? __expf(val(i) - mx) : 0.f;
    local += e[k];
  }
  // exclusive scan of the per-thread sums over the block
  auto block_exclusive = [&](float v, float& total_out) -> float {
    const int lane = tid & 31, w = tid >> 5;
    float inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const float t = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += t;
    }
    __syncthreads();                 // (swarp reuse between calls)
    if (lane == 31) swarp[w] = inc;
    __syncthreads();
    if (w == 0) {
      float x = lane < (blockDim.x >> 5) ? swarp[lane] : 0.f;
      float xi = x;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const float t = __shfl_up_sync(0xffffffffu, xi, o);
        if (lane >= o) xi += t;
      }
      swarp[lane] = xi - x;          // exclusive warp offsets
      if (lane == 31) sbc[0] = xi;   // block total
    }
    __syncthreads();
    total_out = sbc[0];
    return swarp[w] + inc - v;
  };
  float total;
  const float off = block_exclusive(local, total);
  if (tid == 0) { smin[0] = keep; smin[1] = 0x7fffffff; }
  __syncthreads();
  if (top_p < 1.0f) {
    // keep tokens while the mass BEFORE them is <= top_p: the cut is the first rank whose preceding mass exceeds it
    float cum = off;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const int i = i0 + k;
      if (k < per && i < keep && cum / total > top_p) { atomicMin(&smin[0], i); break; }
      cum += e[k];
    }
  }
  __syncthreads();
  const int cut = max(smin[0], 1);
  // mass of the kept prefix at the sampling temperature, then the draw
  float kl = 0.f;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const int i = i0 + k;
    e[k] = (k < per && i < cut) ? __expf((val(i) - mx) * inv_temperature) : 0.f;
    kl += e[k];
  }
  float kept;
  const float koff = block_exclusive(kl, kept);
  const unsigned long long seed = *seed_ptr;
  const uint32_t r = hash_u32(static_cast<uint32_t>(seed) ^ hash_u32(static_cast<uint32_t>(seed >> 32) + 0x9e3779b9u * (pos + 1)) ^
                              hash_u32(0x85ebca6bu * (b + 1)));
  const float u = (static_cast<float>(r >> 8) + 0.5f) * (1.0f / 16777216.0f) * kept;
  {
    float acc = koff;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const int i = i0 + k;
      if (k < per && i < cut) {
        acc += e[k];
        if (acc >= u) { atomicMin(&smin[1], i); break; }
      }
    }
  }
  __syncthreads();
  if (tid == 0) {
    const int pick = smin[1] == 0x7fffffff ? cut - 1 : smin[1];
    const long long tok = 8191u - (skeys[pick] & 0x1fffu);
    ids[static_cast<long long>(b) * ids_stride + pos + 1] = tok + id_offset;
    if (pos + 1 - T >= 0) codes[static_cast<long long>(b) * codes_stride + (pos + 1 - T)] = tok;
  }
}

__global__ void decode_advance_kernel(int* pos_ptr) { *pos_ptr += 1; }

static inline int derr() { return cudaGetLastError() == cudaSuccess ? 0 : -70; }

}  // namespace b200

using namespace b200;

extern "C" int b200_decode_embed(const long long* ids, int ids_stride, const float* W, float* x1, float* x2,
                                 const int* pos_ptr, int B, int D, void* stream) {
  decode_embed_kernel<<<(B * 32 + 255) / 256, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(ids, ids_stride, W, x1,
                                                                                               x2, pos_ptr, B, D);
  return derr();
}

extern "C" int b200_decode_ln_shift(const float* x, const float* x2, const float* w, const float* bias, void* cache,
                                    void* out, const int* pos_ptr, int B, int D, int seq, int T, int F, float eps,
                                    int shift, void* stream) {
  if (D % 128 || D > 1024) return -71;
  DecGeom g{B, D, 0, seq, T, F, 0};
  decode_ln_shift_kernel<<<(B * 32 + 255) / 256, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      x, x2, w, bias, reinterpret_cast<__nv_bfloat16*>(cache), reinterpret_cast<__nv_bfloat16*>(out), pos_ptr, g, eps,
      shift);
  return derr();
}

extern "C" int b200_decode_attn(const void* qkv, void* kcache, void* vcache, void* out, const int* pos_ptr, int B,
                                int H, int seq, int T, int F, int mode, int conv_kernel, void* stream) {
  DecGeom g{B, H * 64, H, seq, T, F, conv_kernel / 2};
  if (T + F + 32 > DEC_MAX_KEYS && mode != 0) return -74;
  decode_attn_kernel<<<B * H, 128, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const __nv_bfloat16*>(qkv), reinterpret_cast<__nv_bfloat16*>(kcache),
      reinterpret_cast<__nv_bfloat16*>(vcache), reinterpret_cast<__nv_bfloat16*>(out), pos_ptr, g, mode);
  return derr();
}

extern "C" int b200_decode_sample(const void* logits, int ld, int n, float temperature, int top_k, float top_p,
                                  const unsigned long long* seed, long long* ids, int ids_stride, long long id_offset,
                                  long long* codes, int codes_stride, const int* pos_ptr, int T, int B, void* stream) {
  if (n > SAMPLE_MAX) return -72;
  const size_t smem = static_cast<size_t>(SAMPLE_MAX) * sizeof(uint32_t);
  static bool configured = false;
  if (!configured) {
    if (cudaFuncSetAttribute(decode_sample_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             SAMPLE_MAX * (int)sizeof(uint32_t)) != cudaSuccess)
      return -73;
    configured = true;
  }
  decode_sample_kernel<<<B, 1024, smem, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const __nv_bfloat16*>(logits), ld, n, 1.0f / fmaxf(temperature, 1e-6f), top_k, top_p, seed, ids,
      ids_stride, id_offset, codes, codes_stride, pos_ptr, T);
  return derr();
}

extern "C" int b200_decode_advance(int* pos_ptr, void* stream) {
  decode_advance_kernel<<<1, 1, 0, reinterpret_cast<cudaStream_t>(stream)>>>(pos_ptr);
  return derr();
}
