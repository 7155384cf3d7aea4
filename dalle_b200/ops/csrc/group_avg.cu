// Fused intra-box gradient averaging + sharded 8-bit LAMB + parameter all-gather over NVLink peer memory
// (SURVEY §8 row 14, §5.8 level 1).  ONE cooperative kernel per rank; no NCCL call on this path.
//
// Replaces, inside an 8xB200 box: nn.DataParallel's per-step parameter broadcast + gradient reduce,
// and the CPU clip + CPULAMB8Bit step (reference lib/training/lamb_8bit.py:83-158) that follows averaging.
//
// Every rank owns a contiguous shard of the 4096-element blocks of the flat parameter buffer (ZeRO-1 style:
// optimizer state exists only for the shard).  All buffers below are symmetric allocations mapped into every
// peer's address space; `*_ptrs[r]` is rank r's copy.
//
//   phase 0  barrier                      - every rank's backward has finished writing its gradient buffer
//   phase 1  reduce-scatter               - for my blocks: g = grad_scale * sum_r grad_r[block]   (P2P loads;
//            optionally through a bf16 payload copy), ||g||^2 and the finite flag of my shard
//   phase 2  exchange {||g||^2, flag}     - P2P stores into every peer's exchange area, barrier
//   phase 3  LAMB moments on my shard     - clip by the GLOBAL norm, dequantise/update/requantise m and v,
//            u = m/(sqrt(v)+eps) + wd*p, per-tensor partial ||u||^2, ||p||^2
//   phase 4  exchange per-tensor partials - tensors straddle shard boundaries; P2P stores + barrier
//   phase 5  apply + all-gather           - p -= lr*trust*u on my shard, written (fp32 master + bf16 shadow) into
//            EVERY rank's buffers with P2P stores; gradients zeroed
//   phase 6  barrier                      - all peers' stores have landed before anyone starts the next forward
//
// Cross-rank barriers: one 32-bit slot per (rank, peer) in the signal pad, monotonically increasing epoch values,
// st.release.sys / ld.acquire.sys, bracketed by grid syncs.  Spins are bounded (trap instead of hanging the box).
#include <cooperative_groups.h>

#include <cstdlib>

#include "b200_api.h"
#include "lamb_device.cuh"

#include <cuda_fp16.h>

namespace cg = cooperative_groups;

namespace b200 {

constexpr int MAX_RANKS = 8;
// hivemind's averaging codecs on the NVLink hop (reference task.py:124-126): a CHUNK is 32
// consecutive 4096-element blocks = 131072 elements = the 512 KiB part hivemind serialises at a time.
//   CODEC_FP16  clamp to +-65504, fp16                                    (tensors below 65537 elements)
//   CODEC_U8H   mu, sigma over the chunk; 256 uniform bins over 6 sigma; a 256-entry codebook of bucket MEANS
//               (of the original values) travels with the codes              (tensors of 65537 elements and more)
// The payload buffer has 2 bytes per element: an fp16 chunk fills its 256 KiB, a U8H chunk uses 1 byte per element of
// codes followed by the 1 KiB codebook.
constexpr int CHUNK_BLOCKS = 32;
constexpr int CODEC_FP16 = 3, CODEC_U8H = 4;
constexpr long long CHUNK_BYTES = 2ll * CHUNK_BLOCKS * 4096;

struct GroupArgs {
  int rank, world;
  float* grad[MAX_RANKS];
  float* param[MAX_RANKS];
  __nv_bfloat16* shadow[MAX_RANKS];
  __nv_bfloat16* payload[MAX_RANKS];  // bf16 copy of the gradients (payload mode 1), else unused
  uint32_t* sig[MAX_RANKS];           // signal pads: sig[r][peer]
  float* xchg[MAX_RANKS];             // exchange areas: xchg[r][peer * xstride + ...]
  int xstride;                        // 2 + 2 * n_tensors (padded)
  int n_tensors;
  int n_blocks;                       // total blocks of the flat buffer
  int blk0, blk1;                     // my shard
  uint32_t epoch;                     // barrier values epoch+1 .. epoch+4 are used by this launch
  int payload_mode;                   // 0 = fp32 P2P loads, 1 = bf16 payload, 2 = absmax int8 payload (int8 codes + one
                                      //     fp32 scale per 4096-element block), 3 = hivemind's codecs, chosen per chunk
                                      //     of 32 blocks by `chunk_codec` (CODEC_FP16 / CODEC_U8H)
  const uint8_t* chunk_codec;         // [ceil(n_blocks / 32)], payload mode 3
  int zero_grad;
  int reduce_only;                    // 1: stop after the reduce phase and give EVERY rank the reduced gradient
  int spin_limit_s;                   // cross-rank barrier: seconds before a missing peer turns into a trap
  // local optimizer state (shard only is touched)
  uint8_t* q1; uint8_t* q2; float* f1; float* f2; float* absmax1; float* absmax2;
  const float* code1; const float* code2;
  const LambBlock* blocks;
  const float* wd_per_tensor;
  float* tsums;                       // local [2*n_tensors] accumulation scratch
  float* stats;                       // local [4]: {sum g^2 (shard), bad flag (shard), global sum, global flag}
  float* trust_out;                   // local [3*n_tensors]
  LambHyper hp;
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

__device__ void rank_barrier(const GroupArgs& a, uint32_t value, cg::grid_group& grid) {
  __threadfence_system();
  grid.sync();
  if (blockIdx.x == 0 && threadIdx.x < a.world) {
    st_release_sys(a.sig[threadIdx.x] + a.rank, value);
    const uint32_t* mine = a.sig[a.rank] + threadIdx.x;
    // a peer may legitimately be late by a long time (data-loader stall, checkpoint write, first-call JIT of another
    // rank): the spin is bounded only to turn a dead peer into a launch failure - a.spin_limit_s seconds, 600 by
    // default (DALLE_B200_GROUP_TIMEOUT_S); NCCL would wait forever here
    const uint64_t t0 = global_timer_ns();
    const uint64_t limit = static_cast<uint64_t>(a.spin_limit_s) * 1000000000ull;
    uint32_t polls = 0;
    while (static_cast<int32_t>(ld_acquire_sys(mine) - value) < 0) {
      if ((++polls & 0xFFFFu) == 0 && global_timer_ns() - t0 > limit) __trap();
    }
  }
  grid.sync();
}

__global__ void __launch_bounds__(LAMB_THREADS)
group_lamb_kernel(const GroupArgs a) {
  cg::grid_group grid = cg::this_grid();
  __shared__ float sc1[256], sc2[256], sred[8], s_book[256];
  sc1[threadIdx.x] = a.code1[threadIdx.x];
  sc2[threadIdx.x] = a.code2[threadIdx.x];
  const int e0 = threadIdx.x * EPT;
  const long long gtid = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const long long gthreads = static_cast<long long>(gridDim.x) * blockDim.x;

  // local scratch must start at zero
  for (long long i = gtid; i < 2 * a.n_tensors; i += gthreads) a.tsums[i] = 0.f;
  if (gtid < 4) a.stats[gtid] = 0.f;
  if (a.payload_mode == 1) {
    // my gradients -> bf16 payload that the shard owners will read over NVLink
    const long long n4 = static_cast<long long>(a.n_blocks) * QBLOCK / 4;
    const float4* src = reinterpret_cast<const float4*>(a.grad[a.rank]);
    uint2* dst = reinterpret_cast<uint2*>(a.payload[a.rank]);
    for (long long i = gtid; i < n4; i += gthreads) {
      const float4 v = src[i];
      dst[i] = make_uint2(pack_bf16x2(v.x, v.y), pack_bf16x2(v.z, v.w));
    }
  }
  if (a.payload_mode == 2) {
    // my gradients -> uniform 8-bit payload: per 4096-element block, code = rint(x * 127 / absmax) and one fp32
    // step; a quarter of the fp32 NVLink bytes (the intra-box counterpart of hivemind's Uniform8BitQuantization)
    int8_t* codes = reinterpret_cast<int8_t*>(a.payload[a.rank]);
    float* steps = reinterpret_cast<float*>(codes + static_cast<long long>(a.n_blocks) * QBLOCK);
    for (int b = blockIdx.x; b < a.n_blocks; b += gridDim.x) {
      const float* src = a.grad[a.rank] + static_cast<long long>(b) * QBLOCK + e0;
      float v[EPT];
      float am = 0.f;
#pragma unroll
      for (int j = 0; j < EPT; j += 4) {
        const float4 t = *reinterpret_cast<const float4*>(src + j);
        v[j] = t.x; v[j + 1] = t.y; v[j + 2] = t.z; v[j + 3] = t.w;
        am = fmaxf(am, fmaxf(fmaxf(fabsf(t.x), fabsf(t.y)), fmaxf(fabsf(t.z), fabsf(t.w))));
        if (!(isfinite(t.x) && isfinite(t.y) && isfinite(t.z) && isfinite(t.w))) am = INFINITY;   // (fmaxf drops NaN)
      }
      am = block_max(am, sred);
      // (a non-finite block keeps its NaN/Inf visible: the step becomes NaN and every code saturates)
      const float step = am > 0.f ? am * (1.0f / 127.0f) : 1.0f;
      const float inv = 1.0f / step;
      uint32_t packed[EPT / 4];
#pragma unroll
      for (int j = 0; j < EPT; j += 4) {
        const int q0 = __float2int_rn(v[j] * inv), q1 = __float2int_rn(v[j + 1] * inv);
        const int q2 = __float2int_rn(v[j + 2] * inv), q3 = __float2int_rn(v[j + 3] * inv);
        packed[j / 4] = (static_cast<uint32_t>(q0) & 0xffu) | ((static_cast<uint32_t>(q1) & 0xffu) << 8) |
                        ((static_cast<uint32_t>(q2) & 0xffu) << 16) | ((static_cast<uint32_t>(q3) & 0xffu) << 24);
      }
      uint32_t* dst = reinterpret_cast<uint32_t*>(codes + static_cast<long long>(b) * QBLOCK + e0);
#pragma unroll
      for (int j = 0; j < EPT / 4; ++j) dst[j] = packed[j];
      if (threadIdx.x == 0) steps[b] = isfinite(am) ? step : am;
    }
  }
  if (a.payload_mode == 3) {
    // my gradients -> hivemind payload, one CTA per chunk
    // bucket sums as 64-bit fixed point of the in-bin residual: shared-memory float atomics are compare-and-swap loops
    // (the first version of this encoder cost 7 ms per step), integer ones are native
    __shared__ unsigned long long h_res[256];
    __shared__ int h_cnt[256];
    const int n_chunks = (a.n_blocks + CHUNK_BLOCKS - 1) / CHUNK_BLOCKS;
    uint8_t* pay = reinterpret_cast<uint8_t*>(a.payload[a.rank]);
    for (int ck = blockIdx.x; ck < n_chunks; ck += gridDim.x) {
      const int b0 = ck * CHUNK_BLOCKS, b1 = min(b0 + CHUNK_BLOCKS, a.n_blocks);
      const float* src = a.grad[a.rank] + static_cast<long long>(b0) * QBLOCK;
      uint8_t* dst = pay + static_cast<long long>(ck) * CHUNK_BYTES;
      if (a.chunk_codec[ck] == CODEC_FP16) {
        for (int b = b0; b < b1; ++b) {
          const float* s16 = src + static_cast<long long>(b - b0) * QBLOCK + e0;
          uint32_t o[EPT / 2];
#pragma unroll
          for (int j = 0; j < EPT; j += 2) {
            // finite values are clamped into the fp16 range (hivemind's Float16Compression); Inf / NaN stay what
            // they are so that the reducer's finite flag still sees them
            const float x0 = s16[j], x1 = s16[j + 1];
            const __half2 hv = __floats2half2_rn(isfinite(x0) ? fminf(fmaxf(x0, -65504.f), 65504.f) : x0,
                                                 isfinite(x1) ? fminf(fmaxf(x1, -65504.f), 65504.f) : x1);
            o[j / 2] = *reinterpret_cast<const uint32_t*>(&hv);
          }
          uint4* d4 = reinterpret_cast<uint4*>(dst + (static_cast<long long>(b - b0) * QBLOCK + e0) * 2);
          d4[0] = make_uint4(o[0], o[1], o[2], o[3]);
          d4[1] = make_uint4(o[4], o[5], o[6], o[7]);
        }
        continue;
      }
      const float n_el = static_cast<float>(b1 - b0) * QBLOCK;
      float sum = 0.f;
      for (int b = b0; b < b1; ++b) {
        const float* s16 = src + static_cast<long long>(b - b0) * QBLOCK + e0;
#pragma unroll
        for (int j = 0; j < EPT; ++j) sum += s16[j];
      }
      const float mu = block_sum(sum, sred) / n_el;
      float ssq_c = 0.f;
      for (int b = b0; b < b1; ++b) {
        const float* s16 = src + static_cast<long long>(b - b0) * QBLOCK + e0;
#pragma unroll
        for (int j = 0; j < EPT; ++j) { const float c = s16[j] - mu; ssq_c += c * c; }
      }
      const float sigma = sqrtf(block_sum(ssq_c, sred) / fmaxf(n_el - 1.f, 1.f));
      float scale = 6.0f * sigma / 256.0f;
      const bool finite = isfinite(scale) && isfinite(mu);
      if (!(scale > 0.f)) scale = 1.0f;
      const float inv = 1.0f / scale;
      h_res[threadIdx.x] = 0ull;
      h_cnt[threadIdx.x] = 0;
      __syncthreads();
      for (int b = b0; b < b1; ++b) {
        const float* s16 = src + static_cast<long long>(b - b0) * QBLOCK + e0;
        uint32_t o[EPT / 4];
#pragma unroll
        for (int j = 0; j < EPT; j += 4) {
          uint32_t w = 0u;
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const float x = s16[j + e];
            const float t = (x - mu) * inv;
            const int q = min(max(__float2int_rn(t) + 128, 0), 255);   // round half to even
            const float r = fminf(fmaxf(t - static_cast<float>(q - 128), -1e6f), 1e6f);   // residual in bin widths
            atomicAdd(&h_res[q], static_cast<unsigned long long>(__float2ll_rn(r * 1048576.f)));
            atomicAdd(&h_cnt[q], 1);
            w |= static_cast<uint32_t>(q) << (8 * e);
          }
          o[j / 4] = w;
        }
        *reinterpret_cast<uint4*>(dst + static_cast<long long>(b - b0) * QBLOCK + e0) = make_uint4(o[0], o[1], o[2], o[3]);
      }
      __syncthreads();
      // codebook = bucket means of the ORIGINAL values; a non-finite chunk stays visible as NaN to the reducer
      float* book = reinterpret_cast<float*>(dst + static_cast<long long>(b1 - b0) * QBLOCK);   // right behind the codes
      {
        const int cnt = h_cnt[threadIdx.x];
        const float mean_r = cnt > 0 ? static_cast<float>(static_cast<double>(static_cast<long long>(h_res[threadIdx.x])) *
                                                          (1.0 / 1048576.0) / cnt) : 0.f;
        const float val = cnt > 0 ? mu + (static_cast<float>(static_cast<int>(threadIdx.x) - 128) + mean_r) * scale : 0.f;
        book[threadIdx.x] = finite ? val : NAN;
      }
      __syncthreads();
    }
  }
  rank_barrier(a, a.epoch + 1, grid);  // phase 0: all gradients (and payloads) are complete

  // ---- phase 1: reduce-scatter into my shard -----------------------------------------------------
  float* g_local = a.grad[a.rank];
  float ssq = 0.f;
  bool bad = false;
  for (int b = a.blk0 + blockIdx.x; b < a.blk1; b += gridDim.x) {
    const long long base = static_cast<long long>(b) * QBLOCK + e0;
    float acc[EPT];
#pragma unroll
    for (int j = 0; j < EPT; ++j) acc[j] = 0.f;
    for (int r = 0; r < a.world; ++r) {
      const int src = (a.rank + r) % a.world;  // start with my own copy, spread peers across ranks
      if (a.payload_mode == 1) {
        const uint4* p = reinterpret_cast<const uint4*>(a.payload[src] + base);
#pragma unroll
        for (int j = 0; j < EPT; j += 8) {
          const uint4 u = p[j / 8];
          float2 f;
          f = unpack_bf16x2(u.x); acc[j] += f.x; acc[j + 1] += f.y;
          f = unpack_bf16x2(u.y); acc[j + 2] += f.x; acc[j + 3] += f.y;
          f = unpack_bf16x2(u.z); acc[j + 4] += f.x; acc[j + 5] += f.y;
          f = unpack_bf16x2(u.w); acc[j + 6] += f.x; acc[j + 7] += f.y;
        }
      } else if (a.payload_mode == 2) {
        const int8_t* codes = reinterpret_cast<const int8_t*>(a.payload[src]);
        const float step = reinterpret_cast<const float*>(codes + static_cast<long long>(a.n_blocks) * QBLOCK)[b];
        const uint32_t* p = reinterpret_cast<const uint32_t*>(codes + base);
#pragma unroll
        for (int j = 0; j < EPT; j += 4) {
          const uint32_t u = p[j / 4];
          acc[j] += step * static_cast<float>(static_cast<int8_t>(u & 0xffu));
          acc[j + 1] += step * static_cast<float>(static_cast<int8_t>((u >> 8) & 0xffu));
          acc[j + 2] += step * static_cast<float>(static_cast<int8_t>((u >> 16) & 0xffu));
          acc[j + 3] += step * static_cast<float>(static_cast<int8_t>(u >> 24));
        }
      } else if (a.payload_mode == 3) {
        const int ck = b / CHUNK_BLOCKS;
        const uint8_t* chunk = reinterpret_cast<const uint8_t*>(a.payload[src]) + static_cast<long long>(ck) * CHUNK_BYTES;
        const long long off = static_cast<long long>(b - ck * CHUNK_BLOCKS) * QBLOCK + e0;
        if (a.chunk_codec[ck] == CODEC_FP16) {
          const uint4* p = reinterpret_cast<const uint4*>(chunk + off * 2);
#pragma unroll
          for (int j = 0; j < EPT; j += 8) {
            const uint4 u = p[j / 8];
            const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&w[e]));
              acc[j + 2 * e] += f.x; acc[j + 2 * e + 1] += f.y;
            }
          }
        } else {
          // the peer's codebook of this chunk (1 KB over NVLink), then the block's codes
          __syncthreads();
          const int ck_blocks = min(CHUNK_BLOCKS, a.n_blocks - ck * CHUNK_BLOCKS);
          s_book[threadIdx.x] = reinterpret_cast<const float*>(chunk + static_cast<long long>(ck_blocks) * QBLOCK)[threadIdx.x];
          const uint4 u = *reinterpret_cast<const uint4*>(chunk + off);
          __syncthreads();
          const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
          for (int j = 0; j < EPT; ++j) acc[j] += s_book[(w[j >> 2] >> (8 * (j & 3))) & 0xffu];
        }
      } else {
        const float4* p = reinterpret_cast<const float4*>(a.grad[src] + base);
#pragma unroll
        for (int j = 0; j < EPT; j += 4) {
          const float4 v = p[j / 4];
          acc[j] += v.x; acc[j + 1] += v.y; acc[j + 2] += v.z; acc[j + 3] += v.w;
        }
      }
    }
#pragma unroll
    for (int j = 0; j < EPT; j += 4) {
      const float4 v = make_float4(acc[j] * a.hp.grad_scale, acc[j + 1] * a.hp.grad_scale,
                                   acc[j + 2] * a.hp.grad_scale, acc[j + 3] * a.hp.grad_scale);
      ssq += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
      bad |= !(isfinite(v.x) && isfinite(v.y) && isfinite(v.z) && isfinite(v.w));
      if (a.reduce_only) {
        for (int r = 0; r < a.world; ++r)
          *reinterpret_cast<float4*>(a.grad[(a.rank + r) % a.world] + base + j) = v;  // all-gather of the mean gradient
      } else {
        *reinterpret_cast<float4*>(g_local + base + j) = v;
      }
    }
  }
  if (a.reduce_only) {
    rank_barrier(a, a.epoch + 2, grid);
    rank_barrier(a, a.epoch + 3, grid);
    rank_barrier(a, a.epoch + 4, grid);  // keep the epoch arithmetic identical to a full step
    return;
  }
  ssq = block_sum(ssq, sred);
  if (threadIdx.x == 0) atomicAdd(&a.stats[0], ssq);
  if (bad) a.stats[1] = 1.f;
  __threadfence();
  grid.sync();

  // ---- phase 2: everyone learns the global norm and the global finite flag -----------------------
  if (blockIdx.x == 0 && threadIdx.x < a.world) {
    float* dst = a.xchg[threadIdx.x] + static_cast<long long>(a.rank) * a.xstride;
    dst[0] = a.stats[0];
    dst[1] = a.stats[1];
  }
  rank_barrier(a, a.epoch + 2, grid);
  float gsum = 0.f, gbad = 0.f;
  {
    const float* mine = a.xchg[a.rank];
    for (int r = 0; r < a.world; ++r) {
      gsum += mine[static_cast<long long>(r) * a.xstride];
      gbad += mine[static_cast<long long>(r) * a.xstride + 1];
    }
  }
  if (gtid == 0) { a.stats[2] = gsum; a.stats[3] = gbad; }
  const bool skip = gbad != 0.f;
  float clip = 1.0f;
  if (a.hp.use_clip) clip = fminf(1.0f, a.hp.max_grad_norm / (sqrtf(gsum) + 1e-6f));

  // ---- phase 3: moments + update direction on my shard --------------------------------------------
  if (!skip) {
    for (int b = a.blk0 + blockIdx.x; b < a.blk1; b += gridDim.x) {
      const LambBlock blk = a.blocks[b];
      const long long base = static_cast<long long>(b) * QBLOCK;
      const float wd = a.wd_per_tensor[blk.tensor];
      const float* p_local = a.param[a.rank];
      float gv[EPT], pv[EPT], m[EPT], v[EPT], u[EPT];
#pragma unroll
      for (int j = 0; j < EPT; j += 4) {
        const float4 x = *reinterpret_cast<const float4*>(g_local + base + e0 + j);
        const float4 y = *reinterpret_cast<const float4*>(p_local + base + e0 + j);
        gv[j] = x.x * clip; gv[j + 1] = x.y * clip; gv[j + 2] = x.z * clip; gv[j + 3] = x.w * clip;
        pv[j] = y.x; pv[j + 1] = y.y; pv[j + 2] = y.z; pv[j + 3] = y.w;
      }
      if (blk.quantized) {
        const float a1 = a.absmax1[blk.absmax_idx], a2 = a.absmax2[blk.absmax_idx];
        const uint4 u1 = *reinterpret_cast<const uint4*>(a.q1 + blk.state_off + e0);
        const uint4 u2 = *reinterpret_cast<const uint4*>(a.q2 + blk.state_off + e0);
        const uint8_t* b1 = reinterpret_cast<const uint8_t*>(&u1);
        const uint8_t* b2 = reinterpret_cast<const uint8_t*>(&u2);
#pragma unroll
        for (int j = 0; j < EPT; ++j) { m[j] = sc1[b1[j]] * a1; v[j] = sc2[b2[j]] * a2; }
      } else {
#pragma unroll
        for (int j = 0; j < EPT; j += 4) {
          const float4 x = *reinterpret_cast<const float4*>(a.f1 + blk.state_off + e0 + j);
          const float4 y = *reinterpret_cast<const float4*>(a.f2 + blk.state_off + e0 + j);
          m[j] = x.x; m[j + 1] = x.y; m[j + 2] = x.z; m[j + 3] = x.w;
          v[j] = y.x; v[j + 1] = y.y; v[j + 2] = y.z; v[j + 3] = y.w;
        }
      }
      float mx1 = 0.f, mx2 = 0.f, su = 0.f, sp = 0.f;
#pragma unroll
      for (int j = 0; j < EPT; ++j) {
        const bool ok = e0 + j < blk.valid;
        m[j] = ok ? a.hp.beta1 * m[j] + (1.f - a.hp.beta1) * gv[j] : 0.f;
        v[j] = ok ? a.hp.beta2 * v[j] + (1.f - a.hp.beta2) * gv[j] * gv[j] : 0.f;
        mx1 = fmaxf(mx1, fabsf(m[j]));
        mx2 = fmaxf(mx2, fabsf(v[j]));
        u[j] = ok ? m[j] / (sqrtf(v[j]) + a.hp.eps) + wd * pv[j] : 0.f;
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
        *reinterpret_cast<uint4*>(a.q1 + blk.state_off + e0) = o1;
        *reinterpret_cast<uint4*>(a.q2 + blk.state_off + e0) = o2;
        if (threadIdx.x == 0) { a.absmax1[blk.absmax_idx] = mx1; a.absmax2[blk.absmax_idx] = mx2; }
      } else {
#pragma unroll
        for (int j = 0; j < EPT; j += 4) {
          *reinterpret_cast<float4*>(a.f1 + blk.state_off + e0 + j) = make_float4(m[j], m[j + 1], m[j + 2], m[j + 3]);
          *reinterpret_cast<float4*>(a.f2 + blk.state_off + e0 + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
        }
      }
#pragma unroll
      for (int j = 0; j < EPT; j += 4)
        *reinterpret_cast<float4*>(g_local + base + e0 + j) = make_float4(u[j], u[j + 1], u[j + 2], u[j + 3]);
      su = block_sum(su, sred);
      sp = block_sum(sp, sred);
      if (threadIdx.x == 0) {
        atomicAdd(&a.tsums[2 * blk.tensor], su);
        atomicAdd(&a.tsums[2 * blk.tensor + 1], sp);
      }
    }
  }
  __threadfence();
  grid.sync();

  // ---- phase 4: per-tensor partial sums to every peer ---------------------------------------------
  for (int r = 0; r < a.world; ++r) {
    float* dst = a.xchg[r] + static_cast<long long>(a.rank) * a.xstride + 2;
    for (long long i = gtid; i < 2 * a.n_tensors; i += gthreads) dst[i] = a.tsums[i];
  }
  rank_barrier(a, a.epoch + 3, grid);

  // ---- phase 5: apply on my shard, write the result into every rank's buffers ---------------------
  for (int b = a.blk0 + blockIdx.x; b < a.blk1; b += gridDim.x) {
    const LambBlock blk = a.blocks[b];
    const long long base = static_cast<long long>(b) * QBLOCK + e0;
    float step = 0.f;
    if (!skip) {
      float su = 0.f, sp = 0.f;
      const float* mine = a.xchg[a.rank] + 2;
      for (int r = 0; r < a.world; ++r) {
        su += mine[static_cast<long long>(r) * a.xstride + 2 * blk.tensor];
        sp += mine[static_cast<long long>(r) * a.xstride + 2 * blk.tensor + 1];
      }
      const float un = sqrtf(su), wn = fminf(sqrtf(sp), a.hp.clamp_value);
      const float trust = (wn != 0.f && un != 0.f) ? wn / un : 1.0f;
      step = a.hp.lr * trust;
      if (threadIdx.x == 0 && a.trust_out) {
        a.trust_out[3 * blk.tensor] = wn;
        a.trust_out[3 * blk.tensor + 1] = un;
        a.trust_out[3 * blk.tensor + 2] = trust;
      }
    }
    float w[EPT];
#pragma unroll
    for (int j = 0; j < EPT; j += 4) {
      const float4 p = *reinterpret_cast<const float4*>(a.param[a.rank] + base + j);
      const float4 u = *reinterpret_cast<const float4*>(g_local + base + j);
      w[j] = p.x - step * u.x; w[j + 1] = p.y - step * u.y; w[j + 2] = p.z - step * u.z; w[j + 3] = p.w - step * u.w;
      if (a.zero_grad) *reinterpret_cast<float4*>(g_local + base + j) = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    for (int r = 0; r < a.world; ++r) {
      const int dst = (a.rank + r) % a.world;
#pragma unroll
      for (int j = 0; j < EPT; j += 4) {
        if (!skip) *reinterpret_cast<float4*>(a.param[dst] + base + j) = make_float4(w[j], w[j + 1], w[j + 2], w[j + 3]);
      }
#pragma unroll
      for (int j = 0; j < EPT; j += 8) {
        uint4 o;
        o.x = pack_bf16x2(w[j], w[j + 1]); o.y = pack_bf16x2(w[j + 2], w[j + 3]);
        o.z = pack_bf16x2(w[j + 4], w[j + 5]); o.w = pack_bf16x2(w[j + 6], w[j + 7]);
        *reinterpret_cast<uint4*>(a.shadow[dst] + base + j) = o;
      }
    }
  }
  // gradients outside my shard were consumed by their owners in phase 1 (all owners passed the phase-2 barrier)
  if (a.zero_grad) {
    for (int b = blockIdx.x; b < a.n_blocks; b += gridDim.x) {
      if (b >= a.blk0 && b < a.blk1) continue;
      float4* p = reinterpret_cast<float4*>(g_local + static_cast<long long>(b) * QBLOCK + e0);
#pragma unroll
      for (int j = 0; j < EPT / 4; ++j) p[j] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
  }
  rank_barrier(a, a.epoch + 4, grid);  // phase 6
}

}  // namespace b200

using namespace b200;

extern "C" int b200_group_lamb_step(const B200Group* d, void* stream) {
  if (d->world < 1 || d->world > MAX_RANKS) return -60;
  GroupArgs a;
  a.rank = d->rank; a.world = d->world;
  for (int r = 0; r < d->world; ++r) {
    a.grad[r] = reinterpret_cast<float*>(d->grad_ptrs[r]);
    a.param[r] = reinterpret_cast<float*>(d->param_ptrs[r]);
    a.shadow[r] = reinterpret_cast<__nv_bfloat16*>(d->shadow_ptrs[r]);
    a.payload[r] = reinterpret_cast<__nv_bfloat16*>(d->payload_ptrs[r]);
    a.sig[r] = reinterpret_cast<uint32_t*>(d->signal_ptrs[r]);
    a.xchg[r] = reinterpret_cast<float*>(d->xchg_ptrs[r]);
  }
  a.xstride = d->xstride; a.n_tensors = d->n_tensors; a.n_blocks = d->n_blocks;
  a.blk0 = d->blk0; a.blk1 = d->blk1; a.epoch = d->epoch; a.payload_mode = d->payload_mode; a.zero_grad = d->zero_grad;
  a.chunk_codec = reinterpret_cast<const uint8_t*>(d->chunk_codec);
  if (a.payload_mode == 3 && a.chunk_codec == nullptr) return -64;
  a.reduce_only = d->reduce_only;
  {
    static const int limit = [] { const char* e = getenv("DALLE_B200_GROUP_TIMEOUT_S"); const int v = e ? atoi(e) : 600; return v > 0 ? v : 600; }();
    a.spin_limit_s = limit;
  }
  a.q1 = reinterpret_cast<uint8_t*>(d->q1); a.q2 = reinterpret_cast<uint8_t*>(d->q2);
  a.f1 = d->f1; a.f2 = d->f2; a.absmax1 = d->absmax1; a.absmax2 = d->absmax2;
  a.code1 = d->code1; a.code2 = d->code2;
  a.blocks = reinterpret_cast<const LambBlock*>(d->blocks);
  a.wd_per_tensor = d->wd_per_tensor; a.tsums = d->tensor_sums; a.stats = d->stats; a.trust_out = d->trust_out;
  a.hp.beta1 = d->beta1; a.hp.beta2 = d->beta2; a.hp.eps = d->eps; a.hp.max_grad_norm = d->max_grad_norm;
  a.hp.clamp_value = d->clamp_value; a.hp.lr = d->lr; a.hp.grad_scale = d->grad_scale;
  a.hp.use_clip = d->max_grad_norm > 0.f;
  if (a.xstride < 2 + 2 * a.n_tensors) return -61;

  int dev = 0, sms = 0, per_sm = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, group_lamb_kernel, LAMB_THREADS, 0) != cudaSuccess ||
      per_sm < 1)
    return -62;
  if (per_sm > 4) per_sm = 4;
  void* args[] = {&a};
  cudaError_t e = cudaLaunchCooperativeKernel(reinterpret_cast<void*>(group_lamb_kernel), dim3(sms * per_sm),
                                              dim3(LAMB_THREADS), args, 0, reinterpret_cast<cudaStream_t>(stream));
  return e == cudaSuccess ? 0 : -63;
}
