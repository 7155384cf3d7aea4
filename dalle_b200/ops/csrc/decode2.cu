// Second-generation single-token decode path (workload of reference inference/run_inference.py:79-92).
//
// A decode step at batch <= 32 is a chain of ~320 tiny dependent operations whose cost is latency, not bytes: the
// first-generation path (decode.cu + the training GEMM at M = 32) spent 15-20 us per GEMM on 12-64 CTAs and ran the
// layer norms as 4-block kernels.  Here:
//
//   dec_gemm_kernel   "skinny" GEMM with the operands swapped: the WEIGHT rows sit on the 128-row M side of
//                     tcgen05.mma, the batch on an N = 32 tile, so no tensor work is padding.  K is split over the
//                     CTAs of a thread-block cluster (<= 8); each CTA has its whole K slice in flight at once (no
//                     ring), accumulates 128x32 in TMEM, and the partial tiles meet through distributed shared memory,
//                     so the epilogue (rotary, residual + layer-scale, GEGLU, bias) runs on complete sums - no atomics,
//                     no zeroed buffers, same bits every run.
//                     The layer norm + token shift that precede the QKV and FF1 GEMMs are that kernel's PROLOGUE: the
//                     producer of the residual stream leaves per-row (sum, sum of squares) behind, the consumer
//                     normalises its own K slice while it builds the swizzled B tile.  That removes both LN launches
//                     per layer (5 kernels per layer instead of 7).
//   dec_attn2_kernel  one CTA per (sample, head): the text keys and the layer type's local keys are contiguous in a
//                     sparsity-aware cache (text + one row / three rows / transposed image), so a handful of bulk
//                     copies bring all K and V rows to shared memory in ONE round trip; scores, softmax and PV then
//                     run out of shared memory.
//
// Every kernel is launched with programmatic dependent launch: it issues its weight loads (which depend on nothing)
// and allocates TMEM, then waits for its predecessor (griddepcontrol.wait) before touching activations - the weight
// stream of kernel n+1 overlaps the tail of kernel n, also inside the captured CUDA graph.
#include "sm100_cg2.cuh"
#include "b200_api.h"

#include <cstdio>

namespace b200 {

int make_tmap_bf16_2d(CUtensorMap* m, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_cols,
                      uint32_t box_rows);

constexpr int DG_NB_MAX = 64;      // batch tile = MMA N: 32 or 64 rows (template parameter NB)
constexpr int DG_THREADS = 128;
constexpr int DG_PART_LD = 132;    // floats per batch row of the partial tile (16-byte aligned, conflict-free writes)
constexpr int DG_MAX_KB = 8;       // K slice <= 512

enum { DG_QKV = 0, DG_RESID = 1, DG_GEGLU = 2, DG_LOGITS = 3 };
enum { DP_TMA = 0, DP_LN = 1 };

struct DecGemmArgs {
  int K, N, ks, ksplit;
  // LN + shift prologue
  const float* x;           // [32, K] fp32 residual stream (row stride K)
  const float* stats_in;    // [32][2]: sum, sum of squares of each row of x
  const float* ln_w;
  const float* ln_b;
  __nv_bfloat16* ring;      // [32][64 slots][K/2]: first half of the normalised channels of the last 64 positions
  const int* pos_ptr;
  int T, F;
  float eps;
  // epilogue
  const float* bias;
  const float* scale;
  float* xout;              // RESID: [32, N] fp32, updated in place
  float* stats_out;         // RESID: [32][2] accumulated with atomics (zeroed at the start of the step)
  __nv_bfloat16* out;       // bf16 outputs [32, ldo]
  int ldo;
  const float2* rot;        // [seq + 1][32] (cos, sin), position-major
  int q_cols;
  float q_scale;
};

__device__ __forceinline__ void griddep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void griddep_launch() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

__device__ __forceinline__ float4 ld_cluster_f4(uint32_t addr) {
  float4 v;
  asm volatile("ld.shared::cluster.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr) : "memory");
  return v;
}

template <int EPI, int PRO, int NB>
__global__ void __launch_bounds__(DG_THREADS)
dec_gemm_kernel(const __grid_constant__ CUtensorMap tmW, const __grid_constant__ CUtensorMap tmX, const DecGemmArgs a) {
  extern __shared__ uint8_t dg_smem_raw[];
  __shared__ __align__(8) uint64_t bar_k[DG_MAX_KB];
  __shared__ __align__(8) uint64_t bar_mma;
  __shared__ uint32_t tmem_slot;
  __shared__ float s_mean[NB], s_rstd[NB];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(dg_smem_raw) + 1023) & ~uintptr_t(1023));
  const int tid = threadIdx.x, warp = tid >> 5;
  const int nkb = a.ks >> 6;
  const uint32_t ksr = cluster_ctarank();
  const int mt = blockIdx.x / a.ksplit;
  const int k0 = static_cast<int>(ksr) * a.ks;
  uint8_t* sW = smem;                                  // nkb tiles of 128 rows x 128 bytes
  constexpr int XT = NB * 128;                         // bytes of one B tile (NB rows x 64 bf16)
  uint8_t* sX = smem + nkb * 16384;                    // nkb tiles of NB rows x 128 bytes
  float* part = reinterpret_cast<float*>(sX + nkb * XT);     // [NB][DG_PART_LD]

  if (warp == 0 && elect_one()) {
    for (int kb = 0; kb < nkb; ++kb) mbar_init(&bar_k[kb], 1);
    mbar_init(&bar_mma, 1);
    fence_barrier_init();
    // weights depend on nothing that runs before this kernel: the whole K slice goes in flight now
    for (int kb = 0; kb < nkb; ++kb) {
      mbar_expect_tx(&bar_k[kb], 16384 + (PRO == DP_TMA ? XT : 0));
      if (EPI == DG_GEGLU) {   // 64 value rows + the 64 gate rows of the same features
        tma_load_2d(sW + kb * 16384, &tmW, &bar_k[kb], k0 + kb * 64, mt * 64);
        tma_load_2d(sW + kb * 16384 + 8192, &tmW, &bar_k[kb], k0 + kb * 64, a.N + mt * 64);
      } else {
        tma_load_2d(sW + kb * 16384, &tmW, &bar_k[kb], k0 + kb * 64, mt * 128);
      }
    }
  }
  if (warp == 1) {
    tmem_alloc(&tmem_slot, NB);
    tmem_relinquish();
  }
  griddep_launch();
  griddep_wait();            // everything below may read what the previous kernels wrote
  if (PRO == DP_TMA) {
    if (warp == 0 && elect_one())
      for (int kb = 0; kb < nkb; ++kb) tma_load_2d(sX + kb * XT, &tmX, &bar_k[kb], k0 + kb * 64, 0);
  } else {
    // ---- layer norm + token shift of this CTA's K slice, written as the 128B-swizzled B tile ----
    // thread <-> one 16-byte chunk (8 channels) of the slice and every (128 / chunks)-th row: the norm weights are
    // loaded once per thread, a warp reads whole contiguous pieces of a row, and the loads of eight rows are in
    // flight together (the kernel is a latency chain: the first version walked 8 dependent round trips here).
    const int pos = __ldg(a.pos_ptr);
    if (tid < NB) {
      const float invK = 1.0f / a.K;
      const float2 st = *reinterpret_cast<const float2*>(a.stats_in + tid * 2);
      const float mean = st.x * invK;
      s_mean[tid] = mean;
      s_rstd[tid] = rsqrtf(fmaxf(st.y * invK - mean * mean, 0.f) + a.eps);
    }
    int src_a = -1, src_b = -1;     // positions feeding channels [0, K/4) and [K/4, K/2)
    if (pos < a.T) {
      src_a = src_b = pos - 1;      // (-1 at the first position)
    } else {
      const int p = pos - a.T, r = p / a.F, c = p - r * a.F;
      src_a = r >= 1 ? pos - a.F : -1;
      src_b = c >= 1 ? pos - 1 : -1;
    }
    const int halfK = a.K >> 1, quartK = a.K >> 2;
    const int nch = a.ks >> 3;                       // chunks per row of the slice: 8 .. 64
    const int ch = tid % nch, r0 = tid / nch;
    const int rstep = DG_THREADS / nch, nrows = NB / rstep;
    const int gc = k0 + ch * 8;
    const float4 w0 = __ldg(reinterpret_cast<const float4*>(a.ln_w + gc));
    const float4 w1 = __ldg(reinterpret_cast<const float4*>(a.ln_w + gc + 4));
    const float4 b0 = __ldg(reinterpret_cast<const float4*>(a.ln_b + gc));
    const float4 b1 = __ldg(reinterpret_cast<const float4*>(a.ln_b + gc + 4));
    const bool shifted = gc < halfK;
    const int src = gc < quartK ? src_a : src_b;
    __syncthreads();                                 // s_mean / s_rstd
    for (int j0 = 0; j0 < nrows; j0 += 8) {
      float4 xa[8], xb[8];
      uint4 rg[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int r = r0 + (j0 + u) * rstep;
        if (j0 + u < nrows) {
          xa[u] = *reinterpret_cast<const float4*>(a.x + static_cast<long long>(r) * a.K + gc);
          xb[u] = *reinterpret_cast<const float4*>(a.x + static_cast<long long>(r) * a.K + gc + 4);
          rg[u] = (shifted && src >= 0)
                      ? *reinterpret_cast<const uint4*>(a.ring + (static_cast<long long>(r) * 64 + (src & 63)) * halfK + gc)
                      : make_uint4(0u, 0u, 0u, 0u);
        }
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int r = r0 + (j0 + u) * rstep;
        if (j0 + u < nrows) {
          const float mean = s_mean[r], rstd = s_rstd[r];
          uint4 o;
          o.x = pack_bf16x2((xa[u].x - mean) * rstd * w0.x + b0.x, (xa[u].y - mean) * rstd * w0.y + b0.y);
          o.y = pack_bf16x2((xa[u].z - mean) * rstd * w0.z + b0.z, (xa[u].w - mean) * rstd * w0.w + b0.w);
          o.z = pack_bf16x2((xb[u].x - mean) * rstd * w1.x + b1.x, (xb[u].y - mean) * rstd * w1.y + b1.y);
          o.w = pack_bf16x2((xb[u].z - mean) * rstd * w1.z + b1.z, (xb[u].w - mean) * rstd * w1.w + b1.w);
          if (shifted) {
            if (mt == 0)
              *reinterpret_cast<uint4*>(a.ring + (static_cast<long long>(r) * 64 + (pos & 63)) * halfK + gc) = o;
            o = rg[u];
          }
          *reinterpret_cast<uint4*>(sX + (ch >> 3) * XT + r * 128 + (((ch & 7) ^ (r & 7)) << 4)) = o;
        }
      }
    }
    fence_proxy_async_smem();
  }
  __syncthreads();           // barriers initialised, TMEM address published, B tile complete
  tc_fence_after();
  const uint32_t tmem = tmem_slot;

  if (warp == 0 && elect_one()) {
    constexpr uint32_t idesc = make_idesc_bf16(128, NB, 0, 0);
    for (int kb = 0; kb < nkb; ++kb) {
      mbar_wait(&bar_k[kb], 0);
      tc_fence_after();
      const uint32_t wa = smem_u32(sW + kb * 16384), xa = smem_u32(sX + kb * XT);
#pragma unroll
      for (int k16 = 0; k16 < 4; ++k16)
        umma_bf16_ss(tmem, make_smem_desc_sw128(wa + k16 * 32, 16, 1024), make_smem_desc_sw128(xa + k16 * 32, 16, 1024),
                     idesc, (kb | k16) != 0);
    }
    umma_commit(&bar_mma);
  }
  // epilogue operands of this thread's first (usually only) group of features: requested now, so that their L2
  // round trip runs under the MMAs and the cluster exchange instead of after them
  const int b0 = tid >> 2, fq = tid & 3;
  const int own = (EPI == DG_GEGLU ? 64 : 128) / a.ksplit;
  const int n_first = mt * (EPI == DG_GEGLU ? 64 : 128) + static_cast<int>(ksr) * own + 4 * fq;
  float4 pre0 = make_float4(0.f, 0.f, 0.f, 0.f), pre1 = pre0, pre2 = pre0;
  if (n_first < a.N) {
    if (EPI == DG_GEGLU) {
      pre0 = __ldg(reinterpret_cast<const float4*>(a.bias + n_first));
      pre1 = __ldg(reinterpret_cast<const float4*>(a.bias + a.N + n_first));
    } else if (EPI == DG_RESID) {
      pre0 = __ldg(reinterpret_cast<const float4*>(a.bias + n_first));
      pre1 = __ldg(reinterpret_cast<const float4*>(a.scale + n_first));
      pre2 = *reinterpret_cast<const float4*>(a.xout + static_cast<long long>(b0) * a.N + n_first);
    } else if (EPI == DG_QKV) {
      const float2* rt = a.rot + static_cast<long long>(__ldg(a.pos_ptr)) * 32 + ((n_first & 63) >> 1);
      const float2 c0 = __ldg(rt), c1 = __ldg(rt + 1);
      pre0 = make_float4(c0.x, c0.y, c1.x, c1.y);
    }
  }
  __syncwarp();
  mbar_wait(&bar_mma, 0);
  tc_fence_after();
  {
#pragma unroll
    for (int jb = 0; jb < NB; jb += 32) {
      uint32_t v[32];
      tmem_ld32(tmem + (static_cast<uint32_t>(warp * 32) << 16) + jb, v);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 32; ++j) part[(jb + j) * DG_PART_LD + tid] = __uint_as_float(v[j]);   // part[batch][feature]
    }
  }
  tc_fence_before();
  cluster_sync_all();        // every CTA's partial tile is in its shared memory

  // ---- reduce over the K splits and finish: this CTA owns 128 / ksplit of the tile's features ----
#pragma unroll 1
  for (int bb = 0; bb < NB; bb += 32) {
  const int b = bb + b0;
  if (EPI == DG_GEGLU) {
    const int nj = 64 / a.ksplit;
    for (int g = 0; g < nj / 16; ++g) {
      const int j = static_cast<int>(ksr) * nj + 16 * g + 4 * fq;
      float4 val = make_float4(0.f, 0.f, 0.f, 0.f), gat = val;
      for (int c = 0; c < a.ksplit; ++c) {
        const float4 pv = ld_cluster_f4(map_to_cta(part + b * DG_PART_LD + j, c));
        const float4 pg = ld_cluster_f4(map_to_cta(part + b * DG_PART_LD + 64 + j, c));
        val.x += pv.x; val.y += pv.y; val.z += pv.z; val.w += pv.w;
        gat.x += pg.x; gat.y += pg.y; gat.z += pg.z; gat.w += pg.w;
      }
      const int n = mt * 64 + j;
      if (n < a.N) {
        const float4 bv = g == 0 ? pre0 : __ldg(reinterpret_cast<const float4*>(a.bias + n));
        const float4 bg = g == 0 ? pre1 : __ldg(reinterpret_cast<const float4*>(a.bias + a.N + n));
        auto act = [](float v, float gt) { return v * (0.5f * gt * (1.0f + erff(gt * 0.70710678118654752f))); };
        uint2 o;
        o.x = pack_bf16x2(act(val.x + bv.x, gat.x + bg.x), act(val.y + bv.y, gat.y + bg.y));
        o.y = pack_bf16x2(act(val.z + bv.z, gat.z + bg.z), act(val.w + bv.w, gat.w + bg.w));
        *reinterpret_cast<uint2*>(a.out + static_cast<long long>(b) * a.ldo + n) = o;
      }
    }
  } else {
    const int nf = 128 / a.ksplit;
    float s0 = 0.f, s1 = 0.f;
    for (int g = 0; g < nf / 16; ++g) {
      const int f = static_cast<int>(ksr) * nf + 16 * g + 4 * fq;
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int c = 0; c < a.ksplit; ++c) {
        const float4 pv = ld_cluster_f4(map_to_cta(part + b * DG_PART_LD + f, c));
        acc.x += pv.x; acc.y += pv.y; acc.z += pv.z; acc.w += pv.w;
      }
      const int n = mt * 128 + f;
      if (n >= a.N) continue;
      if (EPI == DG_QKV) {
        float2 c0 = make_float2(pre0.x, pre0.y), c1 = make_float2(pre0.z, pre0.w);
        if (g != 0) {
          const float2* rt = a.rot + static_cast<long long>(__ldg(a.pos_ptr)) * 32 + ((n & 63) >> 1);
          c0 = __ldg(rt);
          c1 = __ldg(rt + 1);
        }
        const float s = n < a.q_cols ? a.q_scale : 1.0f;
        uint2 o;
        o.x = pack_bf16x2((acc.x * c0.x - acc.y * c0.y) * s, (acc.y * c0.x + acc.x * c0.y) * s);
        o.y = pack_bf16x2((acc.z * c1.x - acc.w * c1.y) * s, (acc.w * c1.x + acc.z * c1.y) * s);
        *reinterpret_cast<uint2*>(a.out + static_cast<long long>(b) * a.ldo + n) = o;
      } else if (EPI == DG_RESID) {
        const float4 bi = g == 0 ? pre0 : __ldg(reinterpret_cast<const float4*>(a.bias + n));
        const float4 sc = g == 0 ? pre1 : __ldg(reinterpret_cast<const float4*>(a.scale + n));
        float4* xp = reinterpret_cast<float4*>(a.xout + static_cast<long long>(b) * a.N + n);
        float4 xo = (g == 0 && bb == 0) ? pre2 : *xp;
        xo.x += sc.x * (acc.x + bi.x); xo.y += sc.y * (acc.y + bi.y);
        xo.z += sc.z * (acc.z + bi.z); xo.w += sc.w * (acc.w + bi.w);
        *xp = xo;
        s0 += (xo.x + xo.y) + (xo.z + xo.w);
        s1 += (xo.x * xo.x + xo.y * xo.y) + (xo.z * xo.z + xo.w * xo.w);
      } else {   // DG_LOGITS (the bias is a slice of the tied vocabulary bias: no 16-byte alignment to rely on)
        const float4 bi = make_float4(__ldg(a.bias + n), __ldg(a.bias + n + 1), __ldg(a.bias + n + 2), __ldg(a.bias + n + 3));
        uint2 o;
        o.x = pack_bf16x2(acc.x + bi.x, acc.y + bi.y);
        o.y = pack_bf16x2(acc.z + bi.z, acc.w + bi.w);
        *reinterpret_cast<uint2*>(a.out + static_cast<long long>(b) * a.ldo + n) = o;
      }
    }
    if (EPI == DG_RESID && a.stats_out != nullptr) {
      s0 += __shfl_xor_sync(0xffffffffu, s0, 1); s1 += __shfl_xor_sync(0xffffffffu, s1, 1);
      s0 += __shfl_xor_sync(0xffffffffu, s0, 2); s1 += __shfl_xor_sync(0xffffffffu, s1, 2);
      if (fq == 0) {
        atomicAdd(a.stats_out + b * 2, s0);
        atomicAdd(a.stats_out + b * 2 + 1, s1);
      }
    }
  }
  }   // batch groups of 32 rows
  cluster_sync_all();        // nobody leaves while a neighbour still reads its partial tile
  if (warp == 1) tmem_dealloc(tmem, NB);
}


// -------------------------------------------------------------------------------------------------
// attention of one new token against the sparsity-aware cache
// -------------------------------------------------------------------------------------------------
constexpr int DA_THREADS = 256;
constexpr int DA_MAX_J = 38;                 // V rows per warp held in registers -> at most 8 * 38 = 304 keys
constexpr int DA_MAX_KEYS = 8 * DA_MAX_J;

struct DecAttnArgs {
  const __nv_bfloat16* qkv;   // [32, 3 * H * 64], rotary applied, q scaled
  __nv_bfloat16* kc;          // [B, H, slots, 64]
  __nv_bfloat16* vc;
  __nv_bfloat16* out;         // [32, H * 64]
  const int* pos_ptr;
  int B, H, T, F, conv_half, mode, slots;
};

// slot of image position (r, c) in the layer's cache: row layers keep one row, the convolution-like layer keeps
// conv_half + 1 rows, column layers keep the whole image TRANSPOSED (a column's keys are then contiguous)
__device__ __forceinline__ int da_slot(const DecAttnArgs& a, int r, int c) {
  if (a.mode == 1) return a.T + c;
  if (a.mode == 2) return a.T + c * a.F + r;
  return a.T + (r % (a.conv_half + 1)) * a.F + c;
}

__global__ void __launch_bounds__(DA_THREADS)
dec_attn2_kernel(const DecAttnArgs a) {
  extern __shared__ __align__(128) uint8_t da_smem[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ float sq[64];
  __shared__ float sp[DA_MAX_KEYS];
  __shared__ int sslot[DA_MAX_KEYS];
  __shared__ float sred[2][8];
  __shared__ float sacc[8][64];
  uint8_t* sK = da_smem;      // [n_keys][128 bytes]
  const int tid = threadIdx.x, w = tid >> 5, lane = tid & 31;
  const int b = blockIdx.x / a.H, h = blockIdx.x - b * a.H;
  const int pos = __ldg(a.pos_ptr);
  const int HD = a.H * 64;
  const long long cbase = (static_cast<long long>(b) * a.H + h) * a.slots * 64;
  const __nv_bfloat16* kc = a.kc + cbase;
  const __nv_bfloat16* vc = a.vc + cbase;

  // ---- the earlier keys this position attends to (the token itself is appended below) ----
  int n_prev, cur_slot;
  if (pos < a.T) {
    n_prev = pos;
    cur_slot = pos;
    for (int i = tid; i < n_prev; i += DA_THREADS) sslot[i] = i;
    if (tid == 0) {
      mbar_init(&bar, 1);
      fence_barrier_init();
      mbar_expect_tx(&bar, n_prev * 128);
      if (n_prev) bulk_copy_g2s(sK, kc, n_prev * 128, &bar);
    }
  } else {
    const int p = pos - a.T, r = p / a.F, c = p - r * a.F;
    cur_slot = da_slot(a, r, c);
    for (int i = tid; i < a.T; i += DA_THREADS) sslot[i] = i;
    if (a.mode == 1 || a.mode == 2) {
      const int nl = a.mode == 1 ? c : r;
      const int s0 = a.mode == 1 ? a.T : a.T + c * a.F;
      n_prev = a.T + nl;
      for (int i = tid; i < nl; i += DA_THREADS) sslot[a.T + i] = s0 + i;
      if (tid == 0) {
        mbar_init(&bar, 1);
        fence_barrier_init();
        mbar_expect_tx(&bar, n_prev * 128);
        bulk_copy_g2s(sK, kc, a.T * 128, &bar);
        if (nl) bulk_copy_g2s(sK + a.T * 128, kc + static_cast<long long>(s0) * 64, nl * 128, &bar);
      }
    } else {
      // window rows r - conv_half .. r, columns c - conv_half .. c + conv_half; in row r only the earlier columns
      const int c_lo = max(0, c - a.conv_half), c_hi = min(a.F - 1, c + a.conv_half);
      const int r_lo = max(0, r - a.conv_half);
      const int wfull = c_hi - c_lo + 1, wlast = c - c_lo;
      const int nl = (r - r_lo) * wfull + wlast;
      n_prev = a.T + nl;
      for (int i = tid; i < nl; i += DA_THREADS) {
        const int jr = i / wfull, jc = i - jr * wfull;
        sslot[a.T + i] = da_slot(a, r_lo + jr, c_lo + jc);
      }
      if (tid == 0) {
        mbar_init(&bar, 1);
        fence_barrier_init();
        mbar_expect_tx(&bar, n_prev * 128);
        bulk_copy_g2s(sK, kc, a.T * 128, &bar);
        for (int rr = r_lo; rr <= r; ++rr) {
          const int cnt = rr < r ? wfull : wlast;
          if (cnt)
            bulk_copy_g2s(sK + (a.T + (rr - r_lo) * wfull) * 128, kc + static_cast<long long>(da_slot(a, rr, c_lo)) * 64,
                          cnt * 128, &bar);
        }
      }
    }
  }
  __syncthreads();
  // ---- V rows of this warp (keys w, w + 8, ...): lane owns dims (2 lane, 2 lane + 1); all loads in flight now ----
  uint32_t vr[DA_MAX_J];
#pragma unroll
  for (int j = 0; j < DA_MAX_J; ++j) {
    const int i = w + 8 * j;
    vr[j] = i < n_prev ? *reinterpret_cast<const uint32_t*>(vc + static_cast<long long>(sslot[i]) * 64 + lane * 2) : 0u;
  }
  griddep_launch();
  griddep_wait();             // the QKV GEMM of this layer has finished
  const __nv_bfloat16* row = a.qkv + static_cast<long long>(b) * 3 * HD + h * 64;
  const uint32_t vcur = *reinterpret_cast<const uint32_t*>(row + 2 * HD + lane * 2);
  if (w == 0) {
    const float2 q = unpack_bf16x2(*reinterpret_cast<const uint32_t*>(row + lane * 2));
    sq[lane * 2] = q.x;
    sq[lane * 2 + 1] = q.y;
  } else if (w == 1) {
    const uint32_t kk = *reinterpret_cast<const uint32_t*>(row + HD + lane * 2);
    *reinterpret_cast<uint32_t*>(sK + n_prev * 128 + lane * 4) = kk;
    *reinterpret_cast<uint32_t*>(a.kc + cbase + static_cast<long long>(cur_slot) * 64 + lane * 2) = kk;
  } else if (w == 2) {
    *reinterpret_cast<uint32_t*>(a.vc + cbase + static_cast<long long>(cur_slot) * 64 + lane * 2) = vcur;
  }
  const int n_keys = n_prev + 1;
  __syncthreads();
  mbar_wait(&bar, 0);

  // ---- scores: thread <-> key ----
  float mx = -1e30f;
  for (int i = tid; i < n_keys; i += DA_THREADS) {
    const uint4* kr = reinterpret_cast<const uint4*>(sK + i * 128);
    float sc = 0.f;
#pragma unroll
    for (int jj = 0; jj < 8; ++jj) {
      const int c8 = (jj + i) & 7;         // rotate the chunk order: neighbouring rows start in different banks
      const uint4 u = kr[c8];
      const float* qq = &sq[c8 * 8];
      float2 f;
      f = unpack_bf16x2(u.x); sc += f.x * qq[0] + f.y * qq[1];
      f = unpack_bf16x2(u.y); sc += f.x * qq[2] + f.y * qq[3];
      f = unpack_bf16x2(u.z); sc += f.x * qq[4] + f.y * qq[5];
      f = unpack_bf16x2(u.w); sc += f.x * qq[6] + f.y * qq[7];
    }
    sc *= 1.4426950408889634f;
    sp[i] = sc;
    mx = fmaxf(mx, sc);
  }
  mx = warp_max(mx);
  if (lane == 0) sred[0][w] = mx;
  __syncthreads();
  mx = sred[0][0];
#pragma unroll
  for (int k = 1; k < 8; ++k) mx = fmaxf(mx, sred[0][k]);
  float l = 0.f;
  for (int i = tid; i < n_keys; i += DA_THREADS) {
    const float pr = exp2f(sp[i] - mx);
    sp[i] = pr;
    l += pr;
  }
  l = warp_sum(l);
  if (lane == 0) sred[1][w] = l;
  __syncthreads();
  l = 0.f;
#pragma unroll
  for (int k = 0; k < 8; ++k) l += sred[1][k];
  // ---- PV from the registers ----
  float a0 = 0.f, a1 = 0.f;
#pragma unroll
  for (int j = 0; j < DA_MAX_J; ++j) {
    const int i = w + 8 * j;
    if (i < n_keys) {
      const float2 vf = unpack_bf16x2(i == n_prev ? vcur : vr[j]);
      const float pr = sp[i];
      a0 += pr * vf.x;
      a1 += pr * vf.y;
    }
  }
  sacc[w][lane * 2] = a0;
  sacc[w][lane * 2 + 1] = a1;
  __syncthreads();
  if (w == 0) {
    float o0 = 0.f, o1 = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) { o0 += sacc[k][lane * 2]; o1 += sacc[k][lane * 2 + 1]; }
    const float inv = 1.0f / l;
    *reinterpret_cast<uint32_t*>(a.out + static_cast<long long>(b) * HD + h * 64 + lane * 2) = pack_bf16x2(o0 * inv, o1 * inv);
  }
}

// x1 = x2 = W[ids[b, pos]]; row statistics of the first layer norm; every other statistics slot of the step zeroed
__global__ void __launch_bounds__(1024)
dec_embed2_kernel(const long long* __restrict__ ids, int ids_stride, const float* __restrict__ W, float* __restrict__ x1,
                  float* __restrict__ x2, float* __restrict__ stats, int n_sites, const int* __restrict__ pos_ptr, int B,
                  int D, int NB) {
  const int lane = threadIdx.x & 31;
  for (int i = NB * 2 + threadIdx.x; i < n_sites * NB * 2; i += 1024) stats[i] = 0.f;
  for (int b = threadIdx.x >> 5; b < NB; b += 32) {   // one warp per row of the batch tile
    float s0 = 0.f, s1 = 0.f;
    const long long id = b < B ? ids[static_cast<long long>(b) * ids_stride + *pos_ptr] : 0;
    for (int c = lane * 4; c < D; c += 128) {
      // (rows past the batch are kept at zero: the residual epilogues update all rows of the tile)
      const float4 v = b < B ? __ldg(reinterpret_cast<const float4*>(W + id * D + c)) : make_float4(0.f, 0.f, 0.f, 0.f);
      *reinterpret_cast<float4*>(x1 + static_cast<long long>(b) * D + c) = v;
      *reinterpret_cast<float4*>(x2 + static_cast<long long>(b) * D + c) = v;
      s0 += (v.x + v.y) + (v.z + v.w);
      s1 += (v.x * v.x + v.y * v.y) + (v.z * v.z + v.w * v.w);
    }
    s0 = warp_sum(s0);
    s1 = warp_sum(s1);
    if (lane == 0) { stats[b * 2] = s0; stats[b * 2 + 1] = s1; }
  }
}

// ---- launchers ----------------------------------------------------------------------------------
template <typename Kern, typename... Args>
static int launch_pdl(Kern kern, dim3 grid, dim3 block, size_t smem, int cluster, bool pdl, cudaStream_t st,
                      Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute attr[2];
  int n = 0;
  if (cluster > 1) {
    attr[n].id = cudaLaunchAttributeClusterDimension;
    attr[n].val.clusterDim.x = cluster; attr[n].val.clusterDim.y = 1; attr[n].val.clusterDim.z = 1;
    ++n;
  }
  if (pdl) {
    attr[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[n].val.programmaticStreamSerializationAllowed = 1;
    ++n;
  }
  cfg.attrs = attr; cfg.numAttrs = n;
  return cudaLaunchKernelEx(&cfg, kern, args...) == cudaSuccess ? 0 : -82;
}

template <int EPI, int PRO, int NB>
static int launch_dec_gemm_nb(const CUtensorMap& tw, const CUtensorMap& tx, const DecGemmArgs& a, int n_tiles, bool pdl,
                              cudaStream_t st) {
  auto kern = dec_gemm_kernel<EPI, PRO, NB>;
  const size_t smem = static_cast<size_t>(a.ks / 64) * (16384 + NB * 128) + NB * DG_PART_LD * 4 + 1024;
  // largest request: a 512-wide K slice (8 k-blocks) - dynamic + ~1 KB of static shared memory must stay under 227 KB
  constexpr int kMaxSmem = DG_MAX_KB * (16384 + NB * 128) + NB * DG_PART_LD * 4 + 1024;
  if (smem > static_cast<size_t>(kMaxSmem)) return -87;
  static bool configured = false;
  if (!configured) {
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxSmem) != cudaSuccess) return -81;
    configured = true;
  }
  return launch_pdl(kern, dim3(n_tiles * a.ksplit), dim3(DG_THREADS), smem, a.ksplit, pdl, st, tw, tx, a);
}
template <int EPI, int PRO>
static int launch_dec_gemm(const CUtensorMap& tw, const CUtensorMap& tx, const DecGemmArgs& a, int n_tiles, bool pdl,
                           cudaStream_t st, int nb) {
  return nb == 64 ? launch_dec_gemm_nb<EPI, PRO, 64>(tw, tx, a, n_tiles, pdl, st)
                  : launch_dec_gemm_nb<EPI, PRO, 32>(tw, tx, a, n_tiles, pdl, st);
}

static bool g_dec_pdl = true;

}  // namespace b200

using namespace b200;

// kind: 0 QKV (LN prologue, rotary epilogue), 1 residual (TMA operand), 2 GEGLU (LN prologue), 3 logits (TMA operand)
extern "C" int b200_decode2_gemm(int kind, const void* W, int N, int K, const void* xin_bf16, const float* x,
                                 const float* stats_in, const float* ln_w, const float* ln_b, void* ring,
                                 const int* pos_ptr, int T, int F, float eps, const float* bias, const float* scale,
                                 float* xout, float* stats_out, void* out, int ldo, const void* rot, int q_cols,
                                 float q_scale, int nb, void* stream) {
  if (K % 64 || (nb != 32 && nb != 64)) return -80;
  int ksplit = 8;
  const int max_split = kind == DG_GEGLU ? 4 : 8;
  // enough CTAs to fill the machine, K slices of at most 512, feature ownership in groups of 16
  const int tiles = kind == DG_GEGLU ? (N + 63) / 64 : (N + 127) / 128;
  while (ksplit > 1 && (ksplit > max_split || K % (64 * ksplit) || (tiles * ksplit > 320 && K / ksplit < 512))) ksplit >>= 1;
  while (K / ksplit > 64 * DG_MAX_KB && ksplit < max_split) ksplit <<= 1;
  if (K % (64 * ksplit) || K / ksplit > 64 * DG_MAX_KB) return -80;
  DecGemmArgs a;
  a.K = K; a.N = N; a.ks = K / ksplit; a.ksplit = ksplit;
  a.x = x; a.stats_in = stats_in; a.ln_w = ln_w; a.ln_b = ln_b; a.ring = reinterpret_cast<__nv_bfloat16*>(ring);
  a.pos_ptr = pos_ptr; a.T = T; a.F = F; a.eps = eps;
  a.bias = bias; a.scale = scale; a.xout = xout; a.stats_out = stats_out;
  a.out = reinterpret_cast<__nv_bfloat16*>(out); a.ldo = ldo;
  a.rot = reinterpret_cast<const float2*>(rot); a.q_cols = q_cols; a.q_scale = q_scale;
  CUtensorMap tw, tx;
  const uint64_t w_rows = kind == DG_GEGLU ? 2ull * N : static_cast<uint64_t>(N);
  int rc = make_tmap_bf16_2d(&tw, W, w_rows, K, K, 64, kind == DG_GEGLU ? 64 : 128);
  if (rc) return rc;
  const bool ln = kind == DG_QKV || kind == DG_GEGLU;
  if (!ln) {
    rc = make_tmap_bf16_2d(&tx, xin_bf16, nb, K, K, 64, nb);
    if (rc) return rc;
  } else {
    tx = tw;
  }
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  switch (kind) {
    case DG_QKV: return launch_dec_gemm<DG_QKV, DP_LN>(tw, tx, a, tiles, g_dec_pdl, st, nb);
    case DG_RESID: return launch_dec_gemm<DG_RESID, DP_TMA>(tw, tx, a, tiles, g_dec_pdl, st, nb);
    case DG_GEGLU: return launch_dec_gemm<DG_GEGLU, DP_LN>(tw, tx, a, tiles, g_dec_pdl, st, nb);
    case DG_LOGITS: return launch_dec_gemm<DG_LOGITS, DP_TMA>(tw, tx, a, tiles, g_dec_pdl, st, nb);
    default: return -83;
  }
}

extern "C" int b200_decode2_attn(const void* qkv, void* kcache, void* vcache, void* out, const int* pos_ptr, int B, int H,
                                 int T, int F, int mode, int conv_kernel, int slots, void* stream) {
  DecAttnArgs a;
  a.qkv = reinterpret_cast<const __nv_bfloat16*>(qkv); a.kc = reinterpret_cast<__nv_bfloat16*>(kcache);
  a.vc = reinterpret_cast<__nv_bfloat16*>(vcache); a.out = reinterpret_cast<__nv_bfloat16*>(out);
  a.pos_ptr = pos_ptr; a.B = B; a.H = H; a.T = T; a.F = F; a.conv_half = conv_kernel / 2; a.mode = mode; a.slots = slots;
  if (mode < 1 || mode > 3) return -84;      // dense causal layers stay on the first-generation kernel
  const int local = mode == 3 ? (a.conv_half + 1) * (2 * a.conv_half + 1) : F;
  const int max_keys = T + local + 1;
  if (max_keys > DA_MAX_KEYS) return -85;
  const size_t smem = static_cast<size_t>(max_keys) * 128;
  static bool configured = false;
  if (!configured) {
    if (cudaFuncSetAttribute(dec_attn2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, DA_MAX_KEYS * 128) != cudaSuccess)
      return -81;
    configured = true;
  }
  return launch_pdl(dec_attn2_kernel, dim3(B * H), dim3(DA_THREADS), smem, 1, g_dec_pdl,
                    reinterpret_cast<cudaStream_t>(stream), a);
}

extern "C" int b200_decode2_embed(const long long* ids, int ids_stride, const float* W, float* x1, float* x2,
                                  float* stats, int n_sites, const int* pos_ptr, int B, int D, int nb, void* stream) {
  if ((nb != 32 && nb != 64) || B > nb || D % 128) return -86;
  dec_embed2_kernel<<<1, 1024, 0, reinterpret_cast<cudaStream_t>(stream)>>>(ids, ids_stride, W, x1, x2, stats, n_sites,
                                                                            pos_ptr, B, D, nb);
  return cudaGetLastError() == cudaSuccess ? 0 : -70;
}

extern "C" int b200_decode2_set_pdl(int on) { g_dec_pdl = on != 0; return 0; }
