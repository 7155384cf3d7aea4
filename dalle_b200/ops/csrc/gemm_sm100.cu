// bf16 GEMM family for sm_100a: TMA -> 128B-swizzled smem ring -> tcgen05.mma (single-thread issue)
// -> fp32 accumulators in TMEM (double buffered) -> tcgen05.ld epilogue with fused element-wise work.
//
// Replaces the library GEMMs the reference model reaches through ATen/cuBLAS (SURVEY §2.5 K4, K9, K11,
// K12, K13 and their autograd twins): forward (NT), data-gradient (NN, B read MN-major) and weight-gradient
// (TN, both operands MN-major, split-K with fp32 red.add into the shared .grad buffer - weight sharing
// means 16 layers accumulate into one gradient, reference task.py:65-66).
//
// Persistent kernel: grid = min(#work units, #SMs); each CTA has 10 warps:
//   warp 0  : TMA producer            (one lane)
//   warp 1  : TMEM allocator + MMA issuer (one lane)
//   warps 2-9: epilogue, warp w owns TMEM lanes 32*(w%4)..+31 (one output row per thread) and one half of the
//              256 accumulator columns
#include <cstring>
#include "sm100_cg2.cuh"
#include "b200_api.h"
#include <cstdlib>

#include <mutex>

namespace b200 {

// internal variant of B200_EPI_RESID for long reductions (K >= 2048): a fourth mainloop stage instead of a second
// staging set, bf16 side output by plain row stores
constexpr int EPI_RESID_DEEP = 7;
constexpr int BM = 128;
constexpr int BN = 256;
constexpr int BK = 64;
constexpr int A_STAGE_BYTES = BM * BK * 2;  // 16 KiB
constexpr int B_STAGE_BYTES = BN * BK * 2;  // 32 KiB
constexpr int STAGE_BYTES = A_STAGE_BYTES + B_STAGE_BYTES;
constexpr int OUT_TILE_BYTES = 32 * 64;                 // 32 rows x 32 bf16, 64B-swizzled TMA-store box
constexpr int OUT_TILES_PER_WARP = 3;
constexpr int OUT_STAGE_BYTES = 8 * OUT_TILES_PER_WARP * OUT_TILE_BYTES;  // 48 KiB for the 8 epilogue warps
// cg = 2 (CTA pairs): each CTA stages its own A tile and HALF of the B tile, 32 KB per stage instead of 48 KB, so
// the same shared memory holds a deeper ring
__host__ __device__ constexpr int gemm_stage_bytes(int cg) { return A_STAGE_BYTES + B_STAGE_BYTES / cg; }
__host__ __device__ constexpr int gemm_stages(int epi, int cg = 1) {
  return ((epi == B200_EPI_GEGLU_BWD || epi == B200_EPI_RESID) ? 3 : 4) * (cg == 2 ? 3 : 2) / 2;   // 3|4 -> 4|6
}
// Epilogue warps per CTA: 4 per TMEM lane quarter x column parts.  (16 warps were measured on the GEGLU backward
// epilogue: no gain.  An ablation showed why: that kernel's time is its 1.3 GB of epilogue loads + stores, whose
// in-flight volume is capped by the 64 KB of staging left beside the mainloop stages, not the epilogue math.)
__host__ __device__ constexpr int gemm_epi_warps(int epi) { return 8; }
// + TMA, MMA warps; the accumulating (wgrad) kernel has two more warps that can sum the A tiles (see GemmDev::a_colsum)
__host__ __device__ constexpr int gemm_sum_warps(int epi) { return epi == B200_EPI_ATOMIC_F32 ? 2 : 0; }
__host__ __device__ constexpr int gemm_threads(int epi) {
  return 64 + 32 * gemm_epi_warps(epi) + 32 * gemm_sum_warps(epi);
}
__host__ __device__ constexpr int gemm_out_tiles(int epi) {   // staging tiles per epilogue warp
  return epi == B200_EPI_GEGLU ? 2 : (epi == B200_EPI_GEGLU_BWD ? (gemm_epi_warps(epi) == 16 ? 2 : 4) : (epi == B200_EPI_RESID ? 5 : 2));
}
__host__ __device__ constexpr int gemm_out_stage_bytes(int epi) {
  return gemm_epi_warps(epi) * gemm_out_tiles(epi) * OUT_TILE_BYTES;
}
__host__ __device__ constexpr int gemm_smem_bytes(int epi, int cg = 1) {
  return gemm_stages(epi, cg) * gemm_stage_bytes(cg) + gemm_out_stage_bytes(epi) + 1024 /*align slack*/ +
         1024 /*barriers, column-sum scratch*/;
}
constexpr int TMEM_COLS = 512;  // 2 accumulator stages x 256 fp32 columns

struct GemmDev {
  int M, N, K;         // output M x N, reduction K
  int m_tiles, n_tiles, splits, kb_total;
  int tile_n_step;     // 256, or 128 for GEGLU pairing
  int second_off;      // row offset of the 2nd 128-row B box (128, or n_half for GEGLU)
  void* out;   long long ldo;
  void* out2;  long long ldo2;
  const float* bias;
  const float* scale;
  const float* res;  long long ldres;
  const float2* rot;  // TRANSPOSED table [32 pairs][rot_len] of (cos, sin): lanes = consecutive rows -> coalesced
  int seq_len;
  int rot_len;
  const int* pos_ptr;  // when set, every row uses this device-resident position (single-token decode)
  int q_cols;          // columns < q_cols are scaled by q_scale (rotary epilogue)
  float q_scale;
  const __nv_bfloat16* aux;  long long ldaux;
  int n_half;
  float alpha;
  float* a_colsum;     // TN only: a_colsum[m] += sum_k A[k, m], taken from the A tiles in shared memory
  int cg;              // 1, or 2: CTA pairs (cta_group::2) - a unit is then a PAIR of vertically adjacent tiles
};

// The i-th work unit of this CTA: output tile (m_blk, n_blk), k-blocks [kb0, kb1).  With CTA pairs the grid is walked
// pair by pair: both CTAs of a pair get the same unit index, the same n_blk and m_blk = 2 * pair_row + rank.
struct GemmUnit { int t, m_blk, n_blk, kb0, kb1; };
__device__ __forceinline__ bool gemm_unit(const GemmDev& g, int i, GemmUnit& un) {
  const int cta = g.cg == 2 ? (blockIdx.x >> 1) : blockIdx.x;
  const int ncta = g.cg == 2 ? (gridDim.x >> 1) : gridDim.x;
  const int m_rows = g.cg == 2 ? (g.m_tiles + 1) / 2 : g.m_tiles;   // tile rows (pairs of tile rows)
  const int u = cta + i * ncta;
  const int tiles = m_rows * g.n_tiles;
  if (u >= tiles * g.splits) return false;
  const int kb_per = (g.kb_total + g.splits - 1) / g.splits;
  // split-major order: the units of one wave share a K range, so every A / B panel of that range is pulled from
  // HBM once and then served by L2 (tile-major order re-read the large operand once per split wave)
  const int split = u / tiles;
  un.t = u - split * tiles;
  un.n_blk = un.t % g.n_tiles;
  un.m_blk = un.t / g.n_tiles;
  if (g.cg == 2) un.m_blk = un.m_blk * 2 + (blockIdx.x & 1);
  un.kb0 = split * kb_per;
  un.kb1 = min(un.kb0 + kb_per, g.kb_total);
  return true;
}

// GELU pieces shared by forward and backward epilogues.  Phi(g) through Abramowitz-Stegun 7.1.26
//   erf(x) = 1 - (a1 t + ... + a5 t^5) exp(-x^2),  t = 1 / (1 + p x),  x >= 0        (|error| <= 1.5e-7)
// with x = |g| / sqrt(2): the exponential is exp(-g^2 / 2), i.e. the same one the density needs, so the pair costs
// two MUFU ops (ex2, rcp) and ~12 FMA-pipe instructions; erff() + __expf() is ~40 and made these epilogues the
// bottleneck of the FF GEMMs (8 epilogue warps cannot hide that much dependent math behind a K=1024 mainloop).
struct GeluParts { float cdf, pdf; };
__device__ __forceinline__ float rcp_approx(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ GeluParts gelu_parts(float g) {
  const float ag = fabsf(g);
  const float e = ex2_approx(g * g * -0.72134752044448170f);           // exp(-g^2 / 2)
  const float t = rcp_approx(fmaf(ag, 0.3275911f * 0.70710678118654752f, 1.0f));
  float p = fmaf(t, 0.5f * 1.061405429f, 0.5f * -1.453152027f);        // 0.5 folded into the coefficients
  p = fmaf(t, p, 0.5f * 1.421413741f);
  p = fmaf(t, p, 0.5f * -0.284496736f);
  p = fmaf(t, p, 0.5f * 0.254829592f);
  const float tail = p * t * e;                                         // 1 - Phi(|g|)
  GeluParts r;
  r.cdf = g >= 0.f ? 1.0f - tail : tail;
  r.pdf = 0.3989422804014327f * e;
  return r;
}
// Two elements at a time on the packed fp32x2 FMA path of sm_100 (FFMA2 / FMUL2 / FADD2 halve the FMA-pipe issue
// slots of the polynomial; MUFU and the sign handling stay scalar).
struct GeluParts2 { float2 cdf, pdf; };
__device__ __forceinline__ float2 splat2(float x) { return make_float2(x, x); }
__device__ __forceinline__ GeluParts2 gelu_parts2(float2 g) {
  const float2 arg = __fmul2_rn(__fmul2_rn(g, g), splat2(-0.72134752044448170f));
  const float2 e = make_float2(ex2_approx(arg.x), ex2_approx(arg.y));
  const float2 den = __ffma2_rn(make_float2(fabsf(g.x), fabsf(g.y)), splat2(0.3275911f * 0.70710678118654752f),
                                splat2(1.0f));
  const float2 t = make_float2(rcp_approx(den.x), rcp_approx(den.y));
  float2 p = __ffma2_rn(t, splat2(0.5f * 1.061405429f), splat2(0.5f * -1.453152027f));
  p = __ffma2_rn(t, p, splat2(0.5f * 1.421413741f));
  p = __ffma2_rn(t, p, splat2(0.5f * -0.284496736f));
  p = __ffma2_rn(t, p, splat2(0.5f * 0.254829592f));
  const float2 tail = __fmul2_rn(__fmul2_rn(p, t), e);                 // 1 - Phi(|g|), in (0, 0.5]
  const float2 h = __ffma2_rn(tail, splat2(-1.0f), splat2(0.5f));      // Phi(|g|) - 0.5 >= 0
  GeluParts2 r;
  r.cdf = __fadd2_rn(splat2(0.5f), make_float2(copysignf(h.x, g.x), copysignf(h.y, g.y)));
  r.pdf = __fmul2_rn(e, splat2(0.3989422804014327f));
  return r;
}
__device__ __forceinline__ float gelu_cdf(float g) { return gelu_parts(g).cdf; }
__device__ __forceinline__ float gelu_erf(float x) { return x * gelu_cdf(x); }

__device__ __forceinline__ void st_bf16x8(__nv_bfloat16* p, const float* v) {
  uint4 u;
  u.x = pack_bf16x2(v[0], v[1]);
  u.y = pack_bf16x2(v[2], v[3]);
  u.z = pack_bf16x2(v[4], v[5]);
  u.w = pack_bf16x2(v[6], v[7]);
  *reinterpret_cast<uint4*>(p) = u;
}
__device__ __forceinline__ void ld_bf16x8(const __nv_bfloat16* p, float* v) {
  const uint4 u = *reinterpret_cast<const uint4*>(p);
  float2 f;
  f = unpack_bf16x2(u.x); v[0] = f.x; v[1] = f.y;
  f = unpack_bf16x2(u.y); v[2] = f.x; v[3] = f.y;
  f = unpack_bf16x2(u.z); v[4] = f.x; v[5] = f.y;
  f = unpack_bf16x2(u.w); v[6] = f.x; v[7] = f.y;
}
__device__ __forceinline__ void red_add_f32x4(float* p, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d)
               : "memory");
}

// Write 32 consecutive values of one row as bf16 (guarded on the right edge).
__device__ __forceinline__ void store_row_bf16(__nv_bfloat16* rowp, int col, int ncols, const float* v) {
  if (col + 32 <= ncols) {
#pragma unroll
    for (int j = 0; j < 32; j += 8) st_bf16x8(rowp + col + j, v + j);
  } else {
#pragma unroll
    for (int j = 0; j < 32; ++j)
      if (col + j < ncols) rowp[col + j] = __float2bfloat16(v[j]);
  }
}

template <int A_MN, int B_MN, int EPI, int CG>
__global__ void __launch_bounds__(gemm_threads(EPI), 1)
gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
            const __grid_constant__ CUtensorMap tmO, const __grid_constant__ CUtensorMap tmO2,
            const __grid_constant__ CUtensorMap tmX, const GemmDev g) {
  constexpr int STAGES = gemm_stages(EPI, CG);
  constexpr int B_BYTES = B_STAGE_BYTES / CG;              // B rows staged per CTA: 256, or 128 of a pair's 256
  constexpr int STAGE_BYTES = gemm_stage_bytes(CG);
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  const uint32_t pad = ((raw_addr + 1023u) & ~1023u) - raw_addr;
  uint8_t* smem = smem_raw + pad;
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + STAGES * A_STAGE_BYTES;
  uint8_t* smem_out = smem + STAGES * STAGE_BYTES;  // per-warp staging tiles for TMA stores (1024-aligned)
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + STAGES * STAGE_BYTES + gemm_out_stage_bytes(EPI));
  uint64_t* full_bar = bars;                 // [STAGES]
  uint64_t* empty_bar = bars + STAGES;       // [STAGES]
  uint64_t* tfull_bar = bars + 2 * STAGES;   // [2]
  uint64_t* tempty_bar = bars + 2 * STAGES + 2;  // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 4);
  uint64_t* aux_bar = bars + 2 * STAGES + 6;     // [epilogue warps][2 sets]: TMA loads of epilogue inputs (aux / residual)
  // CTA pairs: rank 0 (the leader) issues the MMAs of both; the pair's full / accumulator-empty barriers are the
  // leader's, the stage-empty and accumulator-full barriers are signalled in both CTAs by multicast commits
  const uint32_t crank = CG == 2 ? cluster_ctarank() : 0u;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tfull_bar[s], 1);
      mbar_init(&tempty_bar[s], 32 * gemm_epi_warps(EPI) * CG);   // (pairs: both CTAs' epilogue threads)
    }
    if (EPI == B200_EPI_GEGLU_BWD || EPI == B200_EPI_RESID || EPI == EPI_RESID_DEEP || EPI == B200_EPI_ATOMIC_F32)
      for (int s = 0; s < 2 * gemm_epi_warps(EPI); ++s) mbar_init(&aux_bar[s], 1);
    fence_barrier_init();
  }
  if (warp == 1) {
    if (CG == 2) { tmem_alloc_cg2(tmem_slot, TMEM_COLS); tmem_relinquish_cg2(); }
    else { tmem_alloc(tmem_slot, TMEM_COLS); tmem_relinquish(); }
  }
  tc_fence_before();
  __syncthreads();
  if (CG == 2) cluster_sync_all();   // the peer's barriers exist before anything signals them
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;


  if (warp == 0) {
    // elect.sync, not `lane == 0`: the compiler then knows the region runs on one thread and keeps TMA / MMA operands
    // in uniform registers; with `lane == 0` every cp.async.bulk.tensor and tcgen05.mma sat in its own
    // make-uniform loop (309 of them in this file's SASS)
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      GemmUnit un;
      for (int ui = 0; gemm_unit(g, ui, un); ++ui) {
        const int n_blk = un.n_blk, m_blk = un.m_blk;
        const int m0 = m_blk * BM, n0 = n_blk * g.tile_n_step;
        for (int kb = un.kb0; kb < un.kb1; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem_a + stage * A_STAGE_BYTES;
          uint8_t* sb = smem_b + stage * B_BYTES;
          const int k0 = kb * BK;
          if (CG == 2) {
            // my A tile and my half of the pair's B tile; the bytes of both CTAs complete the LEADER's barrier
            if (crank == 0) mbar_expect_tx(&full_bar[stage], 2 * STAGE_BYTES);
            const uint32_t fb = map_to_cta(&full_bar[stage], 0);
            if (A_MN == 0) {
              tma_load_2d_cg2(sa, &tmA, fb, k0, m0);
            } else {
              tma_load_2d_cg2(sa, &tmA, fb, m0, k0);
              tma_load_2d_cg2(sa + 8192, &tmA, fb, m0 + 64, k0);
            }
            if (B_MN == 0) {
              tma_load_2d_cg2(sb, &tmB, fb, k0, n0 + static_cast<int>(crank) * g.second_off);
            } else {
#pragma unroll
              for (int i = 0; i < 2; ++i)
                tma_load_2d_cg2(sb + i * 8192, &tmB, fb, n0 + 128 * static_cast<int>(crank) + 64 * i, k0);
            }
          } else {
            mbar_expect_tx(&full_bar[stage], STAGE_BYTES);
            if (A_MN == 0) {
              tma_load_2d(sa, &tmA, &full_bar[stage], k0, m0);  // box {64 k, 128 m}
            } else {
              tma_load_2d(sa, &tmA, &full_bar[stage], m0, k0);  // box {64 m, 64 k}
              tma_load_2d(sa + 8192, &tmA, &full_bar[stage], m0 + 64, k0);
            }
            if (B_MN == 0) {
              tma_load_2d(sb, &tmB, &full_bar[stage], k0, n0);  // box {64 k, 128 n}
              tma_load_2d(sb + 16384, &tmB, &full_bar[stage], k0, n0 + g.second_off);
            } else {
#pragma unroll
              for (int i = 0; i < 4; ++i) tma_load_2d(sb + i * 8192, &tmB, &full_bar[stage], n0 + 64 * i, k0);
            }
          }
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (crank == 0 && elect_one()) {   // (pairs: the leader issues for both CTAs)
      constexpr uint32_t idesc = make_idesc_bf16(BM * CG, BN, A_MN, B_MN);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      GemmUnit un;
      for (int ui = 0; gemm_unit(g, ui, un); ++ui) {
        const int kb0 = un.kb0, kb1 = un.kb1;
        mbar_wait(&tempty_bar[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + acc * BN;
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem_a + stage * A_STAGE_BYTES);
          const uint32_t sb = smem_u32(smem_b + stage * B_BYTES);
#pragma unroll
          for (int kk = 0; kk < BK / 16; ++kk) {
            // K-major: 8-row groups 1024 B apart, +32 B per 16-element k step inside the swizzle row.
            // MN-major: 64-element MN atoms 8192 B apart (LBO), 8-k-row groups 1024 B apart (SBO),
            //           +2048 B per 16 k rows.
            // (pairs: the descriptors name the leader's tiles; the peer's sit at the same offsets)
            const uint64_t ad = A_MN ? make_smem_desc_sw128(sa + kk * 2048, 8192, 1024)
                                     : make_smem_desc_sw128(sa + kk * 32, 0, 1024);
            const uint64_t bd = B_MN ? make_smem_desc_sw128(sb + kk * 2048, 8192, 1024)
                                     : make_smem_desc_sw128(sb + kk * 32, 0, 1024);
            if (CG == 2) umma_bf16_ss_cg2(tmem_d, ad, bd, idesc, (kb > kb0 || kk > 0) ? 1u : 0u);
            else umma_bf16_ss(tmem_d, ad, bd, idesc, (kb > kb0 || kk > 0) ? 1u : 0u);
          }
          if (CG == 2) umma_commit_cg2(&empty_bar[stage], 0x3);
          else umma_commit(&empty_bar[stage]);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        if (CG == 2) umma_commit_cg2(&tfull_bar[acc], 0x3);
        else umma_commit(&tfull_bar[acc]);
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1;
      }
    }
  } else if (gemm_sum_warps(EPI) > 0 && warp >= 2 + gemm_epi_warps(EPI)) {
    // ------------------------------- A column sums (TN wgrad + bias gradient) -------------------
    // dW = A^T B reads A = d(pre-activation) [K rows, M columns]; the matching bias gradient is the column sum of
    // the same matrix.  Two extra warps fetch A boxes ({64 k} x {64 m}, 8 KB) through their OWN double-buffered
    // TMA loads - L2 hits, the mainloop pulls the same panel at the same time - and sum them.  (Reading the
    // mainloop's stages instead put these warps on the stage-release path: +20 % kernel time even with empty
    // loops.)  The n_tiles units that share an A panel take its k-blocks in turn.
    if (A_MN == 1 && EPI == B200_EPI_ATOMIC_F32 && g.a_colsum) {
      const int box = warp - (2 + gemm_epi_warps(EPI));
      float* s_cs = reinterpret_cast<float*>(bars + 64) + box * 64;   // this warp's 64 column sums of a unit
      uint8_t* buf = smem_out + box * 16384;                          // 2 x 8 KB (the staging space is unused here)
      uint64_t* lbar = aux_bar + box * 2;
      s_cs[lane] = 0.f;
      s_cs[lane + 32] = 0.f;
      __syncwarp();
      uint32_t n_ld = 0, n_use = 0;   // boxes requested / consumed: buffer = n & 1, parity = (n >> 1) & 1
      GemmUnit un;
      for (int ui = 0; gemm_unit(g, ui, un); ++ui) {
        const int n_blk = un.n_blk, m_blk = un.m_blk;
        const int mcol = m_blk * BM + box * 64;
        // first k-block of this unit that falls to it: kb = n_blk (mod n_tiles)
        int kb_first = un.kb0 + ((n_blk - un.kb0 % g.n_tiles) + g.n_tiles) % g.n_tiles;
        float acc[2][8];
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[0][e] = acc[1][e] = 0.f;
        auto request = [&](int kb) {
          if (elect_one()) {
            mbar_expect_tx(&lbar[n_ld & 1], 8192);
            tma_load_2d(buf + (n_ld & 1) * 8192, &tmA, &lbar[n_ld & 1], mcol, kb * BK);
          }
          ++n_ld;
        };
        if (kb_first < un.kb1 && mcol < g.M) {
          request(kb_first);
          for (int kb = kb_first; kb < un.kb1; kb += g.n_tiles) {
            if (kb + g.n_tiles < un.kb1) request(kb + g.n_tiles);   // the other buffer: consumed one step ago
            mbar_wait(&lbar[n_use & 1], (n_use >> 1) & 1);
            // a lane reads the 16-byte cell (lane & 7) of rows (lane >> 3) + 4 i; the swizzle puts logical cell
            // (lane & 7) ^ (row & 7) there, and row & 7 takes two values per lane -> two sets of 8 partial sums
            const uint8_t* tile = buf + (n_use & 1) * 8192 + (lane >> 3) * 128 + ((lane & 7) << 4);
            ++n_use;
#pragma unroll
            for (int i = 0; i < 16; ++i) {
              const uint4 u = *reinterpret_cast<const uint4*>(tile + i * 512);
              float* a_ = acc[i & 1];
              float2 f;
              f = unpack_bf16x2(u.x); a_[0] += f.x; a_[1] += f.y;
              f = unpack_bf16x2(u.y); a_[2] += f.x; a_[3] += f.y;
              f = unpack_bf16x2(u.z); a_[4] += f.x; a_[5] += f.y;
              f = unpack_bf16x2(u.w); a_[6] += f.x; a_[7] += f.y;
            }
            __syncwarp();   // all lanes are done with the buffer before it is requested again
          }
          // unit done: fold the lanes' partials in shared memory, then one global atomic per column
          // (rows past K and columns past M were zero-filled by TMA)
#pragma unroll
          for (int par = 0; par < 2; ++par) {
            const int row7 = ((lane >> 3) + 4 * par) & 7;            // row & 7 of the rows behind acc[par]
            const int cell = (lane & 7) ^ row7;                      // logical 16-byte cell = 8 m columns
#pragma unroll
            for (int e = 0; e < 8; ++e) atomicAdd(&s_cs[cell * 8 + e], acc[par][e]);
          }
          __syncwarp();
#pragma unroll
          for (int e = 0; e < 2; ++e) {
            const int m = mcol + lane + 32 * e;
            const float v = s_cs[lane + 32 * e];
            s_cs[lane + 32 * e] = 0.f;
            if (m < g.M && v != 0.f) atomicAdd(g.a_colsum + m, v);
          }
          __syncwarp();
        }
      }
    }
  } else {
    // ------------------------------- epilogue warps -------------------------------------------
    const int q = warp & 3;
    constexpr int CPARTS = gemm_epi_warps(EPI) / 4;   // column parts of a tile (one warp per lane quarter each)
    constexpr int NCH = BN / 32 / CPARTS;             // 32-column chunks per warp and tile
    const int cpart = (warp - 2) >> 2;
    constexpr int NSETS = (gemm_epi_warps(EPI) == 16) ? 1 : 2;   // staging sets of the TMA-fed epilogues
    constexpr int TILES = gemm_out_tiles(EPI);
    uint8_t* stage = smem_out + (warp - 2) * TILES * OUT_TILE_BYTES;
    // bf16 outputs leave through smem: 32 rows x 32 columns per warp and chunk, 64B-swizzled, one TMA store each
    // (full 32-byte sectors instead of 32 scattered 16-byte row pieces per store instruction)
    auto stage_write = [&](int tile, const float* v32) {
      uint8_t* base = stage + tile * OUT_TILE_BYTES + lane * 64;
      const int sw = (lane >> 1) & 3;
#pragma unroll
      for (int p = 0; p < 4; ++p) {
        uint4 u;
        u.x = pack_bf16x2(v32[8 * p + 0], v32[8 * p + 1]);
        u.y = pack_bf16x2(v32[8 * p + 2], v32[8 * p + 3]);
        u.z = pack_bf16x2(v32[8 * p + 4], v32[8 * p + 5]);
        u.w = pack_bf16x2(v32[8 * p + 6], v32[8 * p + 7]);
        *reinterpret_cast<uint4*>(base + ((p ^ sw) << 4)) = u;
      }
    };
    // (every bulk-group operation of a warp - store, commit, wait - is issued under elect.sync, i.e. by the same
    // elected lane, and the compiler keeps the TMA operands in uniform registers)
    auto stage_reuse = [&]() {  // the previous chunk's stores must have read the staging tiles
      if (elect_one()) tma_store_wait_read();
      __syncwarp();
    };
    auto stage_flush = [&](const CUtensorMap* m, int tile, int x, int y) {
      if (elect_one()) tma_store_2d(m, stage + tile * OUT_TILE_BYTES, x, y);
    };
    auto stage_commit = [&]() {
      if (elect_one()) tma_store_commit();
    };
    uint32_t ring_n = 0;
    auto ring_store = [&](const CUtensorMap* m, const float* v32, int x, int y) {
      const int tile = ring_n % TILES;
      ++ring_n;
      if (elect_one()) tma_store_wait_read_pending<(TILES > 1 ? TILES - 1 : 0)>();
      __syncwarp();
      stage_write(tile, v32);
      fence_proxy_async_smem();
      __syncwarp();
      stage_flush(m, tile, x, y);
      stage_commit();
    };
    uint64_t* my_aux_bar = aux_bar + (warp - 2) * 2;
    uint32_t aux_n = 0;   // chunks consumed so far: set = aux_n & 1, mbarrier parity = (aux_n >> 1) & 1
    auto aux_load = [&](int set, int x, int y) {   // one lane: both aux tiles of a chunk -> staging set
      uint8_t* dst = stage + (2 * set) * OUT_TILE_BYTES;
      mbar_expect_tx(&my_aux_bar[set], 2 * OUT_TILE_BYTES);
      tma_load_2d(dst, &tmO2, &my_aux_bar[set], x, y);
      tma_load_2d(dst + OUT_TILE_BYTES, &tmO2, &my_aux_bar[set], g.n_half + x, y);
    };
    auto res_load = [&](int set, int x, int y) {   // one lane: the fp32 residual tile of a chunk -> staging set
      mbar_expect_tx(&my_aux_bar[set], 2 * OUT_TILE_BYTES);
      tma_load_2d(stage + (2 * set) * OUT_TILE_BYTES, &tmX, &my_aux_bar[set], x, y);
    };
    if (EPI == B200_EPI_GEGLU_BWD || EPI == B200_EPI_RESID || EPI == EPI_RESID_DEEP) {
      GemmUnit u0;   // the first chunk of the CTA's first unit is requested before the loop
      if (gemm_unit(g, 0, u0) && elect_one()) {
        const int x0 = u0.n_blk * g.tile_n_step + cpart * NCH * 32;
        if (x0 < g.N) {
          if (EPI == B200_EPI_RESID || EPI == EPI_RESID_DEEP) res_load(0, x0, u0.m_blk * BM + q * 32);
          else aux_load(0, x0, u0.m_blk * BM + q * 32);
        }
      }
    }
    int acc = 0;
    uint32_t acc_phase = 0;
    GemmUnit un;
    for (int ui = 0; gemm_unit(g, ui, un); ++ui) {
      const int n_blk = un.n_blk, m_blk = un.m_blk;
      const int row = m_blk * BM + q * 32 + lane;
      const int n0 = n_blk * g.tile_n_step;
      const bool row_ok = row < g.M;
      // the accumulator wait spans most of a mainloop: one sleeping poller per warp instead of 32 spinning lanes
      // (issue slots and power go to the MMA / TMA threads)
      if (lane == 0) mbar_wait_backoff(&tfull_bar[acc], acc_phase, 64);
      __syncwarp();
      tc_fence_after();
      const uint32_t taddr = tmem_base + acc * BN + (static_cast<uint32_t>(q * 32) << 16);

      if (EPI == B200_EPI_GEGLU) {
        // accumulator columns [0,128) = value half, [128,256) = gate half (same 128 output columns)
#pragma unroll 1
        for (int c = cpart * (NCH / 2); c < (cpart + 1) * (NCH / 2); ++c) {
          uint32_t ra[32], rg[32];
          tmem_ld32(taddr + c * 32, ra);
          tmem_ld32(taddr + 128 + c * 32, rg);
          tmem_ld_wait();
          const int col = n0 + c * 32;
          if (col < g.n_half) {  // warp-uniform
            float va[32], vg[32], o[32];
#pragma unroll
            for (int j = 0; j < 32; j += 2) {
              const bool c0 = g.bias && col + j < g.n_half, c1 = g.bias && col + j + 1 < g.n_half;
              const float2 ba = make_float2(c0 ? __ldg(g.bias + col + j) : 0.f, c1 ? __ldg(g.bias + col + j + 1) : 0.f);
              const float2 bg = make_float2(c0 ? __ldg(g.bias + g.n_half + col + j) : 0.f,
                                            c1 ? __ldg(g.bias + g.n_half + col + j + 1) : 0.f);
              const float2 a2 = __fadd2_rn(make_float2(__uint_as_float(ra[j]), __uint_as_float(ra[j + 1])), ba);
              const float2 g2 = __fadd2_rn(make_float2(__uint_as_float(rg[j]), __uint_as_float(rg[j + 1])), bg);
              const float2 o2 = __fmul2_rn(__fmul2_rn(a2, g2), gelu_parts2(g2).cdf);
              va[j] = a2.x; va[j + 1] = a2.y;
              vg[j] = g2.x; vg[j + 1] = g2.y;
              o[j] = o2.x; o[j + 1] = o2.y;
            }
            const int y = m_blk * BM + q * 32;
            // the warp's staging tiles form a ring with one commit group per store: a tile is rewritten as soon as
            // the store issued TILES stores ago has read it
            ring_store(&tmO, o, col, y);
            if (g.out2) {
              ring_store(&tmO2, va, col, y);
              ring_store(&tmO2, vg, g.n_half + col, y);
            }
          }
        }
      } else if (EPI == B200_EPI_GEGLU_BWD) {
        // acc = d(geglu out); aux = [value_pre | gate_pre] (bf16).  The aux tiles of a chunk arrive by TMA in the
        // warp's staging buffers (per-lane row loads would cost 32 L1 wavefronts per instruction), are turned into
        // (d value, d gate) in place, 8 columns at a time, and leave through TMA stores.  With two staging sets
        // the next chunk is fetched while this one is processed; with one set (16 epilogue warps) it is fetched as
        // soon as this chunk's stores have read the buffers, and the other warps of the sub-partition cover it.
        const int y = m_blk * BM + q * 32;
        auto prefetch_next = [&](int ci, int col) {
          int ncol = col + 32, ny = y;
          bool have = ci + 1 < NCH;
          if (!have) {
            GemmUnit nx;
            have = gemm_unit(g, ui + 1, nx);
            ncol = nx.n_blk * g.tile_n_step + cpart * NCH * 32;
            ny = nx.m_blk * BM + q * 32;
          }
          if (have && ncol < g.N) {
            stage_reuse();   // the stores that last read that set are done with it
            if (elect_one()) aux_load(aux_n % NSETS, ncol, ny);   // (aux_n already counts the current chunk)
          }
        };
#pragma unroll 1
        for (int ci = 0; ci < NCH; ++ci) {
          const int c = cpart * NCH + ci;
          const int col = n0 + c * 32;
          if (col >= g.N) continue;  // warp-uniform (never prefetched either)
          const int set = aux_n % NSETS;
          const uint32_t par = (aux_n / NSETS) & 1;
          ++aux_n;
          if (NSETS == 2) prefetch_next(ci, col);
          mbar_wait(&my_aux_bar[set], par);
          uint32_t r[32];
          tmem_ld32(taddr + c * 32, r);
          tmem_ld_wait();
          uint8_t* ta_ = stage + (2 * set) * OUT_TILE_BYTES + lane * 64;
          uint8_t* tg_ = ta_ + OUT_TILE_BYTES;
          const int sw = (lane >> 1) & 3;
#pragma unroll
          for (int p8 = 0; p8 < 4; ++p8) {
            uint4* pa = reinterpret_cast<uint4*>(ta_ + ((p8 ^ sw) << 4));
            uint4* pg = reinterpret_cast<uint4*>(tg_ + ((p8 ^ sw) << 4));
            const uint4 ua = *pa, ug = *pg;
            const uint32_t wa[4] = {ua.x, ua.y, ua.z, ua.w}, wg[4] = {ug.x, ug.y, ug.z, ug.w};
            uint32_t oa[4], og[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const int j = p8 * 8 + e * 2;
              const float2 a2 = unpack_bf16x2(wa[e]), g2 = unpack_bf16x2(wg[e]);
              // (rows past M: zero-filled A rows give a zero accumulator, and the TMA store clips them)
              const float2 v2 = make_float2(__uint_as_float(r[j]), __uint_as_float(r[j + 1]));
              const GeluParts2 gp = gelu_parts2(g2);
              const float2 x = __fmul2_rn(__fmul2_rn(v2, g2), gp.cdf);
              const float2 y2 = __fmul2_rn(__fmul2_rn(v2, a2), __ffma2_rn(g2, gp.pdf, gp.cdf));
              oa[e] = pack_bf16x2(x.x, x.y);
              og[e] = pack_bf16x2(y2.x, y2.y);
            }
            *pa = make_uint4(oa[0], oa[1], oa[2], oa[3]);
            *pg = make_uint4(og[0], og[1], og[2], og[3]);
          }
          fence_proxy_async_smem();
          __syncwarp();
          stage_flush(&tmO, 2 * set, col, y);
          stage_flush(&tmO, 2 * set + 1, g.n_half + col, y);
          stage_commit();
          if (NSETS == 1) prefetch_next(ci, col);
        }
      } else if (EPI == EPI_RESID_DEEP) {
        // One staging set per warp (a fourth mainloop stage is worth more than a second set: measured): the next
        // chunk's tile is requested as soon as this chunk's store has read the buffer.  The bf16 copy of
        // (acc + bias) goes out with plain row stores.
        const int y = m_blk * BM + q * 32;
#pragma unroll 1
        for (int ci = 0; ci < NCH; ++ci) {
          const int c = cpart * NCH + ci;
          const int col = n0 + c * 32;
          if (col >= g.N) continue;  // warp-uniform (never prefetched either)
          mbar_wait(&my_aux_bar[0], aux_n & 1);
          ++aux_n;
          uint32_t r[32];
          tmem_ld32(taddr + c * 32, r);
          tmem_ld_wait();
          uint8_t* trow_ = stage + lane * 128;
          float v[32];
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            float4 sc = make_float4(0.f, 0.f, 0.f, 0.f), bb = make_float4(0.f, 0.f, 0.f, 0.f);
            if (col + j < g.N) {   // N % 4 == 0 (checked by the host)
              sc = __ldg(reinterpret_cast<const float4*>(g.scale + col + j));
              if (g.bias) bb = __ldg(reinterpret_cast<const float4*>(g.bias + col + j));
            }
            float4* cell = reinterpret_cast<float4*>(trow_ + ((((j >> 2) ^ (lane & 7))) << 4));
            const float4 rr = *cell;
            v[j] = __uint_as_float(r[j]) + bb.x;
            v[j + 1] = __uint_as_float(r[j + 1]) + bb.y;
            v[j + 2] = __uint_as_float(r[j + 2]) + bb.z;
            v[j + 3] = __uint_as_float(r[j + 3]) + bb.w;
            *cell = make_float4(rr.x + g.alpha * sc.x * v[j], rr.y + g.alpha * sc.y * v[j + 1],
                                rr.z + g.alpha * sc.z * v[j + 2], rr.w + g.alpha * sc.w * v[j + 3]);
          }
          fence_proxy_async_smem();
          __syncwarp();
          stage_flush(&tmO, 0, col, y);
          stage_commit();
          if (g.out2 && row_ok)
            store_row_bf16(reinterpret_cast<__nv_bfloat16*>(g.out2) + (long long)row * g.ldo2, col, g.N, v);
          {   // request the next chunk's residual tile (of this tile, or the first of the CTA's next tile)
            int ncol = col + 32, ny = y;
            bool have = ci + 1 < NCH && ncol < g.N;   // (chunks past N are skipped, so look beyond them)
            if (!have) {
              GemmUnit nx;
              have = gemm_unit(g, ui + 1, nx);
              ncol = nx.n_blk * g.tile_n_step + cpart * NCH * 32;
              ny = nx.m_blk * BM + q * 32;
            }
            if (have && ncol < g.N) {
              stage_reuse();
              if (elect_one()) res_load(0, ncol, ny);
            }
          }
        }
      } else if (EPI == B200_EPI_RESID) {
        // out = res + alpha * scale[col] * (acc + bias[col]); optional bf16 copy of (acc + bias).
        // The fp32 residual tile of a chunk (32 x 32, 128B-swizzled) arrives by TMA in the warp's staging buffers
        // (two sets in ping-pong, fetched one chunk ahead), is updated in place and leaves through a TMA store:
        // full-line traffic instead of 32 row pieces per load/store instruction, and no load-behind-store
        // serialisation when `res` is `out` (the in-place residual stream).
        const int y = m_blk * BM + q * 32;
#pragma unroll 1
        for (int ci = 0; ci < NCH; ++ci) {
          const int c = cpart * NCH + ci;
          const int col = n0 + c * 32;
          {
            int ncol = col + 32, ny = y;
            bool have = ci + 1 < NCH;
            if (!have) {
              GemmUnit nx;
              have = gemm_unit(g, ui + 1, nx);
              ncol = nx.n_blk * g.tile_n_step + cpart * NCH * 32;
              ny = nx.m_blk * BM + q * 32;
            }
            if (have && ncol < g.N) {
              stage_reuse();
              if (elect_one()) res_load((aux_n + 1) & 1, ncol, ny);
            }
          }
          if (col >= g.N) continue;  // warp-uniform (never prefetched either)
          const int set = aux_n & 1;
          mbar_wait(&my_aux_bar[set], (aux_n >> 1) & 1);
          ++aux_n;
          uint32_t r[32];
          tmem_ld32(taddr + c * 32, r);
          tmem_ld_wait();
          uint8_t* trow_ = stage + (2 * set) * OUT_TILE_BYTES + lane * 128;
          float v[32];
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            float4 sc = make_float4(0.f, 0.f, 0.f, 0.f), bb = make_float4(0.f, 0.f, 0.f, 0.f);
            if (col + j < g.N) {   // N % 4 == 0 (checked by the host)
              sc = __ldg(reinterpret_cast<const float4*>(g.scale + col + j));
              if (g.bias) bb = __ldg(reinterpret_cast<const float4*>(g.bias + col + j));
            }
            float4* cell = reinterpret_cast<float4*>(trow_ + ((((j >> 2) ^ (lane & 7))) << 4));
            const float4 rr = *cell;
            v[j] = __uint_as_float(r[j]) + bb.x;
            v[j + 1] = __uint_as_float(r[j + 1]) + bb.y;
            v[j + 2] = __uint_as_float(r[j + 2]) + bb.z;
            v[j + 3] = __uint_as_float(r[j + 3]) + bb.w;
            *cell = make_float4(rr.x + g.alpha * sc.x * v[j], rr.y + g.alpha * sc.y * v[j + 1],
                                rr.z + g.alpha * sc.z * v[j + 2], rr.w + g.alpha * sc.w * v[j + 3]);
          }
          if (g.out2) stage_write(4, v);
          fence_proxy_async_smem();
          __syncwarp();
          stage_flush(&tmO, 2 * set, col, y);
          if (g.out2) stage_flush(&tmO2, 4, col, y);
          stage_commit();
        }
      } else {
#pragma unroll 1
        for (int c = cpart * NCH; c < (cpart + 1) * NCH; ++c) {
          const int col = n0 + c * 32;
          if (col >= g.N) break;  // warp-uniform
          uint32_t r[32];
          tmem_ld32(taddr + c * 32, r);
          tmem_ld_wait();
          constexpr bool kStaged = (EPI == B200_EPI_BF16 || EPI == B200_EPI_ROTARY);
          if (!kStaged && !row_ok) continue;   // staged epilogues keep the warp converged; TMA clips the tails
          const int y = m_blk * BM + q * 32;
          float v[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
          const bool full = col + 32 <= g.N;

          if (EPI == B200_EPI_BF16) {
            if (g.bias) {
#pragma unroll
              for (int j = 0; j < 32; ++j) v[j] += (col + j < g.N) ? __ldg(g.bias + col + j) : 0.f;
            }
            if (g.alpha != 1.0f) {
#pragma unroll
              for (int j = 0; j < 32; ++j) v[j] *= g.alpha;
            }
            ring_store(&tmO, v, col, y);   // two staging tiles: one store stays in flight behind the next chunk
          } else if (EPI == B200_EPI_ROTARY) {
            // interleaved-pair rotary on q, k and v; q additionally scaled (dalle-pytorch applies
            // rotary to v as well)
            const int pos = g.pos_ptr ? __ldg(g.pos_ptr) : row % g.seq_len;
            const float2* rt = g.rot + (long long)((col & 63) >> 1) * g.rot_len + pos;
            const float s = (col < g.q_cols) ? g.q_scale : 1.0f;
            float o[32];
#pragma unroll
            for (int j = 0; j < 32; j += 2) {
              const float2 cs = __ldg(rt + (long long)(j >> 1) * g.rot_len);
              const float x0 = v[j] + ((g.bias && col + j < g.N) ? __ldg(g.bias + col + j) : 0.f);
              const float x1 = v[j + 1] + ((g.bias && col + j + 1 < g.N) ? __ldg(g.bias + col + j + 1) : 0.f);
              o[j] = (x0 * cs.x - x1 * cs.y) * s;
              o[j + 1] = (x1 * cs.x + x0 * cs.y) * s;
            }
            ring_store(&tmO, o, col, y);
          } else if (EPI == B200_EPI_ATOMIC_F32) {
            float* op = reinterpret_cast<float*>(g.out) + (long long)row * g.ldo + col;
            if (g.scale) {
              // split-K form of the in-place residual epilogue (few-row GEMMs of the decode path):
              // out += alpha * scale[col] * (partial + bias[col] in the first split)
#pragma unroll
              for (int j = 0; j < 32; ++j) {
                if (col + j < g.N) {
                  const float bj = (g.bias && un.kb0 == 0) ? __ldg(g.bias + col + j) : 0.f;
                  v[j] = __ldg(g.scale + col + j) * (v[j] + bj);
                } else {
                  v[j] = 0.f;
                }
              }
            }
            if (full) {
#pragma unroll
              for (int j = 0; j < 32; j += 4)
                red_add_f32x4(op + j, g.alpha * v[j], g.alpha * v[j + 1], g.alpha * v[j + 2], g.alpha * v[j + 3]);
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j)
                if (col + j < g.N) atomicAdd(op + j, g.alpha * v[j]);
            }
          } else if (EPI == B200_EPI_F32) {
            float* op = reinterpret_cast<float*>(g.out) + (long long)row * g.ldo + col;
            if (full) {
#pragma unroll
              for (int j = 0; j < 32; j += 4)
                *reinterpret_cast<float4*>(op + j) =
                    make_float4(g.alpha * v[j], g.alpha * v[j + 1], g.alpha * v[j + 2], g.alpha * v[j + 3]);
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j)
                if (col + j < g.N) op[j] = g.alpha * v[j];
            }
          }
        }
      }
      tc_fence_before();
      if (CG == 2 && crank != 0) mbar_arrive_remote(map_to_cta(&tempty_bar[acc], 0));   // the leader's barrier
      else mbar_arrive(&tempty_bar[acc]);
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1;
    }
    if (elect_one()) tma_store_wait_all();  // staged stores must finish before the CTA's smem goes away
  }

  tc_fence_before();
  __syncthreads();
  if (CG == 2) cluster_sync_all();   // nobody leaves (or frees TMEM) while the peer can still signal or read
  if (warp == 1) {
    tc_fence_after();
    if (CG == 2) tmem_dealloc_cg2(tmem_base, TMEM_COLS);
    else tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

// -------------------------------------------------------------------------------------------------
// host side
// -------------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

// A CUtensorMap is a pure function of the encode arguments (it holds the address and the geometry, the driver keeps
// nothing behind it), and a training step asks for the same few hundred maps every step: keep the encoded maps in
// a per-thread direct-mapped table keyed on every argument, so that a steady-state launch costs a hash and a
// compare instead of a driver call per operand.
struct TmapKey {
  const void* ptr;
  uint64_t dim[5];
  uint64_t str[4];
  uint32_t box[5];
  int32_t dtype, rank, swizzle;
};
struct TmapSlot {
  TmapKey key;
  CUtensorMap map;
  bool used;
};
constexpr int kTmapSlots = 2048;

static uint64_t tmap_hash(const TmapKey& k) {
  const uint64_t* w = reinterpret_cast<const uint64_t*>(&k);
  uint64_t h = 0x9E3779B97F4A7C15ull;
  for (size_t i = 0; i < sizeof(TmapKey) / 8; ++i) { h ^= w[i]; h *= 0xFF51AFD7ED558CCDull; h ^= h >> 29; }
  return h;
}

static int encode_cached(CUtensorMap* m, CUtensorMapDataType dtype, int rank, const void* ptr, const uint64_t* dims,
                         const uint64_t* strides_bytes, const uint32_t* box, CUtensorMapSwizzle swizzle) {
  static_assert(sizeof(TmapKey) % 8 == 0, "hashed by 8-byte words");
  static thread_local TmapSlot* table = nullptr;
  if (table == nullptr) table = new TmapSlot[kTmapSlots]();
  TmapKey k;
  memset(&k, 0, sizeof(k));
  k.ptr = ptr; k.dtype = static_cast<int32_t>(dtype); k.rank = rank; k.swizzle = static_cast<int32_t>(swizzle);
  for (int i = 0; i < rank; ++i) { k.dim[i] = dims[i]; k.box[i] = box[i]; }
  for (int i = 0; i + 1 < rank; ++i) k.str[i] = strides_bytes[i];
  TmapSlot& slot = table[tmap_hash(k) & (kTmapSlots - 1)];
  if (slot.used && memcmp(&slot.key, &k, sizeof(k)) == 0) { *m = slot.map; return 0; }
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) return -1;
  cuuint64_t gdim[5];
  cuuint64_t gstr[4];
  cuuint32_t bx[5];
  cuuint32_t estr[5];
  for (int i = 0; i < rank; ++i) { gdim[i] = dims[i]; bx[i] = box[i]; estr[i] = 1; }
  for (int i = 0; i + 1 < rank; ++i) gstr[i] = strides_bytes[i];
  CUresult r = fn(m, dtype, rank, const_cast<void*>(ptr), gdim, gstr, bx, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return -2;
  slot.key = k; slot.map = *m; slot.used = true;
  return 0;
}

// Row-major bf16 matrix [rows, cols] with leading dimension ld (elements); box {box_cols, box_rows}.
int make_tmap_bf16_2d(CUtensorMap* m, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld,
                      uint32_t box_cols, uint32_t box_rows) {
  const uint64_t gdim[2] = {cols, rows}, gstr[1] = {ld * 2};
  const uint32_t box[2] = {box_cols, box_rows};
  return encode_cached(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, ptr, gdim, gstr, box, CU_TENSOR_MAP_SWIZZLE_128B);
}

// Output-side map for the staged epilogue: box {32 columns, 32 rows}, 64-byte swizzle.
int make_tmap_bf16_out(CUtensorMap* m, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld) {
  const uint64_t gdim[2] = {cols, rows}, gstr[1] = {ld * 2};
  const uint32_t box[2] = {32, 32};
  return encode_cached(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, ptr, gdim, gstr, box, CU_TENSOR_MAP_SWIZZLE_64B);
}

// fp32 tile map of the residual epilogue: box {32 columns, 32 rows} = 128-byte rows, 128-byte swizzle.
int make_tmap_f32_tile(CUtensorMap* m, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld) {
  const uint64_t gdim[2] = {cols, rows}, gstr[1] = {ld * 4};
  const uint32_t box[2] = {32, 32};
  return encode_cached(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, ptr, gdim, gstr, box, CU_TENSOR_MAP_SWIZZLE_128B);
}

int make_tmap_bf16_nd(CUtensorMap* m, const void* ptr, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                      const uint32_t* box) {
  if (rank < 1 || rank > 5) return -2;
  return encode_cached(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, rank, ptr, dims, strides_bytes, box,
                       CU_TENSOR_MAP_SWIZZLE_128B);
}

int num_sms() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
  }
  return n;
}

// set by b200_gemm_bf16 before launch<>() (host-side, per call; thread_local: host threads may call concurrently)
static thread_local CUtensorMap g_tm_out, g_tm_out2, g_tm_x;

template <int A_MN, int B_MN, int EPI, int CG>
static int launch_cg(const CUtensorMap& ta, const CUtensorMap& tb, const GemmDev& g, int grid, cudaStream_t st) {
  auto kern = gemm_kernel<A_MN, B_MN, EPI, CG>;
  constexpr int smem = gemm_smem_bytes(EPI, CG);
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return -3;
    configured = true;
  }
  if (CG == 1) {
    kern<<<grid, gemm_threads(EPI), smem, st>>>(ta, tb, g_tm_out, g_tm_out2, g_tm_x, g);
  } else {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(gemm_threads(EPI));
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    if (cudaLaunchKernelEx(&cfg, kern, ta, tb, g_tm_out, g_tm_out2, g_tm_x, g) != cudaSuccess) return -4;
  }
  return cudaGetLastError() == cudaSuccess ? 0 : -4;
}

template <int A_MN, int B_MN, int EPI>
static int launch(const CUtensorMap& ta, const CUtensorMap& tb, const GemmDev& g, int grid, cudaStream_t st) {
  if (g.cg == 2) return launch_cg<A_MN, B_MN, EPI, 2>(ta, tb, g, grid, st);
  return launch_cg<A_MN, B_MN, EPI, 1>(ta, tb, g, grid, st);
}

}  // namespace b200

using namespace b200;

extern "C" int b200_gemm_bf16(const B200Gemm* d_in, void* stream) {
  if (d_in->M <= 0 || d_in->N <= 0 || d_in->K <= 0) return 0;
  B200Gemm dd = *d_in;
  const B200Gemm* d = &dd;
  // Few-row in-place residual GEMMs (decode: M = batch) have only N / 256 output tiles: run them split-K through the
  // accumulating epilogue, which applies LayerScale and adds the bias in the first split.
  const bool resid_atomic = dd.epi == B200_EPI_RESID && dd.mode == B200_GEMM_NT && dd.M <= BM && dd.K >= 1024 &&
                            dd.res == dd.out && dd.out2 == nullptr && dd.ldres == dd.ldo;
  if (resid_atomic) { dd.epi = B200_EPI_ATOMIC_F32; dd.splits = 0; }
  const int a_mn = (d->mode == B200_GEMM_TN) ? 1 : 0;
  const int b_mn = (d->mode == B200_GEMM_NT) ? 0 : 1;
  if ((d->lda % 8) || (d->ldb % 8)) return -10;  // TMA needs 16-byte row strides
  if (reinterpret_cast<uintptr_t>(d->A) % 16 || reinterpret_cast<uintptr_t>(d->B) % 16) return -11;

  CUtensorMap ta, tb;
  int rc;
  if (!a_mn) rc = make_tmap_bf16_2d(&ta, d->A, d->M, d->K, d->lda, 64, 128);
  else       rc = make_tmap_bf16_2d(&ta, d->A, d->K, d->M, d->lda, 64, 64);
  if (rc) return rc;
  const uint64_t b_rows_total = d->epi == B200_EPI_GEGLU ? (uint64_t)2 * d->n_half : (uint64_t)d->N;
  if (!b_mn) rc = make_tmap_bf16_2d(&tb, d->B, b_rows_total, d->K, d->ldb, 64, 128);
  else       rc = make_tmap_bf16_2d(&tb, d->B, d->K, d->N, d->ldb, 64, 64);
  if (rc) return rc;

  GemmDev g;
  g.M = d->M; g.N = d->N; g.K = d->K;
  g.kb_total = (d->K + BK - 1) / BK;
  g.m_tiles = (d->M + BM - 1) / BM;
  if (d->epi == B200_EPI_GEGLU) {
    if (b_mn) return -12;
    g.tile_n_step = 128;
    g.second_off = d->n_half;
    g.n_tiles = (d->n_half + 127) / 128;
  } else {
    g.tile_n_step = BN;
    g.second_off = 128;
    g.n_tiles = (d->N + BN - 1) / BN;
  }
  int splits = d->splits;
  // CTA pairs (cta_group::2).  Measured on the TTT shapes: pairs win where the mainloop is the bound (long K, many
  // tiles: FF2 329 -> 268 us, wgrad FF2 374 -> 294 us, dgrad / wgrad QKV -10 %, GEGLU forward -4 %) and lose where the
  // epilogue is (the leader's MMA then waits for the slower of TWO epilogues: GEGLU backward, K = 1024 residual).
  // B200_GEMM_CG2 = bit mask of (1 << epi) overrides the policy (experiments).
  int cg = 1;
  if (d->M > BM && num_sms() % 2 == 0) {
    static int mask = -2;
    if (mask == -2) { const char* e = getenv("B200_GEMM_CG2"); mask = e ? atoi(e) : -1; }
    bool pair;
    if (mask >= 0) {
      pair = (mask >> d->epi) & 1;
    } else {
      const int n_tiles_all = g.m_tiles * g.n_tiles;
      switch (d->epi) {
        case B200_EPI_GEGLU: pair = true; break;
        case B200_EPI_RESID: pair = d->K >= 2048; break;
        case B200_EPI_BF16: pair = d->K >= 2048; break;
        case B200_EPI_ATOMIC_F32: pair = n_tiles_all >= 64 && !resid_atomic; break;
        default: pair = false;
      }
    }
    if (pair) cg = 2;
  }
  const int tiles = (cg == 2 ? (g.m_tiles + 1) / 2 : g.m_tiles) * g.n_tiles;   // units per split (pairs count once)
  if (splits <= 0) {
    splits = 1;
    if (d->epi == B200_EPI_ATOMIC_F32) {
      // split K so that the (tile, split) units fill whole waves of the persistent grid: efficiency of a choice
      // = units / (waves * #SMs); every extra split costs one more accumulating epilogue per tile
      // (an equal-share stream-K cut of the k-block space measured 1.5x SLOWER here: CTAs at unrelated K
      // offsets lose the L2 reuse of the A / B panels that lock-step units share)
      const int sms = num_sms() / cg;
      const int min_kb = resid_atomic ? 4 : 32;   // k-blocks per unit
      const int max_splits = g.kb_total / min_kb > 0 ? (g.kb_total / min_kb < 16 ? g.kb_total / min_kb : 16) : 1;
      float best = -1.f;
      for (int sp = 1; sp <= max_splits; ++sp) {
        const int units = tiles * sp;
        const int waves = (units + sms - 1) / sms;
        const float score = static_cast<float>(units) / (static_cast<float>(waves) * sms) - 0.004f * sp;
        if (score > best) { best = score; splits = sp; }
      }
    }
  }
  if (splits > 1 && d->epi != B200_EPI_ATOMIC_F32) return -13;
  if (splits > g.kb_total) splits = g.kb_total;
  // every split must own >= 1 k-block
  {
    const int kb_per = (g.kb_total + splits - 1) / splits;
    splits = (g.kb_total + kb_per - 1) / kb_per;
  }
  g.splits = splits;
  g.out = d->out; g.ldo = d->ldo;
  g.out2 = d->out2; g.ldo2 = d->ldo2;
  g.bias = d->bias; g.scale = d->scale;
  g.res = d->res; g.ldres = d->ldres;
  g.rot = reinterpret_cast<const float2*>(d->rot); g.seq_len = d->seq_len > 0 ? d->seq_len : 1;
  g.rot_len = d->rot_len;
  g.pos_ptr = reinterpret_cast<const int*>(d->pos_ptr);
  g.q_cols = d->q_cols; g.q_scale = d->q_scale;
  g.aux = reinterpret_cast<const __nv_bfloat16*>(d->aux); g.ldaux = d->ldaux;
  g.n_half = d->n_half;
  g.alpha = d->alpha;
  g.cg = cg;
  if (d->a_colsum && !(a_mn && d->epi == B200_EPI_ATOMIC_F32)) return -16;
  if (d->epi == B200_EPI_ATOMIC_F32 && !resid_atomic) g.scale = nullptr;
  g.a_colsum = d->a_colsum;

  // output maps for the TMA-store epilogues
  g_tm_out = ta;
  g_tm_out2 = ta;  // placeholders when unused
  g_tm_x = ta;
  if (d->epi == B200_EPI_RESID) {
    if ((d->ldo % 4) || (d->ldres % 4) || (d->N % 4) || reinterpret_cast<uintptr_t>(d->out) % 16 ||
        reinterpret_cast<uintptr_t>(d->res) % 16 || !d->scale)
      return -15;
    rc = make_tmap_f32_tile(&g_tm_out, d->out, d->M, d->N, d->ldo);
    if (rc) return rc;
    rc = make_tmap_f32_tile(&g_tm_x, d->res, d->M, d->N, d->ldres);
    if (rc) return rc;
    if (d->out2) {
      if ((d->ldo2 % 8) || reinterpret_cast<uintptr_t>(d->out2) % 16) return -15;
      rc = make_tmap_bf16_out(&g_tm_out2, d->out2, d->M, d->N, d->ldo2);
      if (rc) return rc;
    }
  }
  {
    const bool staged = d->epi == B200_EPI_BF16 || d->epi == B200_EPI_ROTARY || d->epi == B200_EPI_GEGLU ||
                        d->epi == B200_EPI_GEGLU_BWD;
    if (staged) {
      if ((d->ldo % 8) || reinterpret_cast<uintptr_t>(d->out) % 16) return -15;
      const uint64_t ocols = d->epi == B200_EPI_GEGLU ? (uint64_t)d->n_half
                             : (d->epi == B200_EPI_GEGLU_BWD ? (uint64_t)2 * d->n_half : (uint64_t)d->N);
      rc = make_tmap_bf16_out(&g_tm_out, d->out, d->M, ocols, d->ldo);
      if (rc) return rc;
      if (d->epi == B200_EPI_GEGLU_BWD) {   // the aux operand is read through the second map
        if ((d->ldaux % 8) || reinterpret_cast<uintptr_t>(d->aux) % 16) return -15;
        rc = make_tmap_bf16_out(&g_tm_out2, d->aux, d->M, (uint64_t)2 * d->n_half, d->ldaux);
        if (rc) return rc;
      }
      if (d->epi == B200_EPI_GEGLU && d->out2) {
        if ((d->ldo2 % 8) || reinterpret_cast<uintptr_t>(d->out2) % 16) return -15;
        rc = make_tmap_bf16_out(&g_tm_out2, d->out2, d->M, (uint64_t)2 * d->n_half, d->ldo2);
        if (rc) return rc;
      }
    }
  }
  const int units = tiles * splits;
  const int grid = cg * (units < num_sms() / cg ? units : num_sms() / cg);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);

#define B200_DISPATCH_EPI(AMN, BMN)                                                              \
  switch (d->epi) {                                                                              \
    case B200_EPI_BF16: return launch<AMN, BMN, B200_EPI_BF16>(ta, tb, g, grid, st);             \
    case B200_EPI_F32: return launch<AMN, BMN, B200_EPI_F32>(ta, tb, g, grid, st);               \
    case B200_EPI_ATOMIC_F32: return launch<AMN, BMN, B200_EPI_ATOMIC_F32>(ta, tb, g, grid, st); \
    default: break;                                                                              \
  }
  if (d->mode == B200_GEMM_NT) {
    switch (d->epi) {
      case B200_EPI_ROTARY: return launch<0, 0, B200_EPI_ROTARY>(ta, tb, g, grid, st);
      case B200_EPI_RESID:
        if (d->K >= 2048) return launch<0, 0, EPI_RESID_DEEP>(ta, tb, g, grid, st);
        return launch<0, 0, B200_EPI_RESID>(ta, tb, g, grid, st);
      case B200_EPI_GEGLU: return launch<0, 0, B200_EPI_GEGLU>(ta, tb, g, grid, st);
      default: break;
    }
    B200_DISPATCH_EPI(0, 0)
  } else if (d->mode == B200_GEMM_NN) {
    if (d->epi == B200_EPI_GEGLU_BWD) return launch<0, 1, B200_EPI_GEGLU_BWD>(ta, tb, g, grid, st);
    B200_DISPATCH_EPI(0, 1)
  } else if (d->mode == B200_GEMM_TN) {
    B200_DISPATCH_EPI(1, 1)
  }
#undef B200_DISPATCH_EPI
  return -14;
}
