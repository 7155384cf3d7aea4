// Sparse DALL-E attention forward, fourth generation (text, axial_row, axial_col).
//
// Lessons of the two earlier forwards (profiles/attn3_fwd_timeline_r2.txt): the softmax is bound by the latency of
// its TMEM round trips and exponentials, not by the tensor cores, and a hand-scheduled two-stage pipeline inside one
// big CTA hides that latency worse than the hardware scheduler does with many small CTAs.  So this kernel makes the
// CTA as light as possible and lets FOUR of them share an SM:
//   * 64-key chunks: S is 64 TMEM columns, read ONCE into registers (one wait per chunk);
//   * P goes back into the columns of S as bf16 and is the A operand of the P V MMA straight from TMEM
//     (no shared-memory copy of P: 48 KB of the second generation's 100 KB);
//   * O stays in TMEM for the whole block (64 columns); the reference maximum is only raised - and O rescaled in
//     TMEM - when a chunk exceeds it by more than 2^8, exact either way;
//   * 128 TMEM columns, ~49 KB of shared memory (Q, two K and two V chunk buffers), <= 128 registers.
// One CTA = one (batch, head, 128-query block); warp 0 elects the thread that issues TMA and MMA.
#include "attn_common.cuh"
#include "b200_api.h"

#include <cstdio>
#include <cstdlib>

namespace b200 {

int make_tmap_bf16_2d(CUtensorMap* m, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_cols,
                      uint32_t box_rows);
int attn_make_maps(const void* qkv, const AttnGeom& g, int mode, CUtensorMap* t128, CUtensorMap* ttail,
                   CUtensorMap* taux);

constexpr int AF4_THREADS = 128;
constexpr int AF4_SMEM = 16384 + 4 * 8192 + 128 + 1024;
constexpr uint32_t AF4_TM_O = 64;
constexpr float AF4_RESCALE_LOG2 = 8.0f;

struct AttnFwd4Args {
  AttnGeom g;
  const void* qkv;
  __nv_bfloat16* out;
  float* lse;
  int ld_out;
  int n_text_chunks;   // ceil(T_pad / 64)
  int text_shift;      // 1: the text query blocks start at token 1 (see attn4_fwd), token 0 is written directly
};

template <int IMG_MODE, bool TEXT>   // image pattern (ATT_ROW / ATT_COL); TEXT: this CTA is a text query block
__device__ __forceinline__ void attn_fwd4_body(const CUtensorMap& tm128, const CUtensorMap& tm64, const CUtensorMap& tmTail, const CUtensorMap& tmCol,
                 const AttnFwd4Args& a, const int blk) {
  constexpr int MODE = TEXT ? ATT_TEXT : IMG_MODE;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + (((raw_addr + 1023u) & ~1023u) - raw_addr);
  const AttnGeom& g = a.g;
  uint8_t* sQ = smem;               // 16 KB; reused as the output staging tile
  uint8_t* sK = smem + 16384;       // 2 x 8 KB
  uint8_t* sV = sK + 16384;         // 2 x 8 KB
  uint64_t* bars = reinterpret_cast<uint64_t*>(sV + 16384);
  uint64_t* bar_q = bars + 0;
  uint64_t* bar_k = bars + 1;       // [2]
  uint64_t* bar_v = bars + 3;       // [2]
  uint64_t* bar_s = bars + 5;
  uint64_t* bar_p = bars + 6;
  uint64_t* bar_o = bars + 7;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 8);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int h = blockIdx.y, b = blockIdx.z;
  const int q0 = blk * 128 + (MODE == ATT_TEXT ? a.text_shift : 0);
  const int HD = g.H * 64;
  const int row_base = b * g.seq;
  const int lines = 128 / g.F;          // grid lines per query block
  const int lines64 = 64 / g.F;         // grid lines per 64-key local chunk

  // chunk schedule: text chunks of 64 keys (the last one T_pad mod 64), then two local chunks for image blocks
  const int n_text = (MODE == ATT_TEXT) ? min((min(q0 + 128, g.T) + 63) / 64, a.n_text_chunks) : a.n_text_chunks;
  const int n_chunks = n_text + (MODE == ATT_TEXT ? 0 : 2);

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tm64);
    mbar_init(bar_q, 1);
    mbar_init(&bar_k[0], 1); mbar_init(&bar_k[1], 1);
    mbar_init(&bar_v[0], 1); mbar_init(&bar_v[1], 1);
    mbar_init(bar_s, 1); mbar_init(bar_p, 128); mbar_init(bar_o, 1);
    fence_barrier_init();
  }
  if (warp == 0) {
    tmem_alloc(tmem_slot, 128);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  auto chunk_keys = [&](int c) -> int { return c < n_text ? min(64, g.T_pad - 64 * c) : 64; };
  // K (sect 0) or V (sect 1) of chunk c into buffer c & 1
  auto load_kv = [&](int c, int sect) {
    uint64_t* full = sect == 0 ? &bar_k[c & 1] : &bar_v[c & 1];
    uint8_t* dst = (sect == 0 ? sK : sV) + (c & 1) * 8192;
    const int col = (sect == 0 ? HD : 2 * HD) + h * 64;
    const int nk = chunk_keys(c);
    mbar_expect_tx(full, static_cast<uint32_t>(nk) * 128u);
    if (c < n_text) {
      tma_load_2d(dst, nk == 64 ? &tm64 : &tmTail, full, col, row_base + 64 * c);
    } else {
      const int j = c - n_text;
      if (MODE == ATT_COL) {
        for (int l = 0; l < lines64; ++l)
          tma_load_4d(dst + l * g.F * 128, &tmCol, full, col, blk * lines + j * lines64 + l, 0, b);
      } else {
        tma_load_2d(dst, &tm64, full, col, row_base + g.T + q0 + 64 * j);
      }
    }
  };
  const uint32_t q_addr = smem_u32(sQ), k_addr = smem_u32(sK), v_addr = smem_u32(sV);
  auto issue_s = [&](int c) {
    const int nk = chunk_keys(c);
    mbar_wait(&bar_k[c & 1], (c >> 1) & 1);
    tc_fence_after();
    const uint32_t idesc = make_idesc_bf16(128, nk, 0, 0);
    const uint64_t dq = make_smem_desc_sw128(q_addr, 0, 1024);
    const uint64_t dk = make_smem_desc_sw128(k_addr + (c & 1) * 8192, 0, 1024);
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) umma_bf16_ss(tmem, dq + 2 * kk, dk + 2 * kk, idesc, kk > 0 ? 1u : 0u);
    umma_commit(bar_s);
  };
  auto issue_pv = [&](int c) {
    const int nk = chunk_keys(c);
    mbar_wait(bar_p, c & 1);
    mbar_wait(&bar_v[c & 1], (c >> 1) & 1);
    tc_fence_after();
    constexpr uint32_t idesc = make_idesc_bf16(128, 64, 0, 1);
    const uint64_t dv = make_smem_desc_sw128(v_addr + (c & 1) * 8192, 8192, 1024);
    for (int k16 = 0; k16 < nk / 16; ++k16)
      umma_bf16_ts(tmem + AF4_TM_O, tmem + k16 * 8, dv + 128 * k16, idesc, (c > 0 || k16 > 0) ? 1u : 0u);
    if (c == n_chunks - 1) umma_commit(bar_o);
  };

  if (warp == 0 && elect_one()) {
    mbar_expect_tx(bar_q, 16384);
    if (MODE == ATT_COL) {
      for (int l = 0; l < lines; ++l) tma_load_4d(sQ + l * g.F * 128, &tmCol, bar_q, h * 64, blk * lines + l, 0, b);
    } else {
      tma_load_2d(sQ, &tm128, bar_q, h * 64, row_base + ((MODE == ATT_TEXT) ? q0 : g.T + q0));
    }
    load_kv(0, 0);
    load_kv(0, 1);
    if (n_chunks > 1) load_kv(1, 0);
    mbar_wait(bar_q, 0);
    issue_s(0);
  }
  __syncwarp();

  // ---- one query row per thread
  const int i = warp * 32 + lane;
  const uint32_t trow = tmem + (static_cast<uint32_t>(warp * 32) << 16);
  int tpos;
  uint32_t lmask = 0u;
  if (MODE == ATT_TEXT) {
    tpos = q0 + i;
  } else {
    const int ip = img_query_pos<MODE>(g, blk, i);
    const int ir = ip / g.F, ic = ip - ir * g.F;
    tpos = g.T + ip;
    // local keys of this row: its own line (row or column) up to itself, inside the warp's 32-key window
    const int lo = max(i - (MODE == ATT_ROW ? ic : ir), warp * 32) - warp * 32;
    lmask = (0xFFFFFFFFu >> (31 - lane)) & (0xFFFFFFFFu << lo);
  }
  const bool q_valid = (MODE == ATT_TEXT) ? (tpos < g.T) : (tpos < g.seq);
  const int text_limit = q_valid ? ((MODE == ATT_TEXT) ? tpos + 1 : g.T) : 0;
  if (!q_valid) lmask = 0u;
  float m_used = -1e30f, l_run = 0.f;

  for (int c = 0; c < n_chunks; ++c) {
    const int nk = chunk_keys(c);
    const bool is_text = c < n_text;
    mbar_wait(bar_s, c & 1);
    tc_fence_after();
    // S(c) is complete, and with it every earlier MMA: K buffer c & 1 and V buffer (c + 1) & 1 are free
    if (warp == 0 && elect_one()) {
      if (c + 2 < n_chunks) load_kv(c + 2, 0);
      if (c + 1 < n_chunks) load_kv(c + 1, 1);
    }
    __syncwarp();

    float cmax = -1e30f, lsum0 = 0.f, lsum1 = 0.f;
    if (is_text && nk == 64) {
      // ---- full text chunk: 64 scores in registers, one TMEM read
      uint32_t r0[32], r1[32];
      tmem_ld32(trow, r0);
      tmem_ld32(trow + 32, r1);
      tmem_ld_wait();
      const int nv = text_limit - 64 * c;
      const bool all_vis = __all_sync(0xFFFFFFFFu, nv >= 64);
      if (all_vis) {
#pragma unroll
        for (int j = 0; j < 32; ++j) cmax = fmaxf(cmax, fmaxf(__uint_as_float(r0[j]), __uint_as_float(r1[j])));
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          cmax = fmaxf(cmax, j < nv ? __uint_as_float(r0[j]) : -1e30f);
          cmax = fmaxf(cmax, j + 32 < nv ? __uint_as_float(r1[j]) : -1e30f);
        }
      }
      const float m_c = cmax * kLog2e;
      const bool need = q_valid && (m_c > m_used + AF4_RESCALE_LOG2);
      float factor = 1.0f;
      if (need) { factor = fast_exp2(m_used - m_c); m_used = m_c; l_run *= factor; }
      if (c > 0 && __any_sync(0xFFFFFFFFu, need)) {
#pragma unroll 1
        for (int half = 0; half < 2; ++half) {
          uint32_t o[32];
          tmem_ld32(trow + AF4_TM_O + half * 32, o);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 32; ++j) o[j] = __float_as_uint(__uint_as_float(o[j]) * factor);
          tmem_st32(trow + AF4_TM_O + half * 32, o);
        }
      }
      if (all_vis) {
#pragma unroll
        for (int j = 0; j < 32; j += 2) {
          const float e0 = fast_exp2(__uint_as_float(r0[j]) * kLog2e - m_used);
          const float e1 = fast_exp2(__uint_as_float(r0[j + 1]) * kLog2e - m_used);
          const float e2 = fast_exp2(__uint_as_float(r1[j]) * kLog2e - m_used);
          const float e3 = fast_exp2(__uint_as_float(r1[j + 1]) * kLog2e - m_used);
          lsum0 += e0 + e2; lsum1 += e1 + e3;
          r0[j >> 1] = pack_bf16x2(e0, e1);
          r1[j >> 1] = pack_bf16x2(e2, e3);
        }
      } else {
#pragma unroll
        for (int j = 0; j < 32; j += 2) {
          const float e0 = j < nv ? fast_exp2(__uint_as_float(r0[j]) * kLog2e - m_used) : 0.f;
          const float e1 = j + 1 < nv ? fast_exp2(__uint_as_float(r0[j + 1]) * kLog2e - m_used) : 0.f;
          const float e2 = j + 32 < nv ? fast_exp2(__uint_as_float(r1[j]) * kLog2e - m_used) : 0.f;
          const float e3 = j + 33 < nv ? fast_exp2(__uint_as_float(r1[j + 1]) * kLog2e - m_used) : 0.f;
          lsum0 += e0 + e2; lsum1 += e1 + e3;
          r0[j >> 1] = pack_bf16x2(e0, e1);
          r1[j >> 1] = pack_bf16x2(e2, e3);
        }
      }
      tmem_st16_h(trow, r0);        // every S column is in registers: P may overwrite them
      tmem_st16_h(trow + 16, r1);
    } else {
      // ---- short text tail (16-column groups) or a local chunk (this warp's 32-key window, if it lies in it)
      uint32_t r[32];
      uint32_t vm = 0u;           // visibility of r[0..31]
      int ncols = 0;              // S columns this thread looks at: [col0, col0 + ncols)
      int col0 = 0;
      if (is_text) {
        ncols = nk;               // < 64, multiple of 16; handled 32 at a time below
      } else if (((warp >> 1) == (c - n_text))) {
        ncols = 32; col0 = (warp & 1) * 32; vm = lmask;
      }
      // (text tail: up to 48 columns; two rounds of at most 32)
#pragma unroll 1
      for (int g0 = 0; g0 < ncols; g0 += 32) {
        const int gw = min(32, ncols - g0);
        if (gw == 32) tmem_ld32(trow + col0 + g0, r);
        else tmem_ld16_h(trow + col0 + g0, r);
        tmem_ld_wait();
        if (is_text) {
          const int nv = min(text_limit - (64 * c + g0), gw);
          vm = nv <= 0 ? 0u : (nv >= 32 ? 0xFFFFFFFFu : ((1u << nv) - 1u));
        }
        float gmax = -1e30f;
#pragma unroll
        for (int j = 0; j < 32; ++j) gmax = fmaxf(gmax, ((vm >> j) & 1u) ? __uint_as_float(r[j]) : -1e30f);
        // (a group is at most 32 keys, so the reference may be raised per group)
        const float m_c = gmax * kLog2e;
        const bool need = q_valid && (m_c > m_used + AF4_RESCALE_LOG2);
        float factor = 1.0f;
        if (need) { factor = fast_exp2(m_used - m_c); m_used = m_c; l_run *= factor; lsum0 *= factor; lsum1 *= factor; }
        if (__any_sync(0xFFFFFFFFu, need)) {
          // rescale what this row has accumulated so far: O (if any chunk has been added) ...
          if (c > 0) {
#pragma unroll 1
            for (int half = 0; half < 2; ++half) {
              uint32_t o[32];
              tmem_ld32(trow + AF4_TM_O + half * 32, o);
              tmem_ld_wait();
#pragma unroll
              for (int j = 0; j < 32; ++j) o[j] = __float_as_uint(__uint_as_float(o[j]) * factor);
              tmem_st32(trow + AF4_TM_O + half * 32, o);
            }
          }
          // ... and the P columns of this chunk's earlier group (they were formed against the old reference)
          if (g0 > 0) {
            uint32_t pp[32];
            tmem_ld16_h(trow + col0 / 2, pp);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const float2 f = unpack_bf16x2(pp[j]);
              pp[j] = pack_bf16x2(f.x * factor, f.y * factor);
            }
            tmem_st16_h(trow + col0 / 2, pp);
          }
        }
#pragma unroll
        for (int j = 0; j < 32; j += 2) {
          const float e0 = ((vm >> j) & 1u) ? fast_exp2(__uint_as_float(r[j]) * kLog2e - m_used) : 0.f;
          const float e1 = ((vm >> (j + 1)) & 1u) ? fast_exp2(__uint_as_float(r[j + 1]) * kLog2e - m_used) : 0.f;
          lsum0 += e0; lsum1 += e1;
          r[j >> 1] = pack_bf16x2(e0, e1);
        }
        if (is_text) {
          // P columns of this group; the S columns they cover are already in registers
          if (gw == 32) tmem_st16_h(trow + (g0 >> 1), r);
          else tmem_st8_h(trow + (g0 >> 1), r);
        }
      }
      if (!is_text) {
        // local chunk: the window's P (or zeros) for the chunk's two 32-key halves; a lane only ever reads the S
        // columns of its own window, so writing both halves now is safe
        uint32_t z[32];
#pragma unroll
        for (int j = 0; j < 16; ++j) z[j] = 0u;
        const int my_half = (ncols > 0) ? (warp & 1) : -1;
        if (my_half == 0) { tmem_st16_h(trow, r); tmem_st16_h(trow + 16, z); }
        else if (my_half == 1) { tmem_st16_h(trow, z); tmem_st16_h(trow + 16, r); }
        else { tmem_st16_h(trow, z); tmem_st16_h(trow + 16, z); }
      }
    }
    l_run += lsum0 + lsum1;
    tmem_st_wait();
    tc_fence_before();
    mbar_arrive(bar_p);
    if (warp == 0 && elect_one()) {
      issue_pv(c);
      if (c + 1 < n_chunks) issue_s(c + 1);   // queued behind P V(c): it overwrites the columns P V(c) reads
    }
    __syncwarp();
  }

  // ---- epilogue: O / l -> bf16 through the (now free) Q tile, whole 128-byte lines per store instruction
  mbar_wait(bar_o, 0);
  tc_fence_after();
  {
    uint32_t o0[32], o1[32];
    tmem_ld32(trow + AF4_TM_O, o0);
    tmem_ld32(trow + AF4_TM_O + 32, o1);
    tmem_ld_wait();
    const float inv = l_run > 0.f ? 1.0f / l_run : 0.f;
    uint8_t* stg = sQ + warp * 4096;
#pragma unroll
    for (int c8 = 0; c8 < 8; ++c8) {
      const uint32_t* src = c8 < 4 ? o0 + 8 * c8 : o1 + 8 * (c8 - 4);
      uint4 u;
      u.x = pack_bf16x2(__uint_as_float(src[0]) * inv, __uint_as_float(src[1]) * inv);
      u.y = pack_bf16x2(__uint_as_float(src[2]) * inv, __uint_as_float(src[3]) * inv);
      u.z = pack_bf16x2(__uint_as_float(src[4]) * inv, __uint_as_float(src[5]) * inv);
      u.w = pack_bf16x2(__uint_as_float(src[6]) * inv, __uint_as_float(src[7]) * inv);
      *reinterpret_cast<uint4*>(stg + lane * 128 + ((c8 ^ (lane & 7)) << 4)) = u;
    }
    __syncwarp();
    const int my_row = q_valid ? row_base + tpos : -1;
    const int cq = lane & 7;
#pragma unroll
    for (int it = 0; it < 8; ++it) {
      const int rr = it * 4 + (lane >> 3);
      const int grow = __shfl_sync(0xFFFFFFFFu, my_row, rr);
      const uint4 u = *reinterpret_cast<const uint4*>(stg + rr * 128 + ((cq ^ (rr & 7)) << 4));
      if (grow >= 0) *reinterpret_cast<uint4*>(a.out + static_cast<long long>(grow) * a.ld_out + h * 64 + cq * 8) = u;
    }
    if (q_valid) a.lse[(static_cast<long long>(b) * g.H + h) * g.seq + tpos] = m_used + log2f(l_run);
  }
  if (MODE == ATT_TEXT && a.text_shift && blk == 0 && warp == 1) {
    // token 0 attends to itself only: o = v_0, lse = (q_0 . k_0) log2(e).  Lane = two head dims.
    const __nv_bfloat16* r0 = reinterpret_cast<const __nv_bfloat16*>(a.qkv) + static_cast<long long>(row_base) * 3 * HD + h * 64 + 2 * lane;
    const float2 qf = unpack_bf16x2(*reinterpret_cast<const uint32_t*>(r0));
    const float2 kf = unpack_bf16x2(*reinterpret_cast<const uint32_t*>(r0 + HD));
    const float s = warp_sum(qf.x * kf.x + qf.y * kf.y);
    *reinterpret_cast<uint32_t*>(a.out + static_cast<long long>(row_base) * a.ld_out + h * 64 + 2 * lane) =
        *reinterpret_cast<const uint32_t*>(r0 + 2 * HD);
    if (lane == 0) a.lse[(static_cast<long long>(b) * g.H + h) * g.seq] = s * kLog2e;
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem, 128);
  }
}

// Text and image query blocks are separate launches: one kernel holding both bodies ran 40 % slower (243 vs 174 us at
// batch 32; twice the code in the instruction caches of SMs that then run both paths side by side).
template <int IMG_MODE, bool TEXT>
__global__ void __launch_bounds__(AF4_THREADS, 4)
attn_fwd4_kernel(const __grid_constant__ CUtensorMap tm128, const __grid_constant__ CUtensorMap tm64,
                 const __grid_constant__ CUtensorMap tmTail, const __grid_constant__ CUtensorMap tmCol,
                 const AttnFwd4Args a) {
  attn_fwd4_body<IMG_MODE, TEXT>(tm128, tm64, tmTail, tmCol, a, blockIdx.x);
}

template <int MODE, bool TEXT>
static int launch_fwd4(const CUtensorMap& t128, const CUtensorMap& t64, const CUtensorMap& ttail,
                       const CUtensorMap& tcol, const AttnFwd4Args& a, cudaStream_t st) {
  auto kern = attn_fwd4_kernel<MODE, TEXT>;
  static bool attr_done = false;
  if (!attr_done) {
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, AF4_SMEM) != cudaSuccess) return -21;
    cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    attr_done = true;
  }
  dim3 grid(TEXT ? a.g.n_text_blocks : a.g.n_img_blocks, a.g.H, a.g.B);
  kern<<<grid, AF4_THREADS, AF4_SMEM, st>>>(t128, t64, ttail, tcol, a);
  return cudaGetLastError() == cudaSuccess ? 0 : -22;
}

// 0 = launched, 1 = geometry / pattern not covered by this kernel (caller falls back), < 0 = error
int attn4_fwd(const void* qkv, void* out, float* lse, int B, int H, int seq, int T, int F, int mode,
              cudaStream_t st) {
  if (mode != ATT_ROW && mode != ATT_COL) return 1;
  AttnFwd4Args a;
  AttnGeom& g = a.g;
  g.B = B; g.H = H; g.seq = seq; g.T = T; g.T_pad = (T + 15) / 16 * 16; g.F = F; g.P = F * F;
  g.n_text_blocks = (T + 127) / 128; g.n_img_blocks = g.P / 128; g.conv_half = 0;
  if (g.P % 128 || 32 % F || seq != T + g.P - 1) return 1;
  a.out = reinterpret_cast<__nv_bfloat16*>(out); a.lse = lse; a.ld_out = H * 64; a.qkv = qkv;
  a.n_text_chunks = (g.T_pad + 63) / 64;
  // T = 128 k + 1 (257 here): with blocks starting at token 0 the last block would hold ONE query and still walk every
  // key chunk.  Token 0 sees only itself, so it is written directly and the blocks start at token 1: 3 + 5 instead of
  // 2 + 4 + 5 chunk passes per (sample, head).
  a.text_shift = (T > 1 && T % 128 == 1) ? 1 : 0;
  g.n_text_blocks = (T - a.text_shift + 127) / 128;
  CUtensorMap t128, t64, ttail, tunused, tcol;
  int rc = attn_make_maps(qkv, g, ATT_COL, &t128, &tunused, &tcol);
  if (rc) return rc;
  const uint64_t rows = (uint64_t)B * seq, cols = 3ull * H * 64;
  rc = make_tmap_bf16_2d(&t64, qkv, rows, cols, cols, 64, 64);
  if (rc) return rc;
  const int tail = g.T_pad % 64;
  rc = make_tmap_bf16_2d(&ttail, qkv, rows, cols, cols, 64, tail ? tail : 64);
  if (rc) return rc;
  // text blocks and image blocks write disjoint rows: the two launches run side by side (attn_common.cuh)
  AttnFork* fk = attn_fork_state(2);
  const cudaStream_t side = attn_fork(fk, st);
  rc = launch_fwd4<ATT_ROW, true>(t128, t64, ttail, tcol, a, side);
  const int rc2 = mode == ATT_ROW ? launch_fwd4<ATT_ROW, false>(t128, t64, ttail, tcol, a, st)
                                  : launch_fwd4<ATT_COL, false>(t128, t64, ttail, tcol, a, st);
  attn_join(fk, side, st);
  return rc ? rc : rc2;
}

}  // namespace b200
