// Sparse DALL-E attention backward, fourth generation (backward halves: text, axial_row,
// axial_col; conv_like keeps attn_bwd_sm100.cu).
//
// Same philosophy as attn_fwd4_sm100.cu: light CTAs that the hardware interleaves (two per SM, 16 warps), 64-wide
// score tiles read ONCE from TMEM, and every bf16 intermediate (dS, P^T, dS^T) written back into TMEM and consumed
// by tcgen05.mma as its A operand - no shared-memory copy of P / dS, no fp32 atomics for the shared text keys, no
// full-size finish pass.  The work is split by what stays put:
//
//   attn_bwd4_q_kernel   one CTA = (batch, head, 128-query block), queries on the TMEM lanes.  Per 64-key chunk:
//                        S = Q K^T, dP = dO V^T -> dS = P (dP - delta) -> dQ += dS K (dQ accumulates in TMEM).
//                        Also forms delta = rowsum(dO * O) (kept for the second kernel) and, on CUDA cores, the
//                        dK / dV contributions of the one or two text keys left over by the 128-key tiles (T = 257:
//                        key 256) into a few fp32 rows of `kv_acc`.
//   attn_bwd4_kv_kernel  one CTA = (batch, head, 128-key tile), keys on the TMEM lanes: the text tiles walk ALL query
//                        half-blocks of the head and accumulate dK / dV in TMEM (what used to be 630 MB of red.add
//                        per layer at batch 32), the local tiles walk the two half-blocks of their own query block.
//                        Per 64-query step: S^T = K Q^T, dP^T = V dO^T -> P^T, dS^T -> dV += P^T dO, dK += dS^T Q.
//   attn_bwd4_tail_kernel  converts the few leftover-key rows of kv_acc (inverse rotary, bf16) and clears them.
//
// 256 threads: the two warpgroups split the 64 columns of a tile, thread = (lane row, 32 columns).  A thread writes
// its 16 packed bf16 columns INSIDE the fp32 columns it has just read, so the A operand of the following MMA is two
// 16-column pieces (k-steps at column offsets 0, 8, 32, 40) and no thread overwrites what another still reads.
#include "attn_common.cuh"
#include "b200_api.h"

#include <cstdio>
#include <cstdlib>

namespace b200 {

int make_tmap_bf16_2d(CUtensorMap* m, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_cols,
                      uint32_t box_rows);
int make_tmap_bf16_nd(CUtensorMap* m, const void* ptr, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                      const uint32_t* box);

constexpr int AB4_THREADS = 256;
constexpr int AB4_SMEM = 2 * 16384 + 4 * 8192 + 1024 + 256 + 1024;   // Q, dO | K, V tiles; chunk buffers; stats; barriers
constexpr uint32_t AB4_TM_S = 0, AB4_TM_DP = 64, AB4_TM_ACC0 = 128, AB4_TM_ACC1 = 192;
constexpr int AB4_MAX_TAIL = 4;   // leftover text keys handled on CUDA cores

struct AttnBwd4Args {
  AttnGeom g;
  const float* lse;            // [B,H,seq] log2 domain
  float* delta;                // [B,H,seq]  written by the q kernel, read by the kv kernel
  const __nv_bfloat16* out;    // forward output [B*seq, H*64]
  __nv_bfloat16* dqkv;         // [B*seq, 3*H*64]
  float* kv_acc;               // [B*seq, 2*H*64] fp32, zero on entry and on exit (only the leftover-key rows are used)
  const float2* rot;           // [seq+1][32] (cos, sin), position-major: a token's 32 pairs are 256 contiguous bytes
  const float2* rot_t;         // [32][seq+1] pair-major (tail kernel)
  int rot_len;
  float q_scale;
  int n_text_chunks;           // q kernel: ceil(T_pad / 64) key chunks
  int n_text_tiles;            // kv kernel: 128-key text tiles (full ones, plus a partial one when it is not `tail`)
  int tail0, n_tail;           // leftover text keys [tail0, tail0 + n_tail) handled on CUDA cores (n_tail <= 4)
  int text_shift;              // 1: text query blocks start at token 1; token 0 (sees itself only) is written directly
  const __nv_bfloat16* qkv;
  const __nv_bfloat16* d_out;
};

__device__ __forceinline__ void ab4_red_add_v2(float* p, float a, float b) {
  asm volatile("red.global.add.v2.f32 [%0], {%1, %2};" ::"l"(p), "f"(a), "f"(b) : "memory");
}

// v[0..31] = head dims [d0, d0 + 32) of one token -> inverse rotation, scaled.  rot_pos = the token's 32 (cos, sin)
// pairs (position-major table: 128 contiguous bytes per half head, whatever the stride between the lanes' tokens)
__device__ __forceinline__ void ab4_inv_rotary32(float* v, const float2* rot_pos, int d0, float scale) {
  const float4* t4 = reinterpret_cast<const float4*>(rot_pos + (d0 >> 1));
#pragma unroll
  for (int j4 = 0; j4 < 8; ++j4) {
    const float4 cs = __ldg(t4 + j4);   // (cos, sin) of two consecutive pairs
    const int j = 4 * j4;
    const float x0 = v[j], x1 = v[j + 1], x2 = v[j + 2], x3 = v[j + 3];
    v[j] = (x0 * cs.x + x1 * cs.y) * scale;
    v[j + 1] = (x1 * cs.x - x0 * cs.y) * scale;
    v[j + 2] = (x2 * cs.z + x3 * cs.w) * scale;
    v[j + 3] = (x3 * cs.z - x2 * cs.w) * scale;
  }
}

// 32 rows x 64 bytes of this warp -> global, 64-byte segments per row (staged through 2 KB of shared memory so that a
// store instruction covers eight rows' segments instead of 32 different lines)
__device__ __forceinline__ void ab4_store_rows64(uint8_t* stg, const float* v, int lane, int my_row,
                                                 __nv_bfloat16* base, long long ld, int col) {
#pragma unroll
  for (int c4 = 0; c4 < 4; ++c4) {
    uint4 u;
    u.x = pack_bf16x2(v[8 * c4 + 0], v[8 * c4 + 1]);
    u.y = pack_bf16x2(v[8 * c4 + 2], v[8 * c4 + 3]);
    u.z = pack_bf16x2(v[8 * c4 + 4], v[8 * c4 + 5]);
    u.w = pack_bf16x2(v[8 * c4 + 6], v[8 * c4 + 7]);
    *reinterpret_cast<uint4*>(stg + lane * 64 + ((c4 ^ ((lane >> 1) & 3)) << 4)) = u;
  }
  __syncwarp();
  const int cq = lane & 3;
#pragma unroll
  for (int it = 0; it < 4; ++it) {
    const int rr = it * 8 + (lane >> 2);
    const int grow = __shfl_sync(0xFFFFFFFFu, my_row, rr);
    const uint4 u = *reinterpret_cast<const uint4*>(stg + rr * 64 + ((cq ^ ((rr >> 1) & 3)) << 4));
    if (grow >= 0) *reinterpret_cast<uint4*>(base + static_cast<long long>(grow) * ld + col + cq * 8) = u;
  }
  __syncwarp();
}

// =====================================================================================================================
// q kernel
// =====================================================================================================================
template <int IMG_MODE, bool TEXT>
__device__ __forceinline__ void attn_bwd4_q_body(const CUtensorMap& tmQ128, const CUtensorMap& tmQ64,
                                                 const CUtensorMap& tmQtail, const CUtensorMap& tmQcol,
                                                 const CUtensorMap& tmDO128, const CUtensorMap& tmDOcol,
                                                 const AttnBwd4Args& a, const int blk) {
  constexpr int MODE = TEXT ? ATT_TEXT : IMG_MODE;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + (((raw_addr + 1023u) & ~1023u) - raw_addr);
  const AttnGeom& g = a.g;
  uint8_t* sQ = smem;                 // 16 KB (later: staging of the dQ rows)
  uint8_t* sDO = smem + 16384;        // 16 KB
  uint8_t* sK = smem + 32768;         // 2 x 8 KB
  uint8_t* sV = sK + 16384;           // 2 x 8 KB
  uint64_t* bars = reinterpret_cast<uint64_t*>(sV + 16384 + 1024);
  uint64_t* bar_q = bars + 0;
  uint64_t* bar_k = bars + 1;         // [2]
  uint64_t* bar_v = bars + 3;         // [2]
  uint64_t* bar_s = bars + 5;
  uint64_t* bar_p = bars + 6;
  uint64_t* bar_dq = bars + 7;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 8);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int qw = warp & 3, ch = warp >> 2;      // TMEM lane quarter, column half
  const int h = blockIdx.y, b = blockIdx.z;
  const int q0 = blk * 128 + (MODE == ATT_TEXT ? a.text_shift : 0);
  const int HD = g.H * 64;
  const int row_base = b * g.seq;
  const int lines = 128 / g.F, lines64 = 64 / g.F;

  const int n_text = (MODE == ATT_TEXT) ? min((min(q0 + 128, g.T) + 63) / 64, a.n_text_chunks) : a.n_text_chunks;
  const int n_chunks = n_text + (MODE == ATT_TEXT ? 0 : 2);

  if (threadIdx.x == 0) {
    mbar_init(bar_q, 1);
    mbar_init(&bar_k[0], 1); mbar_init(&bar_k[1], 1);
    mbar_init(&bar_v[0], 1); mbar_init(&bar_v[1], 1);
    mbar_init(bar_s, 1); mbar_init(bar_p, AB4_THREADS); mbar_init(bar_dq, 1);
    fence_barrier_init();
  }
  if (warp == 0) {
    tmem_alloc(tmem_slot, 256);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  auto chunk_keys = [&](int c) -> int { return c < n_text ? min(64, g.T_pad - 64 * c) : 64; };
  auto load_kv = [&](int c, int sect) {
    uint64_t* full = sect == 0 ? &bar_k[c & 1] : &bar_v[c & 1];
    uint8_t* dst = (sect == 0 ? sK : sV) + (c & 1) * 8192;
    const int col = (sect == 0 ? HD : 2 * HD) + h * 64;
    const int nk = chunk_keys(c);
    mbar_expect_tx(full, static_cast<uint32_t>(nk) * 128u);
    if (c < n_text) {
      tma_load_2d(dst, nk == 64 ? &tmQ64 : &tmQtail, full, col, row_base + 64 * c);
    } else {
      const int j = c - n_text;
      if (MODE == ATT_COL) {
        for (int l = 0; l < lines64; ++l)
          tma_load_4d(dst + l * g.F * 128, &tmQcol, full, col, blk * lines + j * lines64 + l, 0, b);
      } else {
        tma_load_2d(dst, &tmQ64, full, col, row_base + g.T + q0 + 64 * j);
      }
    }
  };
  const uint32_t q_addr = smem_u32(sQ), do_addr = smem_u32(sDO), k_addr = smem_u32(sK), v_addr = smem_u32(sV);
  // S = Q K^T and dP = dO V^T of chunk c (both [128 x nk], K-major operands)
  auto issue_s = [&](int c) {
    const int nk = chunk_keys(c);
    mbar_wait(&bar_k[c & 1], (c >> 1) & 1);
    mbar_wait(&bar_v[c & 1], (c >> 1) & 1);
    tc_fence_after();
    const uint32_t idesc = make_idesc_bf16(128, nk, 0, 0);
    const uint64_t dq = make_smem_desc_sw128(q_addr, 0, 1024), ddo = make_smem_desc_sw128(do_addr, 0, 1024);
    const uint64_t dk = make_smem_desc_sw128(k_addr + (c & 1) * 8192, 0, 1024);
    const uint64_t dv = make_smem_desc_sw128(v_addr + (c & 1) * 8192, 0, 1024);
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) umma_bf16_ss(tmem + AB4_TM_S, dq + 2 * kk, dk + 2 * kk, idesc, kk > 0 ? 1u : 0u);
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) umma_bf16_ss(tmem + AB4_TM_DP, ddo + 2 * kk, dv + 2 * kk, idesc, kk > 0 ? 1u : 0u);
    umma_commit(bar_s);
  };
  // dQ += dS K  (A = dS from TMEM, two 16-column pieces; B = K read MN-major: N = 64 dims, K = keys)
  auto issue_dq = [&](int c) {
    const int nk = chunk_keys(c);
    mbar_wait(bar_p, c & 1);
    tc_fence_after();
    constexpr uint32_t idesc = make_idesc_bf16(128, 64, 0, 1);
    const uint64_t dk = make_smem_desc_sw128(k_addr + (c & 1) * 8192, 8192, 1024);
    for (int k16 = 0; k16 < nk / 16; ++k16)
      umma_bf16_ts(tmem + AB4_TM_ACC0, tmem + AB4_TM_DP + (k16 >> 1) * 32 + (k16 & 1) * 8, dk + 128 * k16, idesc,
                   (c > 0 || k16 > 0) ? 1u : 0u);
    if (c == n_chunks - 1) umma_commit(bar_dq);
  };

  if (warp == 0 && elect_one()) {
    mbar_expect_tx(bar_q, 32768);
    if (MODE == ATT_COL) {
      for (int l = 0; l < lines; ++l) {
        tma_load_4d(sQ + l * g.F * 128, &tmQcol, bar_q, h * 64, blk * lines + l, 0, b);
        tma_load_4d(sDO + l * g.F * 128, &tmDOcol, bar_q, h * 64, blk * lines + l, 0, b);
      }
    } else {
      const int t0 = (MODE == ATT_TEXT) ? q0 : g.T + q0;
      tma_load_2d(sQ, &tmQ128, bar_q, h * 64, row_base + t0);
      tma_load_2d(sDO, &tmDO128, bar_q, h * 64, row_base + t0);
    }
    load_kv(0, 0);
    load_kv(0, 1);
    if (n_chunks > 1) { load_kv(1, 0); load_kv(1, 1); }
    mbar_wait(bar_q, 0);
    issue_s(0);
  }
  __syncwarp();

  // ---- thread = (query row i, key columns [32 ch, 32 ch + 32) of every chunk)
  const int i = qw * 32 + lane;
  const uint32_t trow = tmem + (static_cast<uint32_t>(qw * 32) << 16);
  int tpos;
  uint32_t lmask = 0u;
  if (MODE == ATT_TEXT) {
    tpos = q0 + i;
  } else {
    const int ip = img_query_pos<MODE>(g, blk, i);
    const int ir = ip / g.F, ic = ip - ir * g.F;
    tpos = g.T + ip;
    const int lo = max(i - (MODE == ATT_ROW ? ic : ir), qw * 32) - qw * 32;
    lmask = (0xFFFFFFFFu >> (31 - lane)) & (0xFFFFFFFFu << lo);
  }
  const bool q_valid = (MODE == ATT_TEXT) ? (tpos < g.T) : (tpos < g.seq);
  const int text_limit = q_valid ? ((MODE == ATT_TEXT) ? tpos + 1 : g.T) : 0;
  if (!q_valid) lmask = 0u;
  const long long bh = static_cast<long long>(b) * g.H + h;
  const float lse_i = q_valid ? a.lse[bh * g.seq + tpos] : 0.f;
  // delta_i = sum_d dO[i, d] * O[i, d]
  float delta_i = 0.f;
  {
    uint4 ov[8];
    const uint4* op = reinterpret_cast<const uint4*>(a.out + (static_cast<long long>(row_base) + (q_valid ? tpos : 0)) * HD + h * 64);
#pragma unroll
    for (int c8 = 0; c8 < 8; ++c8) ov[c8] = q_valid ? __ldg(op + c8) : make_uint4(0u, 0u, 0u, 0u);
    mbar_wait(bar_q, 0);
#pragma unroll
    for (int c8 = 0; c8 < 8; ++c8) {
      const uint4 dv = *reinterpret_cast<const uint4*>(sDO + sw128_offset(i, c8));
      float2 x, y;
      x = unpack_bf16x2(ov[c8].x); y = unpack_bf16x2(dv.x); delta_i += x.x * y.x + x.y * y.y;
      x = unpack_bf16x2(ov[c8].y); y = unpack_bf16x2(dv.y); delta_i += x.x * y.x + x.y * y.y;
      x = unpack_bf16x2(ov[c8].z); y = unpack_bf16x2(dv.z); delta_i += x.x * y.x + x.y * y.y;
      x = unpack_bf16x2(ov[c8].w); y = unpack_bf16x2(dv.w); delta_i += x.x * y.x + x.y * y.y;
    }
    if (ch == 0 && q_valid) a.delta[bh * g.seq + tpos] = delta_i;
  }
  float p_tail[AB4_MAX_TAIL], ds_tail[AB4_MAX_TAIL];
#pragma unroll
  for (int t = 0; t < AB4_MAX_TAIL; ++t) { p_tail[t] = 0.f; ds_tail[t] = 0.f; }

  for (int c = 0; c < n_chunks; ++c) {
    const int nk = chunk_keys(c);
    const bool is_text = c < n_text;
    mbar_wait(bar_s, c & 1);
    tc_fence_after();
    // S(c), dP(c) complete, and with them dQ(c - 1): the other K buffer and this chunk's V buffer can be refilled
    if (warp == 0 && elect_one()) {
      if (c >= 1 && c + 1 < n_chunks) load_kv(c + 1, 0);
      if (c + 2 < n_chunks) load_kv(c + 2, 1);
    }
    __syncwarp();
    const int mycols = min(32, nk - 32 * ch);   // columns of this chunk that fall to this thread's half
    if (mycols > 0) {
      uint32_t vm;
      if (is_text) {
        const int nv = min(text_limit - (64 * c + 32 * ch), mycols);
        vm = nv <= 0 ? 0u : (nv >= 32 ? 0xFFFFFFFFu : ((1u << nv) - 1u));
      } else {
        vm = (2 * (c - n_text) + ch == qw) ? lmask : 0u;   // the 32 local keys of this half are one warp's window
      }
      uint32_t rs[32], rp[32];
      if (__any_sync(0xFFFFFFFFu, vm != 0u)) {
        if (mycols == 32) { tmem_ld32(trow + AB4_TM_S + 32 * ch, rs); tmem_ld32(trow + AB4_TM_DP + 32 * ch, rp); }
        else { tmem_ld16_h(trow + AB4_TM_S + 32 * ch, rs); tmem_ld16_h(trow + AB4_TM_DP + 32 * ch, rp); }
        tmem_ld_wait();
        if (__all_sync(0xFFFFFFFFu, vm == 0xFFFFFFFFu)) {
#pragma unroll
          for (int j = 0; j < 32; j += 2) {
            const float p0 = fast_exp2(__uint_as_float(rs[j]) * kLog2e - lse_i);
            const float p1 = fast_exp2(__uint_as_float(rs[j + 1]) * kLog2e - lse_i);
            rp[j >> 1] = pack_bf16x2(p0 * (__uint_as_float(rp[j]) - delta_i), p1 * (__uint_as_float(rp[j + 1]) - delta_i));
          }
        } else {
          float pj[2], dj[2];
#pragma unroll
          for (int j = 0; j < 32; j += 2) {
#pragma unroll
            for (int e = 0; e < 2; ++e) {
              pj[e] = ((vm >> (j + e)) & 1u) ? fast_exp2(__uint_as_float(rs[j + e]) * kLog2e - lse_i) : 0.f;
              dj[e] = pj[e] * (__uint_as_float(rp[j + e]) - delta_i);
            }
            // the first columns of the tail chunk are the leftover text keys whose dK / dV are formed below
            if (is_text && c == n_text - 1 && a.n_tail > 0 && ch == 0 && j < AB4_MAX_TAIL) {
              p_tail[j] = pj[0]; ds_tail[j] = dj[0];
              if (j + 1 < AB4_MAX_TAIL) { p_tail[j + 1] = pj[1]; ds_tail[j + 1] = dj[1]; }
            }
            rp[j >> 1] = pack_bf16x2(dj[0], dj[1]);
          }
        }
      } else {
#pragma unroll
        for (int j = 0; j < 16; ++j) rp[j] = 0u;
      }
      // dS (bf16) over the first half of this thread's own dP columns
      if (mycols == 32) tmem_st16_h(trow + AB4_TM_DP + 32 * ch, rp);
      else tmem_st8_h(trow + AB4_TM_DP + 32 * ch, rp);
      tmem_st_wait();
    }
    tc_fence_before();
    mbar_arrive(bar_p);
    if (warp == 0 && elect_one()) {
      issue_dq(c);
      if (c + 1 < n_chunks) issue_s(c + 1);   // behind dQ(c) in the queue: it overwrites what dQ(c) reads
    }
    __syncwarp();
  }

  // ---- leftover text keys: dK[key] += sum_i dS[i, key] Q[i, :], dV[key] += sum_i P[i, key] dO[i, :] over this
  // warp's 32 rows (lane = two head dims), then one red.add per lane and key into kv_acc
  if (a.n_tail > 0 && ch == 0 && n_text == a.n_text_chunks) {
    for (int t = 0; t < a.n_tail; ++t) {
      float ak0 = 0.f, ak1 = 0.f, av0 = 0.f, av1 = 0.f;
      float pt = 0.f, dt = 0.f;
#pragma unroll
      for (int tt = 0; tt < AB4_MAX_TAIL; ++tt)
        if (tt == t) { pt = p_tail[tt]; dt = ds_tail[tt]; }
#pragma unroll 8
      for (int r = 0; r < 32; ++r) {
        const float pr = __shfl_sync(0xFFFFFFFFu, pt, r), dr = __shfl_sync(0xFFFFFFFFu, dt, r);
        const int row = qw * 32 + r;
        const uint32_t off = sw128_offset(row, lane >> 2) + (lane & 3) * 4;
        const float2 qv = unpack_bf16x2(*reinterpret_cast<const uint32_t*>(sQ + off));
        const float2 ov = unpack_bf16x2(*reinterpret_cast<const uint32_t*>(sDO + off));
        ak0 += dr * qv.x; ak1 += dr * qv.y;
        av0 += pr * ov.x; av1 += pr * ov.y;
      }
      float* dst = a.kv_acc + (static_cast<long long>(row_base) + a.tail0 + t) * (2 * HD) + h * 64 + 2 * lane;
      ab4_red_add_v2(dst, ak0, ak1);
      ab4_red_add_v2(dst + HD, av0, av1);
    }
  }

  // ---- dQ: inverse rotary + q scale, bf16, out through the (dead) K / V chunk buffers
  mbar_wait(bar_dq, 0);
  tc_fence_after();
  {
    uint32_t r[32];
    tmem_ld32(trow + AB4_TM_ACC0 + 32 * ch, r);
    tmem_ld_wait();
    float v[32];
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
    if (q_valid) ab4_inv_rotary32(v, a.rot + static_cast<long long>(tpos) * 32, 32 * ch, a.q_scale);
    ab4_store_rows64(sK + warp * 2048, v, lane, q_valid ? row_base + tpos : -1, a.dqkv, 3 * HD, h * 64 + 32 * ch);
  }

  if (MODE == ATT_TEXT && a.text_shift && blk == 0 && warp == 1) {
    // token 0: P = 1 on its single key, so dS = dP - delta = 0: dq_0 = 0 and its dk_0 / dv_0 terms come out of the
    // key/value kernel, which only needs delta_0 = dO_0 . o_0 = dO_0 . v_0.  Lane = two head dims.
    const long long r0 = static_cast<long long>(row_base);
    const float2 vf = unpack_bf16x2(*reinterpret_cast<const uint32_t*>(a.qkv + r0 * 3 * HD + 2 * HD + h * 64 + 2 * lane));
    const float2 df = unpack_bf16x2(*reinterpret_cast<const uint32_t*>(a.d_out + r0 * HD + h * 64 + 2 * lane));
    const float d0 = warp_sum(vf.x * df.x + vf.y * df.y);
    *reinterpret_cast<uint32_t*>(a.dqkv + r0 * 3 * HD + h * 64 + 2 * lane) = 0u;
    if (lane == 0) a.delta[(static_cast<long long>(b) * g.H + h) * g.seq] = d0;
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem, 256);
  }
}

template <int IMG_MODE, bool TEXT>
__global__ void __launch_bounds__(AB4_THREADS, 2)
attn_bwd4_q_kernel(const __grid_constant__ CUtensorMap tmQ128, const __grid_constant__ CUtensorMap tmQ64,
                   const __grid_constant__ CUtensorMap tmQtail, const __grid_constant__ CUtensorMap tmQcol,
                   const __grid_constant__ CUtensorMap tmDO128, const __grid_constant__ CUtensorMap tmDOcol,
                   const AttnBwd4Args a) {
  attn_bwd4_q_body<IMG_MODE, TEXT>(tmQ128, tmQ64, tmQtail, tmQcol, tmDO128, tmDOcol, a, blockIdx.x);
}

// =====================================================================================================================
// kv kernel
// =====================================================================================================================
template <int IMG_MODE>
__global__ void __launch_bounds__(AB4_THREADS, 2)
attn_bwd4_kv_kernel(const __grid_constant__ CUtensorMap tmQ128, const __grid_constant__ CUtensorMap tmQ64,
                    const __grid_constant__ CUtensorMap tmQcol, const __grid_constant__ CUtensorMap tmDO64,
                    const __grid_constant__ CUtensorMap tmDOcol, const AttnBwd4Args a) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + (((raw_addr + 1023u) & ~1023u) - raw_addr);
  const AttnGeom& g = a.g;
  uint8_t* sK = smem;                 // 16 KB key tile (later: staging of the dK / dV rows)
  uint8_t* sV = smem + 16384;         // 16 KB value tile
  uint8_t* sQ = smem + 32768;         // 2 x 8 KB query half-blocks
  uint8_t* sDO = sQ + 16384;          // 2 x 8 KB
  float* s_lse = reinterpret_cast<float*>(sDO + 16384);   // [2][64]
  float* s_del = s_lse + 128;                             // [2][64]
  uint64_t* bars = reinterpret_cast<uint64_t*>(sDO + 16384 + 1024);
  uint64_t* bar_kv = bars + 0;
  uint64_t* bar_q = bars + 1;         // [2]
  uint64_t* bar_s = bars + 3;
  uint64_t* bar_p = bars + 4;
  uint64_t* bar_acc = bars + 5;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 8);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int qw = warp & 3, ch = warp >> 2;
  // 1-D grid, LONG CTAs FIRST: the text key tiles of every (sample, head) - each walks all ~20 query half-steps -
  // take the lowest block indices, the two-step image tiles follow and level the tail of the grid.  (With tiles
  // interleaved per (sample, head) the last wave ended on a few 20-step CTAs while most SMs idled.)
  const int n_bh = g.B * g.H;
  const int lin = blockIdx.x;
  const bool text_tile = lin < a.n_text_tiles * n_bh;
  const int lin2 = text_tile ? lin : lin - a.n_text_tiles * n_bh;
  const int per_bh = text_tile ? a.n_text_tiles : g.n_img_blocks;
  const int bh_lin = lin2 / per_bh;
  const int tile = lin2 - bh_lin * per_bh;                                     // text tile index / image block index
  const int b = bh_lin / g.H, h = bh_lin - b * g.H;
  const int HD = g.H * 64;
  const int row_base = b * g.seq;
  const int lines = 128 / g.F, lines64 = 64 / g.F;
  const int n_img_half = 2 * g.n_img_blocks;
  const int n_text_half = (g.T + 63) / 64;

  // key tile of this CTA and the query half-blocks (64 queries) that can see it
  const int tv0 = text_tile ? 2 * tile : 0;                               // first text half-step (causal)
  const int n_steps = text_tile ? n_img_half + (n_text_half - tv0) : 2;
  // step u -> (is_text_q, half-block index): text tiles take the image half-blocks first, then the text ones
  auto step_text_q = [&](int u) -> bool { return text_tile && u >= n_img_half; };
  auto step_hb = [&](int u) -> int { return text_tile ? (u < n_img_half ? u : tv0 + (u - n_img_half)) : 2 * tile + u; };

  if (threadIdx.x == 0) {
    mbar_init(bar_kv, 1);
    mbar_init(&bar_q[0], 1); mbar_init(&bar_q[1], 1);
    mbar_init(bar_s, 1); mbar_init(bar_p, AB4_THREADS); mbar_init(bar_acc, 1);
    fence_barrier_init();
  }
  if (warp == 0) {
    tmem_alloc(tmem_slot, 256);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  auto load_q = [&](int u) {   // Q and dO of step u into buffer u & 1
    uint64_t* full = &bar_q[u & 1];
    uint8_t* dq = sQ + (u & 1) * 8192;
    uint8_t* dd = sDO + (u & 1) * 8192;
    const int hb = step_hb(u);
    mbar_expect_tx(full, 16384);
    if (step_text_q(u)) {
      tma_load_2d(dq, &tmQ64, full, h * 64, row_base + 64 * hb);
      tma_load_2d(dd, &tmDO64, full, h * 64, row_base + 64 * hb);
    } else if (IMG_MODE == ATT_COL) {
      for (int l = 0; l < lines64; ++l) {
        tma_load_4d(dq + l * g.F * 128, &tmQcol, full, h * 64, hb * lines64 + l, 0, b);
        tma_load_4d(dd + l * g.F * 128, &tmDOcol, full, h * 64, hb * lines64 + l, 0, b);
      }
    } else {
      tma_load_2d(dq, &tmQ64, full, h * 64, row_base + g.T + 64 * hb);
      tma_load_2d(dd, &tmDO64, full, h * 64, row_base + g.T + 64 * hb);
    }
  };
  const uint32_t k_addr = smem_u32(sK), v_addr = smem_u32(sV), q_addr = smem_u32(sQ), do_addr = smem_u32(sDO);
  // S^T = K Q^T, dP^T = V dO^T  ([128 keys x 64 queries]; all operands K-major over the head dim)
  auto issue_s = [&](int u) {
    mbar_wait(&bar_q[u & 1], (u >> 1) & 1);
    tc_fence_after();
    constexpr uint32_t idesc = make_idesc_bf16(128, 64, 0, 0);
    const uint64_t dk = make_smem_desc_sw128(k_addr, 0, 1024), dv = make_smem_desc_sw128(v_addr, 0, 1024);
    const uint64_t dq = make_smem_desc_sw128(q_addr + (u & 1) * 8192, 0, 1024);
    const uint64_t dd = make_smem_desc_sw128(do_addr + (u & 1) * 8192, 0, 1024);
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) umma_bf16_ss(tmem + AB4_TM_S, dk + 2 * kk, dq + 2 * kk, idesc, kk > 0 ? 1u : 0u);
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) umma_bf16_ss(tmem + AB4_TM_DP, dv + 2 * kk, dd + 2 * kk, idesc, kk > 0 ? 1u : 0u);
    umma_commit(bar_s);
  };
  // dV += P^T dO, dK += dS^T Q  (A from TMEM in two 16-column pieces; B read MN-major: N = 64 dims, K = 64 queries)
  auto issue_acc = [&](int u) {
    mbar_wait(bar_p, u & 1);
    tc_fence_after();
    constexpr uint32_t idesc = make_idesc_bf16(128, 64, 0, 1);
    const uint64_t dd = make_smem_desc_sw128(do_addr + (u & 1) * 8192, 8192, 1024);
    const uint64_t dq = make_smem_desc_sw128(q_addr + (u & 1) * 8192, 8192, 1024);
#pragma unroll
    for (int k16 = 0; k16 < 4; ++k16) {
      const uint32_t off = (k16 >> 1) * 32 + (k16 & 1) * 8;
      umma_bf16_ts(tmem + AB4_TM_ACC1, tmem + AB4_TM_S + off, dd + 128 * k16, idesc, (u > 0 || k16 > 0) ? 1u : 0u);
      umma_bf16_ts(tmem + AB4_TM_ACC0, tmem + AB4_TM_DP + off, dq + 128 * k16, idesc, (u > 0 || k16 > 0) ? 1u : 0u);
    }
    if (u == n_steps - 1) umma_commit(bar_acc);
  };

  if (warp == 0 && elect_one()) {
    const int kcol = HD + h * 64, vcol = 2 * HD + h * 64;
    mbar_expect_tx(bar_kv, 32768);
    if (text_tile) {
      tma_load_2d(sK, &tmQ128, bar_kv, kcol, row_base + 128 * tile);
      tma_load_2d(sV, &tmQ128, bar_kv, vcol, row_base + 128 * tile);
    } else if (IMG_MODE == ATT_COL) {
      for (int l = 0; l < lines; ++l) {
        tma_load_4d(sK + l * g.F * 128, &tmQcol, bar_kv, kcol, tile * lines + l, 0, b);
        tma_load_4d(sV + l * g.F * 128, &tmQcol, bar_kv, vcol, tile * lines + l, 0, b);
      }
    } else {
      tma_load_2d(sK, &tmQ128, bar_kv, kcol, row_base + g.T + 128 * tile);
      tma_load_2d(sV, &tmQ128, bar_kv, vcol, row_base + g.T + 128 * tile);
    }
    load_q(0);
    if (n_steps > 1) load_q(1);
    mbar_wait(bar_kv, 0);
    issue_s(0);
  }
  __syncwarp();

  // ---- thread = (key row i of the tile, query columns [32 ch, 32 ch + 32) of every step)
  const int i = qw * 32 + lane;
  const uint32_t trow = tmem + (static_cast<uint32_t>(qw * 32) << 16);
  int ktok;   // token index of this key inside its sample, -1 = no such key
  if (text_tile) {
    ktok = 128 * tile + i;
    if (ktok >= g.T || (a.n_tail > 0 && ktok >= a.tail0)) ktok = -1;   // leftover keys belong to the q kernel
  } else {
    const int ip = img_query_pos<IMG_MODE>(g, tile, i);
    ktok = g.T + ip;
    if (ktok >= g.seq) ktok = -1;
  }
  const long long bh = static_cast<long long>(b) * g.H + h;
  const int line_end = (i / g.F) * g.F + g.F - 1;   // local tile: last local index of this key's grid line

  auto fetch_stat = [&](int u) -> float {   // threads 0..127: lse (0..63) or delta (64..127) of query (tid & 63)
    const int qi = threadIdx.x & 63;
    const int hb = step_hb(u);
    int tp;
    bool ok;
    if (step_text_q(u)) { tp = 64 * hb + qi; ok = tp < g.T; }
    else { tp = g.T + img_query_pos<IMG_MODE>(g, hb >> 1, 64 * (hb & 1) + qi); ok = tp < g.seq; }
    const float* src = threadIdx.x < 64 ? a.lse : a.delta;
    return ok ? __ldg(src + bh * g.seq + tp) : 0.f;
  };
  if (threadIdx.x < 128) (threadIdx.x < 64 ? s_lse : s_del)[threadIdx.x & 63] = fetch_stat(0);

  for (int u = 0; u < n_steps; ++u) {
    const bool tq = step_text_q(u);
    const int hb = step_hb(u);
    // log-sum-exp and delta of the NEXT step's 64 queries (threads 0..63 / 64..127) are fetched now and parked in
    // shared memory at the end of this step: their global-memory latency is off the step's critical path
    float stat_next = 0.f;
    if (threadIdx.x < 128 && u + 1 < n_steps) stat_next = fetch_stat(u + 1);
    mbar_wait(bar_s, u & 1);
    tc_fence_after();
    __syncthreads();   // statistics visible; also: every thread is past step u - 1
    // S^T(u), dP^T(u) complete, and with them the accumulations of step u - 1: its Q / dO buffer can be refilled
    if (warp == 0 && elect_one()) {
      if (u >= 1 && u + 1 < n_steps) load_q(u + 1);
    }
    __syncwarp();
    // visibility of this thread's 32 query columns for its key
    uint32_t vm = 0u;
    if (ktok >= 0) {
      if (tq) {
        // text query tp = 64 hb + 32 ch + j sees text key ktok iff ktok <= tp < T
        const int j_lo = max(ktok - (64 * hb + 32 * ch), 0);
        const int j_hi = min(g.T - 1 - (64 * hb + 32 * ch), 31);
        if (j_lo <= j_hi) vm = (0xFFFFFFFFu << j_lo) & (0xFFFFFFFFu >> (31 - j_hi));
      } else {
        if (text_tile) {
          vm = 0xFFFFFFFFu;   // image queries see every text key
        } else {
          // local tile: query local index 64 (hb & 1) + 32 ch + j sees key i iff same grid line and i <= query
          const int qb = 64 * (hb & 1) + 32 * ch;
          const int j_lo = max(i - qb, 0);
          const int j_hi = min(line_end - qb, 31);
          if (j_lo <= j_hi) vm = (0xFFFFFFFFu << j_lo) & (0xFFFFFFFFu >> (31 - j_hi));
        }
        // the grid has one position more than there are image tokens: the last query of the last block is padding
        if ((hb >> 1) == g.n_img_blocks - 1 && (hb & 1) == 1 && ch == 1) vm &= 0x7FFFFFFFu;
      }
    }
    uint32_t rs[32], rp[32];
    if (__any_sync(0xFFFFFFFFu, vm != 0u)) {
      tmem_ld32(trow + AB4_TM_S + 32 * ch, rs);
      tmem_ld32(trow + AB4_TM_DP + 32 * ch, rp);
      const float4* l4 = reinterpret_cast<const float4*>(s_lse + (u & 1) * 64 + 32 * ch);
      const float4* d4 = reinterpret_cast<const float4*>(s_del + (u & 1) * 64 + 32 * ch);
      tmem_ld_wait();
      const bool all_vis = __all_sync(0xFFFFFFFFu, vm == 0xFFFFFFFFu);
#pragma unroll
      for (int j4 = 0; j4 < 8; ++j4) {
        const float4 ls = l4[j4], dl = d4[j4];
        const float lsv[4] = {ls.x, ls.y, ls.z, ls.w}, dlv[4] = {dl.x, dl.y, dl.z, dl.w};
        float pj[4], dj[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int j = 4 * j4 + e;
          const float pe = fast_exp2(__uint_as_float(rs[j]) * kLog2e - lsv[e]);
          pj[e] = (all_vis || ((vm >> j) & 1u)) ? pe : 0.f;
          dj[e] = pj[e] * (__uint_as_float(rp[j]) - dlv[e]);
        }
        rs[2 * j4] = pack_bf16x2(pj[0], pj[1]);
        rs[2 * j4 + 1] = pack_bf16x2(pj[2], pj[3]);
        rp[2 * j4] = pack_bf16x2(dj[0], dj[1]);
        rp[2 * j4 + 1] = pack_bf16x2(dj[2], dj[3]);
      }
    } else {
#pragma unroll
      for (int j = 0; j < 16; ++j) { rs[j] = 0u; rp[j] = 0u; }
    }
    // P^T / dS^T (bf16) over the first half of this thread's own S^T / dP^T columns
    tmem_st16_h(trow + AB4_TM_S + 32 * ch, rs);
    tmem_st16_h(trow + AB4_TM_DP + 32 * ch, rp);
    tmem_st_wait();
    tc_fence_before();
    mbar_arrive(bar_p);
    if (threadIdx.x < 128 && u + 1 < n_steps) (threadIdx.x < 64 ? s_lse : s_del)[((u + 1) & 1) * 64 + (threadIdx.x & 63)] = stat_next;
    if (warp == 0 && elect_one()) {
      issue_acc(u);
      if (u + 1 < n_steps) issue_s(u + 1);
    }
    __syncwarp();
  }

  // ---- dK / dV of the tile: inverse rotary, bf16, out through the (dead) key / value tiles
  mbar_wait(bar_acc, 0);
  tc_fence_after();
#pragma unroll 1
  for (int sect = 0; sect < 2; ++sect) {   // 0: dK (ACC0), 1: dV (ACC1)
    uint32_t r[32];
    tmem_ld32(trow + (sect == 0 ? AB4_TM_ACC0 : AB4_TM_ACC1) + 32 * ch, r);
    tmem_ld_wait();
    float v[32];
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
    if (ktok >= 0) ab4_inv_rotary32(v, a.rot + static_cast<long long>(ktok) * 32, 32 * ch, 1.0f);
    ab4_store_rows64((sect == 0 ? sK : sV) + warp * 2048, v, lane, ktok >= 0 ? row_base + ktok : -1, a.dqkv, 3 * HD,
                     (sect + 1) * HD + h * 64 + 32 * ch);
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem, 256);
  }
}

// leftover text keys: kv_acc rows (rot domain, fp32) -> inverse rotary -> bf16 k / v gradients; clears the rows.
// thread = one rotary pair of one (batch, key)
__global__ void attn_bwd4_tail_kernel(float* __restrict__ kv_acc, __nv_bfloat16* __restrict__ dqkv,
                                      const float2* __restrict__ rot_t, int rot_len, int B, int seq, int tail0,
                                      int n_tail, int HD) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  const int pairs = HD;   // 2 HD columns / 2
  if (idx >= B * n_tail * pairs) return;
  const int pc = idx % pairs, bt = idx / pairs;
  const int t = tail0 + bt % n_tail, b = bt / n_tail;
  const long long row = static_cast<long long>(b) * seq + t;
  float2* src = reinterpret_cast<float2*>(kv_acc + row * (2 * HD)) + pc;
  const float2 x = *src;
  *src = make_float2(0.f, 0.f);
  const float2 cs = __ldg(rot_t + static_cast<long long>(pc & 31) * rot_len + t);
  __nv_bfloat162 o = __floats2bfloat162_rn(x.x * cs.x + x.y * cs.y, x.y * cs.x - x.x * cs.y);
  *reinterpret_cast<__nv_bfloat162*>(dqkv + row * (3 * HD) + HD + 2 * pc) = o;
}

template <int MODE>
static int launch_bwd4(const CUtensorMap* m, const AttnBwd4Args& a, cudaStream_t st) {
  auto kq = attn_bwd4_q_kernel<MODE, false>;
  auto kqt = attn_bwd4_q_kernel<ATT_ROW, true>;
  auto kkv = attn_bwd4_kv_kernel<MODE>;
  static bool attr_done = false;
  if (!attr_done) {
    if (cudaFuncSetAttribute(kq, cudaFuncAttributeMaxDynamicSharedMemorySize, AB4_SMEM) != cudaSuccess) return -21;
    if (cudaFuncSetAttribute(kqt, cudaFuncAttributeMaxDynamicSharedMemorySize, AB4_SMEM) != cudaSuccess) return -21;
    cudaFuncSetAttribute(kqt, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    if (cudaFuncSetAttribute(kkv, cudaFuncAttributeMaxDynamicSharedMemorySize, AB4_SMEM) != cudaSuccess) return -21;
    cudaFuncSetAttribute(kq, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    cudaFuncSetAttribute(kkv, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    attr_done = true;
  }
  // m: 0 Q128, 1 Q64, 2 Qtail, 3 Qcol, 4 DO128, 5 DO64, 6 DOcol
  // the text-query and image-query kernels touch disjoint query rows (their leftover-key partials meet through
  // red.add): side by side on a forked stream; the key/value kernel needs delta from both and runs after the join
  AttnFork* fk = attn_fork_state(3);
  const cudaStream_t side = attn_fork(fk, st);
  dim3 gqt(a.g.n_text_blocks, a.g.H, a.g.B);
  kqt<<<gqt, AB4_THREADS, AB4_SMEM, side>>>(m[0], m[1], m[2], m[3], m[4], m[6], a);
  const bool bad_t = cudaGetLastError() != cudaSuccess;
  dim3 gq(a.g.n_img_blocks, a.g.H, a.g.B);
  kq<<<gq, AB4_THREADS, AB4_SMEM, st>>>(m[0], m[1], m[2], m[3], m[4], m[6], a);
  const bool bad_i = cudaGetLastError() != cudaSuccess;
  attn_join(fk, side, st);
  if (bad_t || bad_i) return -22;
  dim3 gkv((a.n_text_tiles + a.g.n_img_blocks) * a.g.H * a.g.B);
  kkv<<<gkv, AB4_THREADS, AB4_SMEM, st>>>(m[0], m[1], m[3], m[5], m[6], a);
  if (cudaGetLastError() != cudaSuccess) return -22;
  if (a.n_tail > 0) {
    const int total = a.g.B * a.n_tail * a.g.H * 64;
    attn_bwd4_tail_kernel<<<(total + 255) / 256, 256, 0, st>>>(a.kv_acc, a.dqkv, a.rot_t, a.rot_len, a.g.B, a.g.seq,
                                                               a.tail0, a.n_tail, a.g.H * 64);
    if (cudaGetLastError() != cudaSuccess) return -23;
  }
  return 0;
}

// 0 = launched, 1 = not covered (caller falls back), < 0 = error
int attn4_bwd(const void* qkv, const void* out, const void* d_out, const float* lse, float* delta, void* dqkv,
              float* kv_acc, const void* rot, const void* rot_t, float q_scale, int B, int H, int seq, int T, int F, int mode,
              cudaStream_t st) {
  if (mode != ATT_ROW && mode != ATT_COL) return 1;
  AttnBwd4Args a;
  AttnGeom& g = a.g;
  g.B = B; g.H = H; g.seq = seq; g.T = T; g.T_pad = (T + 15) / 16 * 16; g.F = F; g.P = F * F;
  g.n_text_blocks = (T + 127) / 128; g.n_img_blocks = g.P / 128; g.conv_half = 0;
  if (g.P % 128 || 32 % F || seq != T + g.P - 1) return 1;
  const int HD = H * 64;
  a.lse = lse; a.delta = delta; a.out = reinterpret_cast<const __nv_bfloat16*>(out);
  a.dqkv = reinterpret_cast<__nv_bfloat16*>(dqkv); a.kv_acc = kv_acc;
  a.rot = reinterpret_cast<const float2*>(rot);
  a.rot_t = reinterpret_cast<const float2*>(rot_t); a.rot_len = seq + 1; a.q_scale = q_scale;
  a.n_text_chunks = (g.T_pad + 63) / 64;
  // same block shift as the forward (attn_fwd4_sm100.cu): T = 128 k + 1 -> query blocks start at token 1
  a.text_shift = (T > 1 && T % 128 == 1) ? 1 : 0;
  g.n_text_blocks = (T - a.text_shift + 127) / 128;
  a.qkv = reinterpret_cast<const __nv_bfloat16*>(qkv); a.d_out = reinterpret_cast<const __nv_bfloat16*>(d_out);
  // text keys beyond the full 128-key tiles: a handful go to the CUDA-core path of the q kernel (they must then also
  // be the leading keys of its last 64-key chunk), more get a partial tile of their own
  const int rem = T % 128;
  if (rem > 0 && rem <= AB4_MAX_TAIL && (T - rem) % 64 == 0) { a.n_tail = rem; a.tail0 = T - rem; a.n_text_tiles = T / 128; }
  else { a.n_tail = 0; a.tail0 = T; a.n_text_tiles = (T + 127) / 128; }
  const uint64_t rows = (uint64_t)B * seq, cols = 3ull * HD;
  CUtensorMap m[7];
  int rc = make_tmap_bf16_2d(&m[0], qkv, rows, cols, cols, 64, 128);
  if (!rc) rc = make_tmap_bf16_2d(&m[1], qkv, rows, cols, cols, 64, 64);
  const int tail = g.T_pad % 64;
  if (!rc) rc = make_tmap_bf16_2d(&m[2], qkv, rows, cols, cols, 64, tail ? tail : 64);
  if (!rc) rc = make_tmap_bf16_2d(&m[4], d_out, rows, HD, HD, 64, 128);
  if (!rc) rc = make_tmap_bf16_2d(&m[5], d_out, rows, HD, HD, 64, 64);
  if (rc) return rc;
  {
    const __nv_bfloat16* base = reinterpret_cast<const __nv_bfloat16*>(qkv) + (size_t)T * cols;
    uint64_t dims[4] = {cols, (uint64_t)F, (uint64_t)F, (uint64_t)B};
    uint64_t strides[3] = {cols * 2, cols * 2 * F, cols * 2 * seq};
    uint32_t box[4] = {64, 1, (uint32_t)F, 1};
    rc = make_tmap_bf16_nd(&m[3], base, 4, dims, strides, box);
    if (rc) return rc;
    const __nv_bfloat16* dbase = reinterpret_cast<const __nv_bfloat16*>(d_out) + (size_t)T * HD;
    uint64_t ddims[4] = {(uint64_t)HD, (uint64_t)F, (uint64_t)F, (uint64_t)B};
    uint64_t dstrides[3] = {(uint64_t)HD * 2, (uint64_t)HD * 2 * F, (uint64_t)HD * 2 * seq};
    rc = make_tmap_bf16_nd(&m[6], dbase, 4, ddims, dstrides, box);
    if (rc) return rc;
  }
  return mode == ATT_ROW ? launch_bwd4<ATT_ROW>(m, a, st) : launch_bwd4<ATT_COL>(m, a, st);
}

}  // namespace b200
