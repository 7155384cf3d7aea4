// Sparse DALL-E attention forward, third generation (text, axial_row, axial_col):
// a persistent, warp-specialised, software-pipelined kernel.
//
//   warp 8      query loader     Q tiles into a ring of four 16 KB slots, up to three blocks ahead of the math
//   warp 9      MMA issuer       S = Q K^T and O += P V for two query blocks in flight (two TMEM stages)
//   warp 10     local loader     the 128 local keys / values of an image query block (one buffer per TMEM stage)
//   warp 11     text loader      the text keys / values of the item's (batch, head): loaded ONCE per item, in pieces
//                                (128-key tiles + tail) that are released one by one, so the next item's pieces
//                                arrive while the current item's last blocks are still being computed
//   warps 0-3   softmax group 0  (TMEM stage 0)      warps 4-7   softmax group 1  (TMEM stage 1)
// The single-thread roles sit in the HIGHEST warps on purpose: the warp scheduler of an SM sub-partition prefers the
// highest warp id among eligible warps, and with the issuer in warp 1 (below the math warps it shares a scheduler
// with) each S or P V took 1000+ cycles to issue (profiles/attn3_fwd_timeline_r2.txt).
//
// Scores are read from TMEM twice (row maximum, then probabilities); P goes back into the TMEM columns of S as bf16
// and is the A operand of the P V MMA straight from TMEM - no shared-memory round trip, no proxy fence.  O stays in
// TMEM for the whole block; the running maximum is only raised (and O rescaled in TMEM) when a chunk's maximum
// exceeds it by more than 2^8, so the rescale is rare and the result is exact either way (P and the row sum use the
// same reference).  A completed S commit also proves that the previous P V of the stage is complete (commits cover
// every earlier MMA of the issuing thread), which is what makes the in-place O rescale safe without another barrier.
#include "attn3_common.cuh"
#include "b200_api.h"

#include <cstdio>
#include <cstdlib>

namespace b200 {

int attn_make_maps(const void* qkv, const AttnGeom& g, int mode, CUtensorMap* t128, CUtensorMap* ttail,
                   CUtensorMap* taux);

constexpr int A3F_THREADS = 384;   // 2 softmax groups + 4 single-thread roles
constexpr int A3F_TEXT_BYTES = 272 * 128;            // text keys (or values) of one (batch, head), padded
constexpr int A3F_NQ = 4;                            // query ring slots
constexpr int A3F_SMEM_DATA = 2 * A3F_TEXT_BYTES + A3F_NQ * 16384 + 4 * 16384;
constexpr int A3F_SMEM = A3F_SMEM_DATA + 512 + 1024;
constexpr uint32_t A3F_STAGE_COLS = 208, A3F_TM_O = 144;
constexpr float A3F_RESCALE_LOG2 = 8.0f;

struct Attn3FwdArgs {
  Attn3Plan p;
  __nv_bfloat16* out;
  float* lse;
  int ld_out;
  unsigned long long* trace;   // optional timeline of CTA 0 (tools/attn3_trace.py): 4 roles x 2048 (tag, clock) pairs
};

// role 0: MMA issuer, 1 / 2: softmax groups (first lane), 3: block loader
__device__ __forceinline__ void a3_trace(unsigned long long* tr, int role, uint32_t& n, uint32_t tag) {
  if (tr != nullptr && blockIdx.x == 0 && n < 2048u) {
    tr[(role * 2048 + n) * 2] = tag;
    tr[(role * 2048 + n) * 2 + 1] = clock64();
    ++n;
  }
}

struct A3FBars {
  uint64_t tk_full[3], tk_free[3], tv_full[3], tv_free[3];
  uint64_t q_full[A3F_NQ], q_free[A3F_NQ];
  uint64_t l_full[2], l_free[2];
  uint64_t s_full[2], p_full[2], o_full[2], o_free[2];
  uint32_t tmem_slot;
};

template <int MODE>   // ATT_ROW or ATT_COL: how the image query blocks enumerate the grid
__global__ void __launch_bounds__(A3F_THREADS, 1)
attn3_fwd_kernel(const __grid_constant__ CUtensorMap tm128, const __grid_constant__ CUtensorMap tmTail,
                 const __grid_constant__ CUtensorMap tmCol, const Attn3FwdArgs a) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + (((raw_addr + 1023u) & ~1023u) - raw_addr);
  const Attn3Plan& p = a.p;
  const AttnGeom& g = p.g;
  uint8_t* sTK = smem;
  uint8_t* sTV = sTK + A3F_TEXT_BYTES;
  uint8_t* sQ = sTV + A3F_TEXT_BYTES;          // A3F_NQ x 16 KB
  uint8_t* sLK = sQ + A3F_NQ * 16384;          // 2 x 16 KB
  uint8_t* sLV = sLK + 2 * 16384;              // 2 x 16 KB
  A3FBars* bars = reinterpret_cast<A3FBars*>(sLV + 2 * 16384);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int HD = g.H * 64;
  const int lines = 128 / g.F;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tm128);
    tma_prefetch_desc(&tmTail);
    if (MODE == ATT_COL) tma_prefetch_desc(&tmCol);
    for (int i = 0; i < 3; ++i) {
      mbar_init(&bars->tk_full[i], 1); mbar_init(&bars->tk_free[i], 1);
      mbar_init(&bars->tv_full[i], 1); mbar_init(&bars->tv_free[i], 1);
    }
    for (int i = 0; i < A3F_NQ; ++i) { mbar_init(&bars->q_full[i], 1); mbar_init(&bars->q_free[i], 1); }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&bars->l_full[i], 1); mbar_init(&bars->l_free[i], 129);   // issuer's commit + the group's threads
      mbar_init(&bars->s_full[i], 1); mbar_init(&bars->p_full[i], 128);
      mbar_init(&bars->o_full[i], 1); mbar_init(&bars->o_free[i], 128);
    }
    fence_barrier_init();
  }
  if (warp == 9) {
    tmem_alloc(&bars->tmem_slot, 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = bars->tmem_slot;

  if (warp == 8) {
    // ------------------------------------------------------------------ query loader
    if (elect_one()) {
      uint32_t ntr = 0;
      A3Iter bi;
      bi.init(p, blockIdx.x, gridDim.x);
      for (; bi.valid; bi.next(p)) {
        const int e = bi.g % A3F_NQ, u = bi.g / A3F_NQ;
        if (u > 0) mbar_wait_timed(&bars->q_free[e], (u - 1) & 1);
        a3_trace(a.trace, 3, ntr, (7u << 24) | bi.g);
        int b, h, blk;
        bi.where(p, b, h, blk);
        uint8_t* dst = sQ + e * 16384;
        mbar_expect_tx(&bars->q_full[e], 16384);
        if (bi.is_text(p)) {
          tma_load_2d(dst, &tm128, &bars->q_full[e], h * 64, b * g.seq + blk * 128);
        } else if (MODE == ATT_COL) {
          for (int l = 0; l < lines; ++l)
            tma_load_4d(dst + l * g.F * 128, &tmCol, &bars->q_full[e], h * 64, blk * lines + l, 0, b);
        } else {
          tma_load_2d(dst, &tm128, &bars->q_full[e], h * 64, b * g.seq + g.T + blk * 128);
        }
      }
    }
  } else if (warp == 10) {
    // ------------------------------------------------------------------ local key / value loader
    if (elect_one()) {
      A3Iter bi;
      bi.init(p, blockIdx.x, gridDim.x);
      uint32_t nl0 = 0, nl1 = 0;
      for (; bi.valid; bi.next(p)) {
        if (bi.is_text(p)) continue;
        const int s = bi.g & 1;
        const uint32_t n = s ? nl1 : nl0;
        if (n > 0) mbar_wait_timed(&bars->l_free[s], (n - 1) & 1);
        if (s) ++nl1; else ++nl0;
        int b, h, blk;
        bi.where(p, b, h, blk);
        uint8_t* dk = sLK + s * 16384;
        uint8_t* dv = sLV + s * 16384;
        const int kcol = HD + h * 64, vcol = 2 * HD + h * 64;
        mbar_expect_tx(&bars->l_full[s], 32768);
        if (MODE == ATT_COL) {
          for (int l = 0; l < lines; ++l) {
            tma_load_4d(dk + l * g.F * 128, &tmCol, &bars->l_full[s], kcol, blk * lines + l, 0, b);
            tma_load_4d(dv + l * g.F * 128, &tmCol, &bars->l_full[s], vcol, blk * lines + l, 0, b);
          }
        } else {
          tma_load_2d(dk, &tm128, &bars->l_full[s], kcol, b * g.seq + g.T + blk * 128);
          tma_load_2d(dv, &tm128, &bars->l_full[s], vcol, b * g.seq + g.T + blk * 128);
        }
      }
    }
  } else if (warp == 11) {
    // ------------------------------------------------------------------ text key / value loader (once per item)
    if (elect_one()) {
      int k = 0;
      for (int it = blockIdx.x; it < p.n_items; it += gridDim.x, ++k) {
        const int bh = it < p.n_img_items ? it / p.parts_img : it - p.n_img_items;
        const int b = bh / g.H, h = bh - b * g.H;
        const int kcol = HD + h * 64, vcol = 2 * HD + h * 64;
        for (int pc = 0; pc < p.nkc; ++pc) {
          const int rows = a3_piece_rows(p, pc);
          const CUtensorMap* tm = rows == 128 ? &tm128 : &tmTail;
          if (k > 0) mbar_wait_timed(&bars->tk_free[pc], (k - 1) & 1);
          mbar_expect_tx(&bars->tk_full[pc], rows * 128);
          tma_load_2d(sTK + pc * 16384, tm, &bars->tk_full[pc], kcol, b * g.seq + pc * 128);
          if (k > 0) mbar_wait_timed(&bars->tv_free[pc], (k - 1) & 1);
          mbar_expect_tx(&bars->tv_full[pc], rows * 128);
          tma_load_2d(sTV + pc * 16384, tm, &bars->tv_full[pc], vcol, b * g.seq + pc * 128);
        }
      }
    }
  } else if (warp == 9) {
    // ------------------------------------------------------------------ MMA issuer
    // One thread issues every MMA of the CTA, and a lone thread retires an instruction only every few cycles, so
    // this loop is kept lean (the first version spent ~1700 cycles per S or P V, three times the tensor time:
    // profiles/attn3_fwd_timeline_r2.txt): descriptors are precomputed and advanced by adding to their address
    // field, the per-stage state lives in registers (the two stages are two instances of one struct, never an
    // array indexed at run time), the piece use counters are bit fields of one register.  The thread is chosen
    // with elect.sync rather than `lane == 0`: the compiler then knows the region is single-threaded and keeps the
    // MMA operands in uniform registers instead of wrapping every tcgen05.mma in a make-uniform loop.
    if (elect_one()) {
      struct Stage {
        A3Iter bi;
        int c, nc;                 // current chunk, chunks of the block
        int piece, ntext, local;   // current chunk
        uint32_t np, nb, nl;       // phases consumed of p_full / o_free / l_full
        bool active, want_pv, l_ready;
      };
      Stage st0, st1;
      // [item parity][piece] remaining reads, 5 bits each: keys in the low 15 bits of a half, values in the high
      uint32_t k_left = 0u, v_left = 0u, seen = 0u;   // seen: bit (parity * 8 + piece) keys, (+ 4) values
      int slot_k0 = -1, slot_k1 = -1;
      const uint64_t dq0 = make_smem_desc_sw128(smem_u32(sQ), 0, 1024);
      const uint64_t dtk0 = make_smem_desc_sw128(smem_u32(sTK), 0, 1024);
      const uint64_t dlk0 = make_smem_desc_sw128(smem_u32(sLK), 0, 1024);
      const uint64_t dtv0 = make_smem_desc_sw128(smem_u32(sTV), 8192, 1024);
      const uint64_t dlv0 = make_smem_desc_sw128(smem_u32(sLV), 8192, 1024);
      constexpr uint32_t idesc_pv = make_idesc_bf16(128, 64, 0, 1);
      constexpr uint32_t idesc_s128 = make_idesc_bf16(128, 128, 0, 0);
      uint32_t ntr = 0;

      auto set_chunk = [&](Stage& st) {
        const A3Chunk d = st.bi.chunk(p, st.c);
        st.piece = d.piece; st.ntext = d.ntext; st.local = d.has_local;
      };
      auto start_block = [&](Stage& st) {
        st.active = st.bi.valid;
        if (!st.active) return;
        const int k = st.bi.k, sl = k & 1;
        if ((sl ? slot_k1 : slot_k0) != k) {
          if (sl) slot_k1 = k; else slot_k0 = k;
          uint32_t f = 0u;
          for (int pc = 0; pc < 3; ++pc) f |= static_cast<uint32_t>(st.bi.piece_uses(p, pc)) << (5 * pc);
          const uint32_t keep = sl ? 0x00007FFFu : 0x3FFF8000u;
          k_left = (k_left & keep) | (f << (15 * sl));
          v_left = (v_left & keep) | (f << (15 * sl));
          seen &= sl ? 0x00FFu : 0xFF00u;
        }
        st.c = 0;
        st.nc = st.bi.nchunks(p);
        st.l_ready = false;
        st.want_pv = false;
        set_chunk(st);
      };
      // The two stages are served in strict alternation with blocking waits.  (A polling issuer that let one stage
      // run ahead deadlocked against in-order loaders, and its probing cost more than the MMAs; with alternation
      // the stages stay within a block of each other and every load is consumed in the order it is issued.)
      auto issue_s = [&](Stage& st, const int s) {
        const int g = st.bi.g, k = st.bi.k, sl = k & 1;
        const int e = g % A3F_NQ;
        const uint32_t tm_s = tmem + s * A3F_STAGE_COLS;
        if (st.c == 0) mbar_wait_timed(&bars->q_full[e], (g / A3F_NQ) & 1);
        if (st.ntext > 0 && !((seen >> (sl * 8 + st.piece)) & 1u)) {
          mbar_wait_timed(&bars->tk_full[st.piece], k & 1);
          seen |= 1u << (sl * 8 + st.piece);
        }
        if (st.local && !st.l_ready) {
          mbar_wait_timed(&bars->l_full[s], st.nl & 1);
          st.l_ready = true;
        }
        tc_fence_after();
        const uint64_t dq = dq0 + static_cast<uint64_t>(e * 1024);
        if (st.ntext > 0) {
          const uint64_t dk = dtk0 + static_cast<uint64_t>(st.piece * 1024);
          const uint32_t idesc = st.ntext == 128 ? idesc_s128 : make_idesc_bf16(128, st.ntext, 0, 0);
#pragma unroll
          for (int kk = 0; kk < 4; ++kk) umma_bf16_ss(tm_s, dq + 2 * kk, dk + 2 * kk, idesc, kk > 0 ? 1u : 0u);
          const int sh = 15 * sl + 5 * st.piece;
          k_left -= 1u << sh;
          if (((k_left >> sh) & 31u) == 0u) umma_commit(&bars->tk_free[st.piece]);
        }
        if (st.local) {
          const uint64_t dk = dlk0 + static_cast<uint64_t>(s * 1024);
#pragma unroll
          for (int kk = 0; kk < 4; ++kk)
            umma_bf16_ss(tm_s + st.ntext, dq + 2 * kk, dk + 2 * kk, idesc_s128, kk > 0 ? 1u : 0u);
        }
        umma_commit(&bars->s_full[s]);
        if (st.c == st.nc - 1) umma_commit(&bars->q_free[e]);   // last read of this query tile
        a3_trace(a.trace, 0, ntr, (1u << 24) | (s << 20) | (st.c << 16) | (g & 0xFFFF));
      };
      auto turn = [&](Stage& st, const int s) {
        const int g = st.bi.g, k = st.bi.k, sl = k & 1;
        const uint32_t tm_s = tmem + s * A3F_STAGE_COLS;
        const uint32_t tm_o = tm_s + A3F_TM_O;
        // ---- O (+)= P V of chunk c
        mbar_wait_timed(&bars->p_full[s], st.np & 1);
        ++st.np;
        if (st.c == 0 && st.nb > 0) mbar_wait_timed(&bars->o_free[s], (st.nb - 1) & 1);
        if (st.ntext > 0 && !((seen >> (sl * 8 + 4 + st.piece)) & 1u)) {
          mbar_wait_timed(&bars->tv_full[st.piece], k & 1);
          seen |= 1u << (sl * 8 + 4 + st.piece);
        }
        tc_fence_after();
        uint32_t acc = st.c > 0 ? 1u : 0u;
        if (st.ntext > 0) {
          const uint64_t dv = dtv0 + static_cast<uint64_t>(st.piece * 1024);
          if (st.ntext == 128) {
#pragma unroll
            for (int k16 = 0; k16 < 8; ++k16) {
              umma_bf16_ts(tm_o, tm_s + k16 * 8, dv + 128 * k16, idesc_pv, acc);
              acc = 1u;
            }
          } else {
            for (int k16 = 0; k16 < st.ntext / 16; ++k16) {
              umma_bf16_ts(tm_o, tm_s + k16 * 8, dv + 128 * k16, idesc_pv, acc);
              acc = 1u;
            }
          }
          const int sh = 15 * sl + 5 * st.piece;
          v_left -= 1u << sh;
          if (((v_left >> sh) & 31u) == 0u) umma_commit(&bars->tv_free[st.piece]);
        }
        if (st.local) {
          const uint64_t dv = dlv0 + static_cast<uint64_t>(s * 1024);
          const uint32_t tm_p = tm_s + st.ntext / 2;
#pragma unroll
          for (int k16 = 0; k16 < 8; ++k16) {
            umma_bf16_ts(tm_o, tm_p + k16 * 8, dv + 128 * k16, idesc_pv, acc);
            acc = 1u;
          }
        }
        a3_trace(a.trace, 0, ntr, (2u << 24) | (s << 20) | (st.c << 16) | (g & 0xFFFF));
        if (st.c == st.nc - 1) {
          if (!st.bi.is_text(p)) { umma_commit(&bars->l_free[s]); ++st.nl; }
          umma_commit(&bars->o_full[s]);   // issued last: nothing may still signal this CTA once its groups leave
          ++st.nb;
          st.bi.next(p);
          st.bi.next(p);
          start_block(st);
          if (st.active) issue_s(st, s);
        } else {
          ++st.c;
          set_chunk(st);
          issue_s(st, s);
        }
      };

      st0.np = st0.nb = st0.nl = 0u;
      st1.np = st1.nb = st1.nl = 0u;
      st0.bi.init(p, blockIdx.x, gridDim.x);
      st1.bi = st0.bi;
      st1.bi.next(p);
      start_block(st0);
      if (st0.active) issue_s(st0, 0);
      start_block(st1);
      if (st1.active) issue_s(st1, 1);
      while (st0.active || st1.active) {
        if (st0.active) turn(st0, 0);
        if (st1.active) turn(st1, 1);
      }
    }
  } else {
    // ------------------------------------------------------------------ softmax groups (one query row per thread)
    const int wg = warp >> 2;
    const int qw = warp & 3;
    const int i = qw * 32 + lane;
    const uint32_t trow = tmem + wg * A3F_STAGE_COLS + (static_cast<uint32_t>(qw * 32) << 16);
    const int win0 = qw * 32;
    uint32_t ns = 0, nblk_done = 0, ntr = 0;
    unsigned long long* const tr = (qw == 0 && lane == 0) ? a.trace : nullptr;
    A3Iter bi;
    bi.init(p, blockIdx.x, gridDim.x);
    if (wg == 1) bi.next(p);
    for (; bi.valid; bi.next(p), bi.next(p)) {
      int b, h, blk;
      bi.where(p, b, h, blk);
      const bool is_text = bi.is_text(p);
      int tpos;
      uint32_t lmask = 0u;
      if (is_text) {
        tpos = blk * 128 + i;
      } else {
        const int ip = img_query_pos<MODE>(g, blk, i);
        const int ir = ip / g.F, ic = ip - ir * g.F;
        tpos = g.T + ip;
        // local keys of this row: its own line (row or column) up to itself, inside the warp's 32-key window
        const int lo = max(i - (MODE == ATT_ROW ? ic : ir), win0) - win0;
        lmask = (0xFFFFFFFFu >> (31 - lane)) & (0xFFFFFFFFu << lo);
      }
      const bool q_valid = is_text ? (tpos < g.T) : (tpos < g.seq);
      const int text_limit = q_valid ? (is_text ? tpos + 1 : g.T) : 0;
      if (!q_valid) lmask = 0u;
      const int nchunks = bi.nchunks(p);
      float m_used = -1e30f, l_run = 0.f;

      for (int c = 0; c < nchunks; ++c) {
        const A3Chunk d = bi.chunk(p, c);
        a3_trace(tr, 1 + wg, ntr, (3u << 24) | (c << 16) | (bi.g & 0xFFFF));
        mbar_wait_timed(&bars->s_full[wg], ns & 1);
        ++ns;
        tc_fence_after();
        a3_trace(tr, 1 + wg, ntr, (4u << 24) | (c << 16) | (bi.g & 0xFFFF));
        // The code of this loop is kept SMALL on purpose (one site for the exponentials, loops instead of four-fold
        // unrolling): at 100 KB the first version of this kernel thrashed the instruction caches, which a math
        // warpgroup hides but the lone issuer thread does not - its straight-line bookkeeping then ran at tens of
        // cycles per instruction (profiles/attn3_fwd_timeline_r2.txt).
        // ---- pass 1: row maximum; all TMEM loads of the chunk are issued before a single wait
        float cmax = -1e30f;
        if (d.ntext == 128) {
          uint32_t r0[32], r1[32], r2[32], r3[32];
          tmem_ld32(trow, r0);
          tmem_ld32(trow + 32, r1);
          tmem_ld32(trow + 64, r2);
          tmem_ld32(trow + 96, r3);
          tmem_ld_wait();
          const int nv = text_limit - d.row0;
          if (__all_sync(0xFFFFFFFFu, nv >= 128)) {
            float c0 = -1e30f, c1 = -1e30f;
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              c0 = fmaxf(c0, fmaxf(__uint_as_float(r0[j]), __uint_as_float(r1[j])));
              c1 = fmaxf(c1, fmaxf(__uint_as_float(r2[j]), __uint_as_float(r3[j])));
            }
            cmax = fmaxf(c0, c1);
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              cmax = fmaxf(cmax, j < nv ? __uint_as_float(r0[j]) : -1e30f);
              cmax = fmaxf(cmax, j + 32 < nv ? __uint_as_float(r1[j]) : -1e30f);
              cmax = fmaxf(cmax, j + 64 < nv ? __uint_as_float(r2[j]) : -1e30f);
              cmax = fmaxf(cmax, j + 96 < nv ? __uint_as_float(r3[j]) : -1e30f);
            }
          }
        } else {
          // short chunks (text tail and / or local window): at most two loads
#pragma unroll 1
          for (int g0 = 0; g0 < d.ntext; g0 += 32) {
            uint32_t r[32];
            const int gw = min(32, d.ntext - g0);
            const int nv = min(text_limit - (d.row0 + g0), gw);
            if (gw == 32) tmem_ld32(trow + g0, r);
            else tmem_ld16_h(trow + g0, r);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 32; ++j) cmax = fmaxf(cmax, j < nv ? __uint_as_float(r[j]) : -1e30f);
          }
        }
        if (d.has_local) {
          uint32_t r[32];
          tmem_ld32(trow + d.ntext + win0, r);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 32; ++j) cmax = fmaxf(cmax, ((lmask >> j) & 1u) ? __uint_as_float(r[j]) : -1e30f);
        }
        // ---- raise the reference maximum only when it would otherwise let P grow beyond 2^8
        const float m_c = cmax * kLog2e;
        const bool need = q_valid && (m_c > m_used + A3F_RESCALE_LOG2);
        float factor = 1.0f;
        if (need) {
          factor = fast_exp2(m_used - m_c);
          m_used = m_c;
          l_run *= factor;
        }
        if (c > 0 && __any_sync(0xFFFFFFFFu, need)) {
          // the P V of chunk c - 1 is complete (see the header); none is issued before this thread arrives on p_full
#pragma unroll 1
          for (int half = 0; half < 2; ++half) {
            uint32_t o[32];
            tmem_ld32(trow + A3F_TM_O + half * 32, o);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 32; ++j) o[j] = __float_as_uint(__uint_as_float(o[j]) * factor);
            tmem_st32(trow + A3F_TM_O + half * 32, o);
          }
        }
        // ---- pass 2: probabilities (bf16, packed in place) into the columns of S, row sum.  One loop over groups of
        // 32 columns; the load of the next group is in flight while the current one is exponentiated, and a P store
        // only overwrites S columns that are already in registers.
        float lsum0 = 0.f, lsum1 = 0.f;
        {
          uint32_t cur[32], nxt[32];
          const int ngrp = (d.ntext + 31) >> 5;
          if (ngrp > 0) {
            if (d.ntext >= 32) tmem_ld32(trow, nxt);
            else tmem_ld16_h(trow, nxt);
          }
#pragma unroll 1
          for (int gi = 0; gi < ngrp; ++gi) {
            const int g0 = gi * 32;
            const int gw = min(32, d.ntext - g0);
            const int nv = min(text_limit - (d.row0 + g0), gw);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 32; ++j) cur[j] = nxt[j];
            if (gi + 1 < ngrp) {
              if (d.ntext - (g0 + 32) >= 32) tmem_ld32(trow + g0 + 32, nxt);
              else tmem_ld16_h(trow + g0 + 32, nxt);
            }
            if (__all_sync(0xFFFFFFFFu, nv >= 32)) {
#pragma unroll
              for (int j = 0; j < 32; j += 2) {
                const float e0 = fast_exp2(__uint_as_float(cur[j]) * kLog2e - m_used);
                const float e1 = fast_exp2(__uint_as_float(cur[j + 1]) * kLog2e - m_used);
                lsum0 += e0; lsum1 += e1;
                cur[j >> 1] = pack_bf16x2(e0, e1);
              }
            } else {
#pragma unroll
              for (int j = 0; j < 32; j += 2) {
                const float e0 = j < nv ? fast_exp2(__uint_as_float(cur[j]) * kLog2e - m_used) : 0.f;
                const float e1 = j + 1 < nv ? fast_exp2(__uint_as_float(cur[j + 1]) * kLog2e - m_used) : 0.f;
                lsum0 += e0; lsum1 += e1;
                cur[j >> 1] = pack_bf16x2(e0, e1);
              }
            }
            if (gw == 32) tmem_st16_h(trow + (g0 >> 1), cur);
            else tmem_st8_h(trow + (g0 >> 1), cur);
          }
          if (d.has_local) {
            tmem_ld32(trow + d.ntext + win0, cur);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 32; j += 2) {
              const float e0 = ((lmask >> j) & 1u) ? fast_exp2(__uint_as_float(cur[j]) * kLog2e - m_used) : 0.f;
              const float e1 = ((lmask >> (j + 1)) & 1u) ? fast_exp2(__uint_as_float(cur[j + 1]) * kLog2e - m_used) : 0.f;
              lsum0 += e0; lsum1 += e1;
              cur[j >> 1] = pack_bf16x2(e0, e1);
            }
#pragma unroll
            for (int j = 0; j < 16; ++j) nxt[j] = 0u;
            const uint32_t pl = trow + (d.ntext >> 1);
            // the local S columns of this lane (its warp's window) are in registers: the P columns may go over them
#pragma unroll 1
            for (int w = 0; w < 4; ++w) {
              if (w == qw) tmem_st16_h(pl + w * 16, cur);
              else tmem_st16_h(pl + w * 16, nxt);
            }
          }
        }
        const float lsum = lsum0 + lsum1;
        l_run += lsum;
        tmem_st_wait();
        tc_fence_before();
        mbar_arrive(&bars->p_full[wg]);
        a3_trace(tr, 1 + wg, ntr, (5u << 24) | (c << 16) | (bi.g & 0xFFFF));
      }

      // ---- epilogue: O / l -> bf16, straight to global (one 128-byte line per thread)
      mbar_wait_timed(&bars->o_full[wg], nblk_done & 1);
      ++nblk_done;
      tc_fence_after();
      a3_trace(tr, 1 + wg, ntr, (6u << 24) | (bi.g & 0xFFFF));
      uint32_t o0[32], o1[32];
      tmem_ld32(trow + A3F_TM_O, o0);
      tmem_ld32(trow + A3F_TM_O + 32, o1);
      tmem_ld_wait();
      tc_fence_before();
      mbar_arrive(&bars->o_free[wg]);
      {
        // O / l -> bf16 through this warp's 4 KB of the stage's local-key buffer (free: every MMA of the block is
        // complete, and the buffer is handed back to the loader only after the arrivals below), then out with
        // whole 128-byte lines per store instruction: a thread writing its own row 16 bytes at a time touched 32
        // different lines per instruction and made the epilogue the longest phase of a block.
        const float inv = l_run > 0.f ? 1.0f / l_run : 0.f;
        uint8_t* stg = sLK + wg * 16384 + qw * 4096;
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
        const int my_row = q_valid ? b * g.seq + tpos : -1;
        const int cq = lane & 7;
#pragma unroll
        for (int it = 0; it < 8; ++it) {
          const int rr = it * 4 + (lane >> 3);
          const int grow = __shfl_sync(0xFFFFFFFFu, my_row, rr);
          const uint4 u = *reinterpret_cast<const uint4*>(stg + rr * 128 + ((cq ^ (rr & 7)) << 4));
          if (grow >= 0)
            *reinterpret_cast<uint4*>(a.out + static_cast<long long>(grow) * a.ld_out + h * 64 + cq * 8) = u;
        }
        __syncwarp();
        if (!is_text) mbar_arrive(&bars->l_free[wg]);
        if (q_valid) a.lse[(static_cast<long long>(b) * g.H + h) * g.seq + tpos] = m_used + log2f(l_run);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 9) {
    tc_fence_after();
    tmem_dealloc(tmem, 512);
  }
}

unsigned long long* g_attn3_trace = nullptr;

static int a3_sm_count() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
  }
  return n;
}

template <int MODE>
static int launch_fwd3(const CUtensorMap& t128, const CUtensorMap& ttail, const CUtensorMap& tcol,
                       const Attn3FwdArgs& a, cudaStream_t st) {
  auto kern = attn3_fwd_kernel<MODE>;
  static bool attr_done = false;
  if (!attr_done) {
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, A3F_SMEM) != cudaSuccess) return -21;
    attr_done = true;
  }
  const int grid = a.p.n_items < a3_sm_count() ? a.p.n_items : a3_sm_count();
  kern<<<grid, A3F_THREADS, A3F_SMEM, st>>>(t128, ttail, tcol, a);
  return cudaGetLastError() == cudaSuccess ? 0 : -22;
}

// 0 = launched, 1 = geometry not covered by this kernel (caller falls back), < 0 = error
int attn3_fwd(const void* qkv, void* out, float* lse, int B, int H, int seq, int T, int F, int mode,
              cudaStream_t st) {
  if (mode != ATT_ROW && mode != ATT_COL) return 1;
  Attn3FwdArgs a;
  if (attn3_make_plan(&a.p, B, H, seq, T, F)) return 1;
  a.out = reinterpret_cast<__nv_bfloat16*>(out); a.lse = lse; a.ld_out = H * 64;
  a.trace = g_attn3_trace;
  CUtensorMap t128, ttail, tcol;
  int rc = attn_make_maps(qkv, a.p.g, ATT_COL, &t128, &ttail, &tcol);
  if (rc) return rc;
  if (a.p.tail_pad > 0 && a.p.tail_pad != a.p.g.T_pad % 128) return 1;
  return mode == ATT_ROW ? launch_fwd3<ATT_ROW>(t128, ttail, tcol, a, st) : launch_fwd3<ATT_COL>(t128, ttail, tcol, a, st);
}

}  // namespace b200

// Developer hook (SURVEY 5.1 tracing): device buffer of 4 x 2048 x 2 u64 that CTA 0 of the pipelined attention
// forward fills with (tag, SM clock) pairs; nullptr switches it off.  Read back with tools/attn3_trace.py.
extern "C" void b200_attn3_set_trace(void* buf) { b200::g_attn3_trace = reinterpret_cast<unsigned long long*>(buf); }
