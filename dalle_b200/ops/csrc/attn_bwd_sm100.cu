// Sparse DALL-E attention backward on tcgen05 (backward halves).
//
// One CTA = one (batch, head, 128-query block); it walks the 128-key chunks that block can see
// (text chunks, then its local chunk(s)) and per chunk runs five MMAs:
//     S = Q K^T        dP = dO V^T                         (recompute / dgrad of PV)
//     P = exp2(S*log2e - lse) (masked),  dS = P * (dP - delta)   (softmax warps, TMEM -> regs -> swizzled smem)
//     dQ += dS K       dK = dS^T Q       dV = P^T dO        (accumulators in TMEM)
// dQ is written once at the end (inverse rotary + q scale applied, bf16, straight into d_qkv).
// dK/dV of an axial block's own local chunk are exclusive to the CTA -> inverse rotary + bf16 store.
// dK/dV of text keys (and of conv_like local keys) are shared between CTAs -> fp32 red.add into `kv_acc`
// (rot-domain), which b200_attn_bwd_finish converts (inverse rotary, bf16) and clears.
#include "attn_common.cuh"
#include "b200_api.h"

#include <cstdlib>

namespace b200 {

int make_tmap_bf16_2d(CUtensorMap* m, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_cols,
                      uint32_t box_rows);
int make_tmap_bf16_nd(CUtensorMap* m, const void* ptr, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                      const uint32_t* box);
int attn_make_maps(const void* qkv, const AttnGeom& g, int mode, CUtensorMap* t128, CUtensorMap* ttail,
                   CUtensorMap* taux);
int attn4_bwd(const void* qkv, const void* out, const void* d_out, const float* lse, float* delta, void* dqkv,
              float* kv_acc, const void* rot, const void* rot_t, float q_scale, int B, int H, int seq, int T, int F, int mode,
              cudaStream_t st);

constexpr int ATTB_THREADS = 320;   // TMA warp, MMA warp, 8 worker warps

struct AttnBwdArgs {
  AttnGeom g;
  const float* lse;     // [B,H,seq] log2 domain
  const __nv_bfloat16* out;  // forward output [B*seq, H*64] (delta = rowsum(dO * O) is formed in the kernel)
  __nv_bfloat16* dqkv;  // [B*seq, 3*H*64]
  float* kv_acc;        // [B*seq, 2*H*64] fp32, zero on entry
  const float2* rot;    // [seq+1][32] (cos, sin)
  const float2* rot_t;  // [32][seq+1] pair-major copy (coalesced for thread-per-row access)
  int rot_len;
  float q_scale;
};

__device__ __forceinline__ void red_add_v4(float* p, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// v[0..31] = 32 consecutive head dims starting at d0 (0 or 32) -> inverse rotation by the position's angles
__device__ __forceinline__ void inv_rotary32(float* v, const float2* rot_t_pos, int rot_len, int d0, float scale) {
#pragma unroll
  for (int j = 0; j < 32; j += 2) {
    const float2 cs = __ldg(rot_t_pos + static_cast<long long>((d0 + j) >> 1) * rot_len);
    const float x0 = v[j], x1 = v[j + 1];
    v[j] = (x0 * cs.x + x1 * cs.y) * scale;
    v[j + 1] = (x1 * cs.x - x0 * cs.y) * scale;
  }
}

__device__ __forceinline__ void store32_bf16(__nv_bfloat16* p, const float* v) {
#pragma unroll
  for (int j = 0; j < 32; j += 8) {
    uint4 u;
    u.x = pack_bf16x2(v[j], v[j + 1]);
    u.y = pack_bf16x2(v[j + 2], v[j + 3]);
    u.z = pack_bf16x2(v[j + 4], v[j + 5]);
    u.w = pack_bf16x2(v[j + 6], v[j + 7]);
    *reinterpret_cast<uint4*>(p + j) = u;
  }
}

template <int MODE>
__global__ void __launch_bounds__(ATTB_THREADS, 1)
attn_bwd_kernel(const __grid_constant__ CUtensorMap tmQKV, const __grid_constant__ CUtensorMap tmQKVcol,
                const __grid_constant__ CUtensorMap tmDO, const __grid_constant__ CUtensorMap tmDOcol,
                const AttnBwdArgs a) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + (((raw_addr + 1023u) & ~1023u) - raw_addr);
  const AttnGeom& g = a.g;
  uint8_t* sQ = smem;
  uint8_t* sDO = smem + 16384;
  uint8_t* sK = smem + 32768;        // 2 stages x 16 KiB
  uint8_t* sV = smem + 65536;        // 2 stages x 16 KiB
  uint8_t* sP = smem + 98304;        // 32 KiB  [128 q][128 keys] as two 64-key K-major tiles
  uint8_t* sDS = smem + 131072;      // 32 KiB
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 163840);
  uint64_t* bar_q = bars + 0;
  uint64_t* kv_full = bars + 1;   // [2]
  uint64_t* kv_empty = bars + 3;  // [2]
  uint64_t* bar_s = bars + 5;
  uint64_t* bar_pds = bars + 6;
  uint64_t* bar_mma2 = bars + 7;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 8);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int blk = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
  const int q0 = blk * 128;
  const int HD = g.H * 64;

  // chunk schedule
  const int n_text_chunks = (MODE == ATT_TEXT) ? min(blk + 1, (g.T + 127) / 128) : (g.T + 127) / 128;
  const int n_local_chunks = (MODE == ATT_TEXT) ? 0 : (MODE == ATT_CONV ? 2 : 1);
  const int n_chunks = n_text_chunks + n_local_chunks;

  if (warp == 0 && lane == 0) {
    mbar_init(bar_q, 1);
    for (int s = 0; s < 2; ++s) { mbar_init(&kv_full[s], 1); mbar_init(&kv_empty[s], 1); }
    mbar_init(bar_s, 1);
    mbar_init(bar_pds, 256);
    mbar_init(bar_mma2, 1);
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  constexpr uint32_t TM_S = 0, TM_DP = 128, TM_DQ = 256, TM_DK = 320, TM_DV = 384;

  if (warp == 0) {
    // elect.sync (not `lane == 0`): single-thread region known to the compiler -> TMA / MMA operands stay in uniform
    // registers instead of a make-uniform loop around every instruction
    if (elect_one()) {
      const int row_base = b * g.seq;
      const int lines = 128 / g.F;
      // Q and dO
      mbar_expect_tx(bar_q, 32768);
      if (MODE == ATT_COL) {
        for (int l = 0; l < lines; ++l) {
          tma_load_4d(sQ + l * g.F * 128, &tmQKVcol, bar_q, h * 64, blk * lines + l, 0, b);
          tma_load_4d(sDO + l * g.F * 128, &tmDOcol, bar_q, h * 64, blk * lines + l, 0, b);
        }
      } else {
        const int t0 = (MODE == ATT_TEXT) ? q0 : g.T + q0;
        tma_load_2d(sQ, &tmQKV, bar_q, h * 64, row_base + t0);
        tma_load_2d(sDO, &tmDO, bar_q, h * 64, row_base + t0);
      }
      for (int c = 0; c < n_chunks; ++c) {
        const int s = c & 1;
        const uint32_t ph = (c >> 1) & 1;
        mbar_wait_backoff(&kv_empty[s], ph ^ 1, 64);
        mbar_expect_tx(&kv_full[s], 32768);
        uint8_t* dk = sK + s * 16384;
        uint8_t* dv = sV + s * 16384;
        const int kcol = HD + h * 64, vcol = 2 * HD + h * 64;
        if (c < n_text_chunks) {
          tma_load_2d(dk, &tmQKV, &kv_full[s], kcol, row_base + c * 128);
          tma_load_2d(dv, &tmQKV, &kv_full[s], vcol, row_base + c * 128);
        } else {
          const int lc = c - n_text_chunks;
          if (MODE == ATT_COL) {
            for (int l = 0; l < lines; ++l) {
              tma_load_4d(dk + l * g.F * 128, &tmQKVcol, &kv_full[s], kcol, blk * lines + l, 0, b);
              tma_load_4d(dv + l * g.F * 128, &tmQKVcol, &kv_full[s], vcol, blk * lines + l, 0, b);
            }
          } else {
            const int t0 = (MODE == ATT_CONV) ? g.T + q0 - 2 * g.F + lc * 128 : g.T + q0;
            tma_load_2d(dk, &tmQKV, &kv_full[s], kcol, row_base + t0);
            tma_load_2d(dv, &tmQKV, &kv_full[s], vcol, row_base + t0);
          }
        }
      }
    }
  } else if (warp == 1) {
    if (elect_one()) {
      mbar_wait(bar_q, 0);
      const uint32_t q_addr = smem_u32(sQ), do_addr = smem_u32(sDO), p_addr = smem_u32(sP), ds_addr = smem_u32(sDS);
      constexpr uint32_t id_dq = make_idesc_bf16(128, 64, 0, 1);
      constexpr uint32_t id_dkv = make_idesc_bf16(128, 64, 1, 1);
      for (int c = 0; c < n_chunks; ++c) {
        const int s = c & 1;
        const uint32_t ph = (c >> 1) & 1;
        mbar_wait(&kv_full[s], ph);
        tc_fence_after();
        const uint32_t k_addr = smem_u32(sK + s * 16384), v_addr = smem_u32(sV + s * 16384);
        // a short text tail (T mod 128 keys, padded to 16) only needs that many score columns / dQ k-steps
        const int nkeys = (c < n_text_chunks) ? min(128, g.T_pad - c * 128) : 128;
        const uint32_t id_sn = make_idesc_bf16(128, nkeys, 0, 0);
#pragma unroll
        for (int kk = 0; kk < 4; ++kk)
          umma_bf16_ss(tmem + TM_S, make_smem_desc_sw128(q_addr + kk * 32, 0, 1024),
                       make_smem_desc_sw128(k_addr + kk * 32, 0, 1024), id_sn, kk > 0);
#pragma unroll
        for (int kk = 0; kk < 4; ++kk)
          umma_bf16_ss(tmem + TM_DP, make_smem_desc_sw128(do_addr + kk * 32, 0, 1024),
                       make_smem_desc_sw128(v_addr + kk * 32, 0, 1024), id_sn, kk > 0);
        umma_commit(bar_s);
        mbar_wait_backoff(bar_pds, c & 1);
        tc_fence_after();
        for (int k16 = 0; k16 < nkeys / 16; ++k16) {
          // dQ += dS K   (A = dS K-major over keys; B = K read MN-major: N = 64 dims, K = keys)
          umma_bf16_ss(tmem + TM_DQ, make_smem_desc_sw128(ds_addr + (k16 >> 2) * 16384 + (k16 & 3) * 32, 0, 1024),
                       make_smem_desc_sw128(k_addr + k16 * 2048, 8192, 1024), id_dq, (c > 0 || k16 > 0));
        }
        {
#pragma unroll
          for (int k16 = 0; k16 < 8; ++k16) {
            // dV = P^T dO, dK = dS^T Q  (A read MN-major: M = keys, K = queries; B MN-major)
            umma_bf16_ss(tmem + TM_DV, make_smem_desc_sw128(p_addr + k16 * 2048, 16384, 1024),
                         make_smem_desc_sw128(do_addr + k16 * 2048, 8192, 1024), id_dkv, k16 > 0);
            umma_bf16_ss(tmem + TM_DK, make_smem_desc_sw128(ds_addr + k16 * 2048, 16384, 1024),
                         make_smem_desc_sw128(q_addr + k16 * 2048, 8192, 1024), id_dkv, k16 > 0);
          }
        }
        umma_commit(&kv_empty[s]);
        umma_commit(bar_mma2);
      }
    }
  } else {
    // ---- 8 worker warps: warps w and w+4 share TMEM lane quarter (w & 3) and split the column work ----
    const int qw = warp & 3;
    const int part = (warp - 2) >> 2;   // 0: column groups 0,1 / dK / dQ dims 0-31   1: groups 2,3 / dV / dims 32-63
    const int i = qw * 32 + lane;
    const uint32_t trow = tmem + (static_cast<uint32_t>(qw * 32) << 16);
    int tpos, ip = 0, ir = 0, ic = 0;
    bool q_valid;
    if (MODE == ATT_TEXT) {
      tpos = q0 + i;
      q_valid = tpos < g.T;
    } else {
      ip = img_query_pos<MODE>(g, blk, i);
      ir = ip / g.F;
      ic = ip - ir * g.F;
      tpos = g.T + ip;
      q_valid = tpos < g.seq;
    }
    const long long bh = static_cast<long long>(b) * g.H + h;
    const float lse_i = q_valid ? a.lse[bh * g.seq + tpos] : 0.f;
    // delta_i = sum_d dO[i, d] * O[i, d]: O row straight from global (one 128-byte line per thread), dO row from the
    // swizzled smem tile the MMAs use
    float delta_i = 0.f;
    {
      uint4 ov[8];
      const uint4* op = reinterpret_cast<const uint4*>(a.out + (static_cast<long long>(b) * g.seq + (q_valid ? tpos : 0)) * HD + h * 64);
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
    }
    const int win0 = qw * 32;
    // validity bits of the local keys this row can see, relative to win0 (same construction as the forward)
    uint32_t lmask[3] = {0u, 0u, 0u};
    if (MODE == ATT_ROW || MODE == ATT_COL) {
      const int lo = max(i - (MODE == ATT_ROW ? ic : ir), win0) - win0;
      lmask[0] = (0xFFFFFFFFu >> (31 - lane)) & (0xFFFFFFFFu << lo);
    }
    if (MODE == ATT_CONV) {
      for (int dr = 0; dr <= g.conv_half; ++dr)
        for (int dc = -g.conv_half; dc <= g.conv_half; ++dc) {
          const int kr = ir - dr, kc = ic + dc;
          const int kp = kr * g.F + kc;
          const int j = kp - (q0 - 2 * g.F) - win0;
          if (kr >= 0 && kc >= 0 && kc < g.F && kp <= ip && j >= 0 && j < 96) lmask[j >> 5] |= 1u << (j & 31);
        }
    }
    const int text_limit = (MODE == ATT_TEXT) ? min(tpos + 1, g.T) : g.T;
    uint8_t* pbase = sP + part * 16384 + i * 128;   // this warp's column groups live in 64-key tile `part`
    uint8_t* dsbase = sDS + part * 16384 + i * 128;
    const int sw = i & 7;

    for (int c = 0; c < n_chunks; ++c) {
      const bool is_text = c < n_text_chunks;
      const int lc = c - n_text_chunks;
      const int nkeys = is_text ? min(128, g.T_pad - c * 128) : 128;   // keys the chunk's MMAs cover
      mbar_wait(bar_s, c & 1);
      tc_fence_after();
      // ---- P and dS for this chunk ----
#pragma unroll
      for (int t = 0; t < 2; ++t) {
        const int j0 = (part * 2 + t) * 32;
        if (j0 >= nkeys) continue;   // beyond a short text tail: columns are not read by the dQ MMA
        uint32_t vm = 0u;
        if (q_valid) {
          if (is_text) {
            const int nv = text_limit - (c * 128 + j0);
            vm = nv <= 0 ? 0u : (nv >= 32 ? 0xFFFFFFFFu : (1u << nv) - 1u);
          } else if (MODE == ATT_ROW || MODE == ATT_COL) {
            vm = (j0 == win0) ? lmask[0] : 0u;
          } else if (MODE == ATT_CONV) {
            const int jrel = lc * 128 + j0 - win0;
            vm = (jrel >= 0 && jrel < 96) ? lmask[jrel >> 5] : 0u;
          }
        }
        float p[32], ds[32];
        if (__any_sync(0xffffffffu, vm != 0u)) {
          uint32_t rs[32], rp[32];
          tmem_ld32(trow + TM_S + j0, rs);
          tmem_ld32(trow + TM_DP + j0, rp);
          tmem_ld_wait();
          if (__all_sync(0xffffffffu, vm == 0xFFFFFFFFu)) {
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              p[j] = fast_exp2(__uint_as_float(rs[j]) * kLog2e - lse_i);
              ds[j] = p[j] * (__uint_as_float(rp[j]) - delta_i);
            }
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              const float pe = ((vm >> j) & 1u) ? fast_exp2(__uint_as_float(rs[j]) * kLog2e - lse_i) : 0.f;
              p[j] = pe;
              ds[j] = pe * (__uint_as_float(rp[j]) - delta_i);
            }
          }
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j) { p[j] = 0.f; ds[j] = 0.f; }
        }
#pragma unroll
        for (int gq = 0; gq < 4; ++gq) {
          const uint32_t off = static_cast<uint32_t>(((t * 4 + gq) ^ sw) << 4);
          uint4 u;
          u.x = pack_bf16x2(p[gq * 8 + 0], p[gq * 8 + 1]);
          u.y = pack_bf16x2(p[gq * 8 + 2], p[gq * 8 + 3]);
          u.z = pack_bf16x2(p[gq * 8 + 4], p[gq * 8 + 5]);
          u.w = pack_bf16x2(p[gq * 8 + 6], p[gq * 8 + 7]);
          *reinterpret_cast<uint4*>(pbase + off) = u;
          u.x = pack_bf16x2(ds[gq * 8 + 0], ds[gq * 8 + 1]);
          u.y = pack_bf16x2(ds[gq * 8 + 2], ds[gq * 8 + 3]);
          u.z = pack_bf16x2(ds[gq * 8 + 4], ds[gq * 8 + 5]);
          u.w = pack_bf16x2(ds[gq * 8 + 6], ds[gq * 8 + 7]);
          *reinterpret_cast<uint4*>(dsbase + off) = u;
        }
      }
      fence_proxy_async_smem();
      tc_fence_before();
      mbar_arrive(bar_pds);

      // ---- dK (part 0) / dV (part 1) of this chunk: thread = key row i of the chunk ----
      mbar_wait(bar_mma2, c & 1);
      tc_fence_after();
      const bool direct = !is_text && (MODE == ATT_ROW || MODE == ATT_COL);
      // token index of key row `ri` of the chunk, -1 if it must not be written
      auto key_tok = [&](int ri) -> int {
        if (is_text) {
          const int kt = c * 128 + ri;
          return kt < g.T ? kt : -1;
        }
        const int kp = q0 - 2 * g.F + lc * 128 + ri;   // (conv_like local chunks; axial local chunks are `direct`)
        return (kp >= 0 && kp < g.P - 1 && kp < q0 + 128) ? g.T + kp : -1;
      };
      const int ktok = direct ? (q_valid ? tpos : -1) : key_tok(i);   // axial: the block's own tokens
      if (__any_sync(0xffffffffu, ktok >= 0)) {
        const bool is_v = part == 1;
#pragma unroll 1
        for (int half = 0; half < 2; ++half) {
          const int d0 = half * 32;
          uint32_t r[32];
          tmem_ld32(trow + (is_v ? TM_DV : TM_DK) + d0, r);
          tmem_ld_wait();
          if (direct) {
            if (ktok < 0) continue;
            float v[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
            const long long grow = static_cast<long long>(b) * g.seq + ktok;
            inv_rotary32(v, a.rot_t + ktok, a.rot_len, d0, 1.0f);
            store32_bf16(a.dqkv + grow * (3 * HD) + (is_v ? 2 : 1) * HD + h * 64 + d0, v);
          } else {
            // shared keys: fp32 red.add into kv_acc.  Straight from the TMEM layout (lane = key row) one red.v4
            // touches 32 different rows = 32 L1 wavefronts; the warp's own slice of the P / dS buffers (dead
            // until the next chunk) transposes the 32 x 32 tile so that an instruction covers four whole
            // 128-byte row segments instead.
            uint8_t* stg = (half == 0 ? sP : sDS) + part * 16384 + qw * 4096;
#pragma unroll
            for (int c8 = 0; c8 < 8; ++c8)
              *reinterpret_cast<uint4*>(stg + lane * 128 + ((c8 ^ (lane & 7)) << 4)) =
                  make_uint4(r[4 * c8], r[4 * c8 + 1], r[4 * c8 + 2], r[4 * c8 + 3]);
            __syncwarp();
            const int cq = lane & 7;
#pragma unroll
            for (int it = 0; it < 8; ++it) {
              const int rr = (lane >> 3) + 4 * it;
              const int kt = key_tok(qw * 32 + rr);
              const uint4 u = *reinterpret_cast<const uint4*>(stg + rr * 128 + ((cq ^ (rr & 7)) << 4));
              if (kt >= 0) {
                float* dst = a.kv_acc + (static_cast<long long>(b) * g.seq + kt) * (2 * HD) + (is_v ? HD : 0) +
                             h * 64 + d0 + cq * 4;
                red_add_v4(dst, __uint_as_float(u.x), __uint_as_float(u.y), __uint_as_float(u.z),
                           __uint_as_float(u.w));
              }
            }
            __syncwarp();
          }
        }
      }
      tc_fence_before();
    }

    // ---- dQ (each part one half of the head dims) ----
    {
      uint32_t r[32];
      tmem_ld32(trow + TM_DQ + part * 32, r);
      tmem_ld_wait();
      if (q_valid) {
        float v[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
        inv_rotary32(v, a.rot_t + tpos, a.rot_len, part * 32, a.q_scale);
        store32_bf16(a.dqkv + (static_cast<long long>(b) * g.seq + tpos) * (3 * HD) + h * 64 + part * 32, v);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem, 512);
  }
}

// kv_acc (fp32, rot domain) -> inverse rotary -> bf16 into d_qkv k/v sections; clears kv_acc.
// rows_per_sample = T (axial: only text keys went through the accumulator) or seq (conv_like).
// thread = 8 consecutive columns (4 rotary pairs) of one row.
constexpr int FIN_ROWS = 4;   // rows per thread: all loads of the four rows are issued before the first use
__global__ void __launch_bounds__(256)
attn_bwd_finish_kernel(float* __restrict__ kv_acc, __nv_bfloat16* __restrict__ dqkv, const float2* __restrict__ rot,
                       int B, int seq, int rows_per_sample, int HD) {
  const int groups_per_row = (2 * HD) / 8;
  const unsigned n_rows = static_cast<unsigned>(B) * rows_per_sample;
  const unsigned row_groups = (n_rows + FIN_ROWS - 1) / FIN_ROWS;
  const unsigned idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= row_groups * groups_per_row) return;
  const unsigned gcol = idx % groups_per_row, rg = idx / groups_per_row;
  const int col = gcol * 8;
  float4 a[FIN_ROWS], c[FIN_ROWS], r01[FIN_ROWS], r23[FIN_ROWS];
  long long grow[FIN_ROWS];
  bool ok[FIN_ROWS];
#pragma unroll
  for (int j = 0; j < FIN_ROWS; ++j) {
    const unsigned rr = rg * FIN_ROWS + j;
    ok[j] = rr < n_rows;
    const unsigned rs = ok[j] ? rr : 0u;
    const int t = rs % rows_per_sample, b = rs / rows_per_sample;
    grow[j] = static_cast<long long>(b) * seq + t;
    const float4* src = reinterpret_cast<const float4*>(kv_acc + grow[j] * (2 * HD) + col);
    a[j] = src[0];
    c[j] = src[1];
    const float4* rt = reinterpret_cast<const float4*>(rot + static_cast<long long>(t) * 32 + ((col & 63) >> 1));
    r01[j] = __ldg(rt);        // (cos0, sin0, cos1, sin1)
    r23[j] = __ldg(rt + 1);    // (cos2, sin2, cos3, sin3)
  }
#pragma unroll
  for (int j = 0; j < FIN_ROWS; ++j) {
    if (!ok[j]) continue;
    float4* src = reinterpret_cast<float4*>(kv_acc + grow[j] * (2 * HD) + col);
    src[0] = make_float4(0.f, 0.f, 0.f, 0.f);
    src[1] = make_float4(0.f, 0.f, 0.f, 0.f);
    uint4 o;
    o.x = pack_bf16x2(a[j].x * r01[j].x + a[j].y * r01[j].y, a[j].y * r01[j].x - a[j].x * r01[j].y);
    o.y = pack_bf16x2(a[j].z * r01[j].z + a[j].w * r01[j].w, a[j].w * r01[j].z - a[j].z * r01[j].w);
    o.z = pack_bf16x2(c[j].x * r23[j].x + c[j].y * r23[j].y, c[j].y * r23[j].x - c[j].x * r23[j].y);
    o.w = pack_bf16x2(c[j].z * r23[j].z + c[j].w * r23[j].w, c[j].w * r23[j].z - c[j].z * r23[j].w);
    *reinterpret_cast<uint4*>(dqkv + grow[j] * (3 * HD) + HD + col) = o;
  }
}

template <int MODE>
static int launch_bwd(const CUtensorMap& t1, const CUtensorMap& t2, const CUtensorMap& t3, const CUtensorMap& t4,
                      const AttnBwdArgs& a, int nblocks, cudaStream_t st) {
  auto kern = attn_bwd_kernel<MODE>;
  const int smem = 163840 + 128 + 1024;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  if (e != cudaSuccess) return -21;
  dim3 grid(nblocks, a.g.H, a.g.B);
  kern<<<grid, ATTB_THREADS, smem, st>>>(t1, t2, t3, t4, a);
  return cudaGetLastError() == cudaSuccess ? 0 : -22;
}

}  // namespace b200

using namespace b200;

extern "C" int b200_attn_bwd(const void* qkv, const void* out, const void* d_out, const float* lse, float* delta,
                             void* dqkv, float* kv_acc, const void* rot, const void* rot_t, float q_scale, int B, int H, int seq, int T,
                             int F, int mode, int conv_kernel, void* stream) {
  AttnGeom g;
  g.B = B; g.H = H; g.seq = seq; g.T = T; g.T_pad = (T + 15) / 16 * 16; g.F = F; g.P = F * F;
  g.n_text_blocks = (T + 127) / 128; g.n_img_blocks = g.P / 128; g.conv_half = conv_kernel / 2;
  if (g.P % 128 || 32 % F || seq != T + g.P - 1) return -30;   // grid sides 16 and 32, as in the forward
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  // text + axial patterns: the fourth-generation kernels (attn_bwd4_sm100.cu); DALLE_B200_ATTN_BWD=2 keeps this file's
  // kernel for A/B runs
  static const int which = [] { const char* e = getenv("DALLE_B200_ATTN_BWD"); return e ? atoi(e) : 4; }();
  if (which == 4) {
    const int rc4 = attn4_bwd(qkv, out, d_out, lse, delta, dqkv, kv_acc, rot, rot_t, q_scale, B, H, seq, T, F, mode, st);
    if (rc4 <= 0) return rc4;
  }
  const int HD = H * 64;
  AttnBwdArgs a;
  a.g = g; a.lse = lse; a.out = reinterpret_cast<const __nv_bfloat16*>(out); (void)delta; a.dqkv = reinterpret_cast<__nv_bfloat16*>(dqkv); a.kv_acc = kv_acc;
  a.rot = reinterpret_cast<const float2*>(rot); a.q_scale = q_scale;
  a.rot_t = reinterpret_cast<const float2*>(rot_t); a.rot_len = seq + 1;

  CUtensorMap tq, ttail, tqcol, tdo, tdocol;
  int rc = attn_make_maps(qkv, g, ATT_COL, &tq, &ttail, &tqcol);
  if (rc) return rc;
  const uint64_t rows = (uint64_t)B * seq;
  rc = make_tmap_bf16_2d(&tdo, d_out, rows, HD, HD, 64, 128);
  if (rc) return rc;
  {
    const __nv_bfloat16* base = reinterpret_cast<const __nv_bfloat16*>(d_out) + (size_t)T * HD;
    uint64_t dims[4] = {(uint64_t)HD, (uint64_t)F, (uint64_t)F, (uint64_t)B};
    uint64_t strides[3] = {(uint64_t)HD * 2, (uint64_t)HD * 2 * F, (uint64_t)HD * 2 * seq};
    uint32_t box[4] = {64, 1, (uint32_t)F, 1};
    rc = make_tmap_bf16_nd(&tdocol, base, 4, dims, strides, box);
    if (rc) return rc;
  }
  rc = launch_bwd<ATT_TEXT>(tq, tqcol, tdo, tdocol, a, g.n_text_blocks, st);
  if (rc) return rc;
  switch (mode) {
    case ATT_ROW: rc = launch_bwd<ATT_ROW>(tq, tqcol, tdo, tdocol, a, g.n_img_blocks, st); break;
    case ATT_COL: rc = launch_bwd<ATT_COL>(tq, tqcol, tdo, tdocol, a, g.n_img_blocks, st); break;
    case ATT_CONV: rc = launch_bwd<ATT_CONV>(tq, tqcol, tdo, tdocol, a, g.n_img_blocks, st); break;
    default: return -31;
  }
  if (rc) return rc;
  const int rows_per_sample = (mode == ATT_CONV) ? seq : T;
  const long long total =
      (static_cast<long long>(B) * rows_per_sample + FIN_ROWS - 1) / FIN_ROWS * (2 * HD / 8);
  if (total >= (1ll << 32)) return -24;
  const int threads = 256;
  const long long blocks = (total + threads - 1) / threads;
  attn_bwd_finish_kernel<<<static_cast<unsigned>(blocks), threads, 0, st>>>(
      kv_acc, reinterpret_cast<__nv_bfloat16*>(dqkv), a.rot, B, seq, rows_per_sample, HD);
  return cudaGetLastError() == cudaSuccess ? 0 : -23;
}
