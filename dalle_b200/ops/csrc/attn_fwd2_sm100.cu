// Sparse DALL-E attention forward, second generation: flash-style walk over key chunks with an
// online softmax so that a CTA needs only 256 TMEM columns and ~100 KB of shared memory -> TWO CTAs per SM, one
// CTA's softmax overlapping the other's TMA loads and MMAs (the first-generation kernel in attn_fwd_sm100.cu held
// the whole score row in TMEM, one CTA per SM, and spent a third of its samples spinning on load barriers:
// profiles/ncu_summary_r1_hot_kernels_B16.txt).
//
// One CTA = one (batch, head, 128-query block).  Key chunks:
//   text chunks of 128 keys, then ONE merged chunk [text tail (T mod 128, padded to 16) | 128 local keys],
//   conv_like adds a last chunk with the remaining 2F halo keys.
// Per chunk: S = Q K_c^T (tcgen05, TMEM cols [0,144)) -> softmax warps: running max / sum, P (bf16) -> swizzled smem
// -> O_c = P V_c (TMEM cols [160,224), fresh accumulator) -> merged into the per-thread fp32 output registers.
#include "attn_common.cuh"
#include "b200_api.h"

#include <cstdio>
#include <cstdlib>

namespace b200 {

int make_tmap_bf16_2d(CUtensorMap* m, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_cols,
                      uint32_t box_rows);
int make_tmap_bf16_nd(CUtensorMap* m, const void* ptr, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                      const uint32_t* box);
int attn_make_maps(const void* qkv, const AttnGeom& g, int mode, CUtensorMap* t128, CUtensorMap* ttail,
                   CUtensorMap* taux);
int attn3_fwd(const void* qkv, void* out, float* lse, int B, int H, int seq, int T, int F, int mode, cudaStream_t st);
int attn4_fwd(const void* qkv, void* out, float* lse, int B, int H, int seq, int T, int F, int mode, cudaStream_t st);

constexpr int AF2_THREADS = 128;
constexpr int AF2_MAXK = 144;                       // keys per chunk
constexpr int AF2_KV_BYTES = AF2_MAXK * 128;        // 18 KiB
constexpr int AF2_P_BYTES = 3 * 16384;              // three 64-key K-major tiles
constexpr int AF2_SMEM = 16384 + 2 * AF2_KV_BYTES + AF2_P_BYTES + 128 + 1024;
constexpr uint32_t AF2_TM_S = 0, AF2_TM_O = 160;

struct AttnFwd2Args {
  AttnGeom g;
  __nv_bfloat16* out;
  float* lse;
  int ld_out;
};

struct ChunkDesc {
  int n_text, key0;    // text segment: keys [key0, key0 + n_text) (n_text multiple of 16; validity checked per key)
  int n_local, j0;     // local segment: local key indices [j0, j0 + n_local)
};

template <int MODE>
__device__ __forceinline__ int af2_num_chunks(const AttnGeom& g, int blk) {
  const int nT = (g.T + 127) / 128;
  if (MODE == ATT_TEXT) return min(blk + 1, nT);
  const int tail = g.T_pad - 128 * (nT - 1);
  const bool merge = tail != 128 && tail + 128 <= AF2_MAXK;  // short tail rides with the local keys
  return (merge ? nT - 1 : nT) + 1 + (MODE == ATT_CONV ? 1 : 0);
}

template <int MODE>
__device__ __forceinline__ ChunkDesc af2_chunk(const AttnGeom& g, int c) {
  const int nT = (g.T + 127) / 128;
  const int tail = g.T_pad - 128 * (nT - 1);
  ChunkDesc d;
  d.n_text = 0; d.key0 = 0; d.n_local = 0; d.j0 = 0;
  if (MODE == ATT_TEXT) {
    d.key0 = 128 * c;
    d.n_text = (c == nT - 1) ? tail : 128;
    return d;
  }
  const bool merge = tail != 128 && tail + 128 <= AF2_MAXK;
  const int n_text_chunks = merge ? nT - 1 : nT;
  if (c < n_text_chunks) { d.key0 = 128 * c; d.n_text = (c == nT - 1) ? tail : 128; return d; }
  if (c == n_text_chunks) {
    if (merge) { d.key0 = 128 * (nT - 1); d.n_text = tail; }
    d.n_local = 128; d.j0 = 0;
    return d;
  }
  d.n_local = 2 * g.F; d.j0 = 128;   // conv_like halo remainder
  return d;
}

template <int MODE>
__global__ void __launch_bounds__(AF2_THREADS, 2)
attn_fwd2_kernel(const __grid_constant__ CUtensorMap tm128, const __grid_constant__ CUtensorMap tmTail,
                 const __grid_constant__ CUtensorMap tmAux, const AttnFwd2Args a) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + (((raw_addr + 1023u) & ~1023u) - raw_addr);
  const AttnGeom& g = a.g;
  uint8_t* sQ = smem;
  uint8_t* sK = smem + 16384;
  uint8_t* sV = sK + AF2_KV_BYTES;
  uint8_t* sP = sV + AF2_KV_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sP + AF2_P_BYTES);
  uint64_t* bar_q = bars + 0;
  uint64_t* bar_k = bars + 1;
  uint64_t* bar_v = bars + 2;
  uint64_t* bar_s = bars + 3;
  uint64_t* bar_p = bars + 4;
  uint64_t* bar_o = bars + 5;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 8);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int blk = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
  const int q0 = blk * 128;
  const int n_chunks = af2_num_chunks<MODE>(g, blk);

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tm128);
    tma_prefetch_desc(&tmTail);
    tma_prefetch_desc(&tmAux);
    mbar_init(bar_q, 1); mbar_init(bar_k, 1); mbar_init(bar_v, 1); mbar_init(bar_s, 1);
    mbar_init(bar_p, 128); mbar_init(bar_o, 1);
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

  // Thread 0 is also the TMA producer and the MMA issuer (single-thread instructions); a 128-thread CTA keeps the
  // register allocation at 4 warps so that two CTAs share an SM.
  const int row_base = b * g.seq;
  const int lines = 128 / g.F;
  auto load_kv = [&](int c, int sect) {
    const ChunkDesc d = af2_chunk<MODE>(g, c);
    uint64_t* full = sect == 0 ? bar_k : bar_v;
    uint8_t* dst = sect == 0 ? sK : sV;
    const int col = (sect == 0 ? g.H * 64 : 2 * g.H * 64) + h * 64;
    mbar_expect_tx(full, static_cast<uint32_t>(d.n_text + d.n_local) * 128u);
    if (d.n_text == 128) tma_load_2d(dst, &tm128, full, col, row_base + d.key0);
    else if (d.n_text > 0) tma_load_2d(dst, &tmTail, full, col, row_base + d.key0);
    uint8_t* dl = dst + d.n_text * 128;
    if (d.n_local > 0) {
      if (MODE == ATT_ROW) {
        tma_load_2d(dl, &tm128, full, col, row_base + g.T + q0);
      } else if (MODE == ATT_COL) {
        for (int l = 0; l < lines; ++l) tma_load_4d(dl + l * g.F * 128, &tmAux, full, col, blk * lines + l, 0, b);
      } else if (MODE == ATT_CONV) {
        if (d.j0 == 0) tma_load_2d(dl, &tm128, full, col, row_base + g.T + q0 - 2 * g.F);
        else tma_load_2d(dl, &tmAux, full, col, row_base + g.T + q0 - 2 * g.F + 128);  // box rows = 2F
      }
    }
  };
  // single-thread TMA / MMA issue goes through elect.sync on warp 0 (the compiler then keeps the operands in uniform
  // registers; `threadIdx.x == 0` put a make-uniform loop around every instruction)
  if (warp == 0 && elect_one()) {
    mbar_expect_tx(bar_q, 16384);
    if (MODE == ATT_COL) {
      for (int l = 0; l < lines; ++l) tma_load_4d(sQ + l * g.F * 128, &tmAux, bar_q, h * 64, blk * lines + l, 0, b);
    } else {
      tma_load_2d(sQ, &tm128, bar_q, h * 64, row_base + ((MODE == ATT_TEXT) ? q0 : g.T + q0));
    }
    load_kv(0, 0);
    load_kv(0, 1);
  }
  const uint32_t q_addr = smem_u32(sQ), k_addr = smem_u32(sK), v_addr = smem_u32(sV), p_addr = smem_u32(sP);
  auto issue_s = [&](int c) {       // thread 0
    const ChunkDesc d = af2_chunk<MODE>(g, c);
    if (c == 0) mbar_wait(bar_q, 0);
    mbar_wait(bar_k, c & 1);
    tc_fence_after();
    const uint32_t idesc_s = make_idesc_bf16(128, d.n_text + d.n_local, 0, 0);
#pragma unroll
    for (int kk = 0; kk < 4; ++kk)
      umma_bf16_ss(tmem + AF2_TM_S, make_smem_desc_sw128(q_addr + kk * 32, 0, 1024),
                   make_smem_desc_sw128(k_addr + kk * 32, 0, 1024), idesc_s, kk > 0 ? 1u : 0u);
    umma_commit(bar_s);
  };
  auto issue_pv = [&](int c) {      // thread 0, after every softmax thread arrived on bar_p
    const ChunkDesc d = af2_chunk<MODE>(g, c);
    mbar_wait(bar_p, c & 1);
    mbar_wait(bar_v, c & 1);
    tc_fence_after();
    const uint32_t idesc_pv = make_idesc_bf16(128, 64, 0, 1);
    for (int k16 = 0; k16 < (d.n_text + d.n_local) / 16; ++k16)
      umma_bf16_ss(tmem + AF2_TM_O, make_smem_desc_sw128(p_addr + (k16 >> 2) * 16384 + (k16 & 3) * 32, 0, 1024),
                   make_smem_desc_sw128(v_addr + k16 * 2048, 8192, 1024), idesc_pv, k16 > 0 ? 1u : 0u);
    umma_commit(bar_o);
  };
  if (warp == 0 && elect_one()) issue_s(0);
  __syncwarp();

#include "attn_fwd2_softmax.inc"

  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem, 256);
  }
}

template <int MODE>
static int launch_fwd2(const CUtensorMap& t128, const CUtensorMap& ttail, const CUtensorMap& taux,
                       const AttnFwd2Args& a, int nblocks, cudaStream_t st) {
  auto kern = attn_fwd2_kernel<MODE>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, AF2_SMEM);
  if (e != cudaSuccess) return -21;
  cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
  dim3 grid(nblocks, a.g.H, a.g.B);
  kern<<<grid, AF2_THREADS, AF2_SMEM, st>>>(t128, ttail, taux, a);
  return cudaGetLastError() == cudaSuccess ? 0 : -22;
}

int attn_make_maps(const void* qkv, const AttnGeom& g, int mode, CUtensorMap* t128, CUtensorMap* ttail,
                   CUtensorMap* taux) {
  const uint64_t rows = (uint64_t)g.B * g.seq, cols = 3ull * g.H * 64;
  int rc = make_tmap_bf16_2d(t128, qkv, rows, cols, cols, 64, 128);
  if (rc) return rc;
  const int tail = g.T_pad % 128;
  rc = make_tmap_bf16_2d(ttail, qkv, rows, cols, cols, 64, tail ? tail : 128);
  if (rc) return rc;
  if (mode == ATT_COL) {
    // image part viewed as [b][r][c][feature]; one box = one image column (F tokens, stride F rows)
    const __nv_bfloat16* base = reinterpret_cast<const __nv_bfloat16*>(qkv) + (size_t)g.T * cols;
    // the last image position has no token: rows r = F-1 expose F-1 columns only for the last sample -> the
    // tensor map is declared over the full grid; reads past the allocation are avoided by bounding dim 3 (batch)
    // and letting the final element fall into the next sample / OOB fill.
    uint64_t dims[4] = {cols, (uint64_t)g.F, (uint64_t)g.F, (uint64_t)g.B};
    uint64_t strides[3] = {cols * 2, cols * 2 * g.F, cols * 2 * g.seq};
    uint32_t box[4] = {64, 1, (uint32_t)g.F, 1};
    rc = make_tmap_bf16_nd(taux, base, 4, dims, strides, box);
  } else {
    rc = make_tmap_bf16_2d(taux, qkv, rows, cols, cols, 64, 2 * g.F);
  }
  return rc;
}

}  // namespace b200

using namespace b200;

extern "C" int b200_attn_fwd(const void* qkv, void* out, float* lse, int B, int H, int seq, int T, int F, int mode,
                             int conv_kernel, void* stream) {
  AttnGeom g;
  g.B = B; g.H = H; g.seq = seq; g.T = T; g.T_pad = (T + 15) / 16 * 16; g.F = F; g.P = F * F;
  g.n_text_blocks = (T + 127) / 128; g.n_img_blocks = g.P / 128; g.conv_half = conv_kernel / 2;
  if (g.P % 128 || 128 % F || seq != T + g.P - 1) return -30;
  // the softmax warps assume that a query's local keys sit inside its own 32-key window: grid sides 16 and 32
  // (the reference fixes 32, task.py:26).  Wider grids are refused here and in the backward, not approximated.
  if (32 % F) {
    fprintf(stderr, "b200_attn_fwd: image grid side %d is not supported by the sm_100a kernels (16 or 32)\n", F);
    return -30;
  }
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  // text + axial patterns have three forward kernels; DALLE_B200_ATTN_FWD = 2 | 3 | 4 picks one for A/B runs:
  //   4  (default) many light CTAs, 64-key chunks, P and O in TMEM (attn_fwd4_sm100.cu)
  //   3  persistent software-pipelined kernel (attn3_fwd_sm100.cu; slower, see profiles/attn3_fwd_timeline_r2.txt)
  //   2  this file
  static const int which = [] { const char* e = getenv("DALLE_B200_ATTN_FWD"); return e ? atoi(e) : 4; }();
  if (which == 3 || which == 4) {
    const int rc3 = which == 3 ? attn3_fwd(qkv, out, lse, B, H, seq, T, F, mode, st)
                               : attn4_fwd(qkv, out, lse, B, H, seq, T, F, mode, st);
    if (rc3 <= 0) return rc3;
  }
  AttnFwd2Args a;
  a.g = g; a.out = reinterpret_cast<__nv_bfloat16*>(out); a.lse = lse; a.ld_out = H * 64;
  CUtensorMap t128, ttail, taux;
  int rc = attn_make_maps(qkv, g, ATT_TEXT, &t128, &ttail, &taux);
  if (rc) return rc;
  rc = launch_fwd2<ATT_TEXT>(t128, ttail, taux, a, g.n_text_blocks, st);
  if (rc) return rc;
  rc = attn_make_maps(qkv, g, mode, &t128, &ttail, &taux);
  if (rc) return rc;
  switch (mode) {
    case ATT_ROW: return launch_fwd2<ATT_ROW>(t128, ttail, taux, a, g.n_img_blocks, st);
    case ATT_COL: return launch_fwd2<ATT_COL>(t128, ttail, taux, a, g.n_img_blocks, st);
    case ATT_CONV: return launch_fwd2<ATT_CONV>(t128, ttail, taux, a, g.n_img_blocks, st);
    default: return -31;
  }
}
