// Memory-bound kernels around the GEMMs: tied-embedding gather / scatter-add, LayerNorm fused with the DALL-E
// token shift (forward + backward), LayerScale/residual backward prep, column sums, masked-segment cross-entropy.
// Replaces ATen embedding / layer_norm / pad+chunk+cat / mul / cross_entropy calls of the reference model
// (SURVEY K1, K2, K3, K10, K13).  One warp per row, 128-bit accesses, fp32 math.
#include "sm100_ptx.cuh"
#include "b200_api.h"

namespace b200 {

int num_sms();  // gemm_sm100.cu

constexpr int MAXV = 8;  // float4 chunks per lane: D <= 1024

struct ShiftGeom {
  int seq, T, F, shift;
  int split, B;  // split > 0: the bf16 side is stored [all samples' first `split` tokens | all remaining tokens]
};
__device__ __forceinline__ long long permuted_row(const ShiftGeom& s, int bidx, int t) {
  if (s.split <= 0) return static_cast<long long>(bidx) * s.seq + t;
  return t < s.split ? static_cast<long long>(bidx) * s.split + t
                     : static_cast<long long>(s.B) * s.split + static_cast<long long>(bidx) * (s.seq - s.split) +
                           (t - s.split);
}

// source row (token index inside the sample) that feeds channel `col` of token t after the shift, or -1 for zero
__device__ __forceinline__ int shift_src(const ShiftGeom& s, int t, int col, int D) {
  if (!s.shift) return t;
  if (t < s.T) return (col < D / 2) ? (t >= 1 ? t - 1 : -1) : t;
  const int p = t - s.T, r = p / s.F, c = p - r * s.F;
  if (col < D / 4) return r >= 1 ? t - s.F : -1;
  if (col < D / 2) return c >= 1 ? t - 1 : -1;
  return t;
}
// destination row whose channel `col` is fed by token t (the transpose of shift_src), or -1 if none
__device__ __forceinline__ int shift_dst(const ShiftGeom& s, int t, int col, int D) {
  if (!s.shift) return t;
  if (t < s.T) return (col < D / 2) ? (t + 1 < s.T ? t + 1 : -1) : t;
  const int p = t - s.T, r = p / s.F, c = p - r * s.F;
  if (col < D / 4) return (r + 1 < s.F && t + s.F < s.seq) ? t + s.F : -1;
  if (col < D / 2) return (c + 1 < s.F && t + 1 < s.seq) ? t + 1 : -1;
  return t;
}

// Per-row routing of the shift, computed once per row (the helpers above divide by the grid side on every call):
// channels [0, D/4) of token t feed token dst[0], channels [D/4, D/2) feed dst[1], the rest stay (-1: dropped);
// zero[q] tells that nothing shifts INTO quarter q of token t.
struct ShiftRoute {
  int dst[2];
  bool zero[2];
};
// walks consecutive rows of the [B*seq] token matrix keeping (sample, token, grid row, grid column) without divisions
struct RowWalk {
  int row, bidx, t, ir, ic;
  __device__ __forceinline__ void init(const ShiftGeom& s, int row0) {
    row = row0;
    bidx = row0 / s.seq;
    t = row0 - bidx * s.seq;
    const int p = t >= s.T ? t - s.T : 0;
    ir = p / s.F;
    ic = p - ir * s.F;
  }
  __device__ __forceinline__ void next(const ShiftGeom& s) {
    ++row;
    ++t;
    if (t == s.seq) {
      t = 0; ++bidx; ir = 0; ic = 0;
    } else if (t > s.T) {
      if (++ic == s.F) { ic = 0; ++ir; }
    }
  }
};
__device__ __forceinline__ ShiftRoute shift_route(const ShiftGeom& s, const RowWalk& rw) {
  ShiftRoute r;
  const int t = rw.t;
  if (!s.shift) {
    r.dst[0] = r.dst[1] = t;
    r.zero[0] = r.zero[1] = false;
  } else if (t < s.T) {
    r.dst[0] = r.dst[1] = (t + 1 < s.T) ? t + 1 : -1;
    r.zero[0] = r.zero[1] = (t == 0);
  } else {
    r.dst[0] = (rw.ir + 1 < s.F && t + s.F < s.seq) ? t + s.F : -1;
    r.dst[1] = (rw.ic + 1 < s.F && t + 1 < s.seq) ? t + 1 : -1;
    r.zero[0] = rw.ir == 0;
    r.zero[1] = rw.ic == 0;
  }
  return r;
}
// row of the bf16 side (see ShiftGeom::split) as a 32-bit index
__device__ __forceinline__ int permuted_row32(const ShiftGeom& s, int bidx, int t) {
  if (s.split <= 0) return bidx * s.seq + t;
  return t < s.split ? bidx * s.split + t : s.B * s.split + bidx * (s.seq - s.split) + (t - s.split);
}

// ------------------------------------------------------------------------------------------------
// LayerNorm fused with the token shift, forward.  128 threads per SOURCE row: the row is read once, its statistics
// are reduced in the group, and the normalised values are scattered to the rows the shift sends them to (each
// output element has exactly one writer: its source row, or - where the shift pads with zeros - the row itself).
// ------------------------------------------------------------------------------------------------
constexpr int LNF_GROUPS = 2;
__global__ void __launch_bounds__(128 * LNF_GROUPS, 3)
ln_shift_fwd_kernel(const float* __restrict__ x, const float* __restrict__ x2, const float* __restrict__ w,
                    const float* __restrict__ b, __nv_bfloat16* __restrict__ out, float* __restrict__ mean,
                    float* __restrict__ rstd, int M, int D, float eps, ShiftGeom sg, int ring_rows) {
  // persistent blocks of 2 row groups x 128 threads (<= 8 channels per thread), each group walking a contiguous
  // range of rows with its loads three rows ahead.  (A warp-per-row variant with 8 float4 per lane measured 1.8x slower: too few loads in flight.)
  __shared__ float sred[LNF_GROUPS][2][2][4];   // [group][row parity][mean / var pass][warp]
  const int grp = threadIdx.x >> 7, t128 = threadIdx.x & 127, wsub = t128 >> 5;
  const int nper = D / 512;
  bool act[2];
  int colk[2], quarter[2];
  float4 ww[2], bb[2];
#pragma unroll
  for (int k = 0; k < 2; ++k) {
    colk[k] = k * 512 + t128 * 4;
    quarter[k] = colk[k] < D / 4 ? 0 : (colk[k] < D / 2 ? 1 : 2);
    act[k] = (k < nper) || (nper == 0 && k == 0 && colk[k] < D);
    ww[k] = act[k] ? __ldg(reinterpret_cast<const float4*>(w + colk[k])) : make_float4(0.f, 0.f, 0.f, 0.f);
    bb[k] = act[k] ? __ldg(reinterpret_cast<const float4*>(b + colk[k])) : make_float4(0.f, 0.f, 0.f, 0.f);
  }
  auto load_row = [&](int row, float4* v) {
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      v[k] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (act[k] && row < M) {
        const long long off = static_cast<long long>(row) * D + colk[k];
        v[k] = *reinterpret_cast<const float4*>(x + off);
        if (x2) {
          const float4 u = *reinterpret_cast<const float4*>(x2 + off);
          v[k] = make_float4(0.5f * (v[k].x + u.x), 0.5f * (v[k].y + u.y), 0.5f * (v[k].z + u.z),
                             0.5f * (v[k].w + u.w));
        }
      }
    }
  };
  const int groups_total = gridDim.x * LNF_GROUPS;
  const int rows_per = (M + groups_total - 1) / groups_total;
  const int row_begin = min(M, (blockIdx.x * LNF_GROUPS + grp) * rows_per), row_end = min(M, row_begin + rows_per);
  RowWalk rw;
  rw.init(sg, row_begin);
  const float inv_d = 1.0f / D;
  int it = 0;
  int refill_row = -1;
  uint8_t* refill_dst = nullptr;
  uint64_t* refill_bar = nullptr;
  auto process = [&](const float4* v) {
    const int row = rw.row;
    float* red = &sred[grp][it & 1][0][0];
    float s = (v[0].x + v[0].y) + (v[0].z + v[0].w) + (v[1].x + v[1].y) + (v[1].z + v[1].w);
    s = warp_sum(s);
    if ((threadIdx.x & 31) == 0) red[wsub] = s;
    asm volatile("bar.sync %0, 128;" ::"r"(grp + 1) : "memory");
    if (refill_row >= 0 && t128 == 0) {   // bulk-ring path: every thread of the group has read this row's slot
      mbar_expect_tx(refill_bar, D * 4);
      bulk_copy_g2s(refill_dst, x + static_cast<long long>(refill_row) * D, D * 4, refill_bar);
    }
    const float mu = ((red[0] + red[1]) + (red[2] + red[3])) * inv_d;
    float q = 0.f;
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      if (act[k]) {
        const float a = v[k].x - mu, c = v[k].y - mu, d = v[k].z - mu, e = v[k].w - mu;
        q += a * a + c * c + d * d + e * e;
      }
    }
    q = warp_sum(q);
    if ((threadIdx.x & 31) == 0) red[4 + wsub] = q;
    asm volatile("bar.sync %0, 128;" ::"r"(grp + 1) : "memory");
    const float rs = rsqrtf(((red[4] + red[5]) + (red[6] + red[7])) * inv_d + eps);
    if (t128 == 0) {
      mean[row] = mu;
      rstd[row] = rs;
    }
    const ShiftRoute rt = shift_route(sg, rw);
    const int r_self = permuted_row32(sg, rw.bidx, rw.t);
    const float nmr = -mu * rs;
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      if (act[k]) {
        const int dtok = quarter[k] == 0 ? rt.dst[0] : (quarter[k] == 1 ? rt.dst[1] : rw.t);
        if (dtok >= 0) {
          uint2 o;
          o.x = pack_bf16x2(fmaf(fmaf(v[k].x, rs, nmr), ww[k].x, bb[k].x), fmaf(fmaf(v[k].y, rs, nmr), ww[k].y, bb[k].y));
          o.y = pack_bf16x2(fmaf(fmaf(v[k].z, rs, nmr), ww[k].z, bb[k].z), fmaf(fmaf(v[k].w, rs, nmr), ww[k].w, bb[k].w));
          *reinterpret_cast<uint2*>(out + static_cast<long long>(permuted_row32(sg, rw.bidx, dtok)) * D + colk[k]) = o;
        }
        const bool z = quarter[k] == 0 ? rt.zero[0] : (quarter[k] == 1 ? rt.zero[1] : false);
        if (z)   // nothing shifts into these channels of row t: they are zero
          *reinterpret_cast<uint2*>(out + static_cast<long long>(r_self) * D + colk[k]) = make_uint2(0u, 0u);
      }
    }
    ++it;
  };
  if (ring_rows > 0) {
    // Rows arrive by bulk async copies into a shared-memory ring, `ring_rows` rows ahead: the bytes in flight live
    // in shared memory instead of registers (a register ring three rows deep left HBM half idle).  A slot is
    // refilled right after the group's first barrier of the row that used it (every thread has read it by then).
    extern __shared__ __align__(128) uint8_t ring_raw[];
    uint8_t* ring = ring_raw + static_cast<size_t>(grp) * ring_rows * D * 4;
    uint64_t* rbar = reinterpret_cast<uint64_t*>(ring_raw + static_cast<size_t>(LNF_GROUPS) * ring_rows * D * 4) +
                     grp * ring_rows;
    if (t128 == 0) {
      for (int s_ = 0; s_ < ring_rows; ++s_) mbar_init(&rbar[s_], 1);
      fence_barrier_init();
    }
    asm volatile("bar.sync %0, 128;" ::"r"(grp + 1) : "memory");
    if (row_begin < row_end) {
      if (t128 == 0) {
        for (int s_ = 0; s_ < ring_rows && row_begin + s_ < row_end; ++s_) {
          mbar_expect_tx(&rbar[s_], D * 4);
          bulk_copy_g2s(ring + static_cast<size_t>(s_) * D * 4, x + static_cast<long long>(row_begin + s_) * D, D * 4,
                        &rbar[s_]);
        }
      }
      int slot = 0;
      uint32_t par = 0;
      for (; rw.row < row_end; rw.next(sg)) {
        mbar_wait(&rbar[slot], par);
        float4 v[2];
#pragma unroll
        for (int k = 0; k < 2; ++k)
          v[k] = act[k] ? *reinterpret_cast<const float4*>(ring + static_cast<size_t>(slot) * D * 4 + colk[k] * 4)
                        : make_float4(0.f, 0.f, 0.f, 0.f);
        refill_row = rw.row + ring_rows < row_end ? rw.row + ring_rows : -1;
        refill_dst = ring + static_cast<size_t>(slot) * D * 4;
        refill_bar = &rbar[slot];
        process(v);
        if (++slot == ring_rows) { slot = 0; par ^= 1; }
      }
    }
    return;
  }
  // register path (final norm over the mean of two streams): ring of four register buffers, loads three rows ahead
  float4 v[4][2];
  if (row_begin < row_end) {   // (uniform per row group, and the group barriers are only used inside)
    load_row(row_begin, v[0]);
    load_row(row_begin + 1 < row_end ? row_begin + 1 : M, v[1]);
    load_row(row_begin + 2 < row_end ? row_begin + 2 : M, v[2]);
    bool done = false;
    while (!done) {
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        if (!done) {
          load_row(rw.row + 3 < row_end ? rw.row + 3 : M, v[(k + 3) & 3]);
          process(v[k]);
          rw.next(sg);
          done = rw.row >= row_end;
        }
      }
    }
  }
}

// Warp-per-row variant for the layer norms inside the stack (single input stream).  One warp owns a contiguous range
// of rows and a private shared-memory ring that bulk async copies fill LNW_RING rows ahead, so the per-row cost is
// pure warp work (no block barriers, ~3x fewer instructions per byte than 128 threads per row) and the bytes in
// flight do not occupy registers.
constexpr int LNW_WARPS = 8, LNW_RING = 1;
__global__ void __launch_bounds__(32 * LNW_WARPS, 4)
ln_shift_fwd_warp_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ b,
                         __nv_bfloat16* __restrict__ out, float* __restrict__ mean, float* __restrict__ rstd, int M,
                         int D, float eps, ShiftGeom sg) {
  extern __shared__ __align__(128) uint8_t ring_raw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nv = D / 128;
  const size_t row_bytes = static_cast<size_t>(D) * 4;
  uint8_t* ring = ring_raw + static_cast<size_t>(warp) * LNW_RING * row_bytes;
  uint64_t* rbar = reinterpret_cast<uint64_t*>(ring_raw + static_cast<size_t>(LNW_WARPS) * LNW_RING * row_bytes) +
                   warp * LNW_RING;
  float* s_w = reinterpret_cast<float*>(ring_raw + static_cast<size_t>(LNW_WARPS) * LNW_RING * row_bytes +
                                        LNW_WARPS * LNW_RING * 8);
  float* s_b = s_w + D;
  for (int c = threadIdx.x; c < D; c += blockDim.x) { s_w[c] = w[c]; s_b[c] = b[c]; }
  if (lane == 0) {
    for (int s_ = 0; s_ < LNW_RING; ++s_) mbar_init(&rbar[s_], 1);
    fence_barrier_init();
  }
  __syncthreads();
  // rows are dealt round-robin to the warps of the grid: at any moment the grid reads one contiguous window of x
  const int warps_total = gridDim.x * LNW_WARPS;
  const int row_begin = blockIdx.x * LNW_WARPS + warp, row_end = M;
  if (row_begin >= row_end) return;
  if (lane == 0) {
    for (int s_ = 0; s_ < LNW_RING && row_begin + s_ * warps_total < row_end; ++s_) {
      mbar_expect_tx(&rbar[s_], static_cast<uint32_t>(row_bytes));
      bulk_copy_g2s(ring + s_ * row_bytes, x + static_cast<long long>(row_begin + s_ * warps_total) * D,
                    static_cast<uint32_t>(row_bytes), &rbar[s_]);
    }
  }
  const float inv_d = 1.0f / D;
  int slot = 0;
  uint32_t par = 0;
  for (int row = row_begin; row < row_end; row += warps_total) {
    RowWalk rw;
    rw.init(sg, row);
    mbar_wait(&rbar[slot], par);
    float4 v[MAXV];
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < MAXV; ++k) {
      v[k] = k < nv ? *reinterpret_cast<const float4*>(ring + slot * row_bytes + (k * 128 + lane * 4) * 4)
                    : make_float4(0.f, 0.f, 0.f, 0.f);
      s += (v[k].x + v[k].y) + (v[k].z + v[k].w);
    }
    __syncwarp();   // every lane has read the slot: refill it with the row LNW_RING ahead
    if (lane == 0 && rw.row + LNW_RING * warps_total < row_end) {
      mbar_expect_tx(&rbar[slot], static_cast<uint32_t>(row_bytes));
      bulk_copy_g2s(ring + slot * row_bytes, x + static_cast<long long>(rw.row + LNW_RING * warps_total) * D,
                    static_cast<uint32_t>(row_bytes), &rbar[slot]);
    }
    const float mu = warp_sum(s) * inv_d;
    float q = 0.f;
#pragma unroll
    for (int k = 0; k < MAXV; ++k) {
      if (k < nv) {
        const float a = v[k].x - mu, c = v[k].y - mu, d = v[k].z - mu, e = v[k].w - mu;
        q += a * a + c * c + d * d + e * e;
      }
    }
    const float rs = rsqrtf(warp_sum(q) * inv_d + eps);
    if (lane == 0) {
      mean[rw.row] = mu;
      rstd[rw.row] = rs;
    }
    const ShiftRoute rt = shift_route(sg, rw);
    __nv_bfloat16* const o_self = out + static_cast<long long>(permuted_row32(sg, rw.bidx, rw.t)) * D;
    __nv_bfloat16* const o_q0 = rt.dst[0] >= 0 ? out + static_cast<long long>(permuted_row32(sg, rw.bidx, rt.dst[0])) * D : nullptr;
    __nv_bfloat16* const o_q1 = rt.dst[1] >= 0 ? out + static_cast<long long>(permuted_row32(sg, rw.bidx, rt.dst[1])) * D : nullptr;
    const float nmr = -mu * rs;
#pragma unroll
    for (int k = 0; k < MAXV; ++k) {
      if (k < nv) {
        const int col = k * 128 + lane * 4;
        const int quarter = col < D / 4 ? 0 : (col < D / 2 ? 1 : 2);
        __nv_bfloat16* const dst = quarter == 0 ? o_q0 : (quarter == 1 ? o_q1 : o_self);
        if (dst) {
          const float4 ww = *reinterpret_cast<const float4*>(s_w + col), bb = *reinterpret_cast<const float4*>(s_b + col);
          uint2 o;
          o.x = pack_bf16x2(fmaf(fmaf(v[k].x, rs, nmr), ww.x, bb.x), fmaf(fmaf(v[k].y, rs, nmr), ww.y, bb.y));
          o.y = pack_bf16x2(fmaf(fmaf(v[k].z, rs, nmr), ww.z, bb.z), fmaf(fmaf(v[k].w, rs, nmr), ww.w, bb.w));
          *reinterpret_cast<uint2*>(dst + col) = o;
        }
        const bool z = quarter == 0 ? rt.zero[0] : (quarter == 1 ? rt.zero[1] : false);
        if (z) *reinterpret_cast<uint2*>(o_self + col) = make_uint2(0u, 0u);   // nothing shifts into these channels
      }
    }
    if (++slot == LNW_RING) { slot = 0; par ^= 1; }
  }
}

// Backward of shift(LN(x)), one pass.  dh: gradient of the shifted output (bf16).  Per row t: dn = gather of dh
// through the transposed shift; dx = rstd * (w*dn - mean(w*dn) - xhat * mean(w*dn*xhat)).
//   mode 0: dx_out[t] += dx           (reversible block: accumulate into the running stream gradient)
//   mode 1: dx_out[t] = dx2_out[t] = 0.5 * dx   (final norm over the mean of the two streams)
// Persistent blocks of 4 row groups x 128 threads: a group walks rows with a grid stride and a thread owns the same
// <= 8 columns on every row, so dw / db accumulate in registers and leave through one shared-memory reduction +
// global atomics per block.  All loads of a row (dh, x, the old dx) are issued before the first use.
// Optional tail of the mode-0 kernel: the stream gradient it has just completed (y = dx_out) is exactly what the
// next block of the backward walk starts from, so its LayerScale/residual prologue is done on the values in
// registers:  d_o = bf16(y * scale),  dscale += colsum(y * o),  dbias += colsum(y * scale)   (resid_bwd_prep)
struct LnBwdNext {
  const __nv_bfloat16* o;   // that block's output before LayerScale (kept from the forward); nullptr: no tail
  const float* scale;
  __nv_bfloat16* d_o;
  float* dscale;
  float* dbias;
};
// Two row groups of 128 threads per block; three blocks per SM (<= 85 registers) without the tail, two (<= 128
// registers) with it - sized so that neither variant spills.
template <int LNB_GROUPS, bool TAIL, bool RING>
__global__ void __launch_bounds__(128 * LNB_GROUPS, TAIL ? 2 : 3)
ln_shift_bwd_kernel(const __nv_bfloat16* __restrict__ dh, const float* __restrict__ x, const float* __restrict__ x2,
                    const float* __restrict__ mean, const float* __restrict__ rstd, const float* __restrict__ w,
                    float* __restrict__ dx_out, float* __restrict__ dx2_out, float* __restrict__ dw,
                    float* __restrict__ db, int M, int D, ShiftGeom sg, int mode, LnBwdNext nx, int ring_rows) {
  __shared__ float sred[LNB_GROUPS][2][4][2];   // [group][parity][warp][s1, s2]
  __shared__ float s_acc[LNB_GROUPS - 1][2][128 * MAXV];
  const int grp = threadIdx.x >> 7;            // row group of the block
  const int t128 = threadIdx.x & 127;          // thread inside the row group
  const int wsub = t128 >> 5;
  const int nper = D / 512;                    // float4 chunks per thread (D = 1024 -> 2, D = 512 -> 1, 256 -> half)
  bool act[2];
  int colk[2];
  float4 ww[2];
#pragma unroll
  for (int k = 0; k < 2; ++k) {
    colk[k] = k * 512 + t128 * 4;
    act[k] = (k < nper) || (nper == 0 && k == 0 && colk[k] < D);
    ww[k] = act[k] ? __ldg(reinterpret_cast<const float4*>(w + colk[k])) : make_float4(0.f, 0.f, 0.f, 0.f);
  }
  float4 aw[2], ab[2];
  aw[0] = aw[1] = ab[0] = ab[1] = make_float4(0.f, 0.f, 0.f, 0.f);
  float4 nsc[2], as2[2], ab2[2];   // tail: next block's LayerScale, its dscale / dbias partial sums
#pragma unroll
  for (int k = 0; k < 2; ++k) {
    as2[k] = ab2[k] = make_float4(0.f, 0.f, 0.f, 0.f);
    nsc[k] = (TAIL && act[k]) ? __ldg(reinterpret_cast<const float4*>(nx.scale + colk[k]))
                              : make_float4(0.f, 0.f, 0.f, 0.f);
  }
  const int groups_total = gridDim.x * LNB_GROUPS;
  const int rows_per = (M + groups_total - 1) / groups_total;   // every row group owns a contiguous range of rows
  const int row_begin = min(M, (blockIdx.x * LNB_GROUPS + grp) * rows_per), row_end = min(M, row_begin + rows_per);
  // Optional shared-memory ring (mode 0): a row's inputs - x, the old dx, the three shifted pieces of dh, the tail's o -
  // arrive by bulk async copies `ring_rows` rows ahead; the bytes in flight then cost no registers and more blocks
  // fit on an SM.  Slot layout: [x : 4D][dx : 4D][dh : 2D][o : 2D] bytes.
  extern __shared__ __align__(128) uint8_t ring_raw[];
  const size_t slot_bytes = static_cast<size_t>(D) * (TAIL ? 12 : 10);
  uint8_t* ring = ring_raw + static_cast<size_t>(grp) * ring_rows * slot_bytes;
  uint64_t* rbar = reinterpret_cast<uint64_t*>(ring_raw + static_cast<size_t>(LNB_GROUPS) * ring_rows * slot_bytes) +
                   grp * ring_rows;
  auto issue_row = [&](const RowWalk& w_, int slot) {   // one thread of the group
    const ShiftRoute rt_ = shift_route(sg, w_);
    uint8_t* dst = ring + slot * slot_bytes;
    const __nv_bfloat16* p_self = dh + static_cast<long long>(permuted_row32(sg, w_.bidx, w_.t)) * D;
    const uint32_t nq = (rt_.dst[0] >= 0 ? 1u : 0u) + (rt_.dst[1] >= 0 ? 1u : 0u);
    mbar_expect_tx(&rbar[slot], static_cast<uint32_t>(D) * (TAIL ? 11u : 9u) + nq * (D / 2));
    bulk_copy_g2s(dst, x + static_cast<long long>(w_.row) * D, D * 4, &rbar[slot]);
    bulk_copy_g2s(dst + D * 4, dx_out + static_cast<long long>(w_.row) * D, D * 4, &rbar[slot]);
    if (rt_.dst[0] >= 0)
      bulk_copy_g2s(dst + D * 8, dh + static_cast<long long>(permuted_row32(sg, w_.bidx, rt_.dst[0])) * D, D / 2, &rbar[slot]);
    if (rt_.dst[1] >= 0)
      bulk_copy_g2s(dst + D * 8 + D / 2, dh + static_cast<long long>(permuted_row32(sg, w_.bidx, rt_.dst[1])) * D + D / 4,
                    D / 2, &rbar[slot]);
    bulk_copy_g2s(dst + D * 9, p_self + D / 2, D, &rbar[slot]);
    if (TAIL) bulk_copy_g2s(dst + D * 10, nx.o + static_cast<long long>(w_.row) * D, D * 2, &rbar[slot]);
  };
  RowWalk rw, rl;
  rw.init(sg, row_begin);
  rl = rw;
  int slot = 0;
  uint32_t rpar = 0;
  if (RING) {
    if (t128 == 0) {
      for (int s_ = 0; s_ < ring_rows; ++s_) mbar_init(&rbar[s_], 1);
      fence_barrier_init();
      for (int s_ = 0; s_ < ring_rows && rl.row < row_end; ++s_, rl.next(sg)) issue_row(rl, s_);
    }
    asm volatile("bar.sync %0, 128;" ::"r"(grp + 1) : "memory");
  }
  for (int it = 0; rw.row < row_end; rw.next(sg), ++it) {
    const int row = rw.row;
    const ShiftRoute rt = shift_route(sg, rw);
    const long long r_self = static_cast<long long>(permuted_row32(sg, rw.bidx, rw.t)) * D;
    const long long r_q0 = rt.dst[0] >= 0 ? static_cast<long long>(permuted_row32(sg, rw.bidx, rt.dst[0])) * D : -1;
    const long long r_q1 = rt.dst[1] >= 0 ? static_cast<long long>(permuted_row32(sg, rw.bidx, rt.dst[1])) * D : -1;
    const float mu = mean[row], rs = rstd[row];
    uint2 u[2], uo[2];
    float4 v[2], v2[2], old[2];
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      u[k] = uo[k] = make_uint2(0u, 0u);
      v[k] = v2[k] = old[k] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (act[k] && RING) {
        if (k == 0) mbar_wait(&rbar[slot], rpar);
        const int col = colk[k];
        const uint8_t* sl = ring + slot * slot_bytes;
        const long long src = col < D / 4 ? r_q0 : (col < D / 2 ? r_q1 : r_self);
        if (src >= 0) u[k] = *reinterpret_cast<const uint2*>(sl + D * 8 + col * 2);
        v[k] = *reinterpret_cast<const float4*>(sl + col * 4);
        old[k] = *reinterpret_cast<const float4*>(sl + D * 4 + col * 4);
        if (TAIL) uo[k] = *reinterpret_cast<const uint2*>(sl + D * 10 + col * 2);
      } else if (act[k]) {
        const int col = colk[k];
        const long long src = col < D / 4 ? r_q0 : (col < D / 2 ? r_q1 : r_self);
        if (src >= 0) u[k] = *reinterpret_cast<const uint2*>(dh + src + col);
        v[k] = *reinterpret_cast<const float4*>(x + static_cast<long long>(row) * D + col);
        if (x2) v2[k] = *reinterpret_cast<const float4*>(x2 + static_cast<long long>(row) * D + col);
        if (mode == 0) old[k] = *reinterpret_cast<const float4*>(dx_out + static_cast<long long>(row) * D + col);
        if (TAIL) uo[k] = *reinterpret_cast<const uint2*>(nx.o + static_cast<long long>(row) * D + col);
      }
    }
    float4 dn[2], xh[2];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      const float2 a = unpack_bf16x2(u[k].x), c = unpack_bf16x2(u[k].y);
      float4 vv = v[k];
      if (x2) vv = make_float4(0.5f * (vv.x + v2[k].x), 0.5f * (vv.y + v2[k].y), 0.5f * (vv.z + v2[k].z),
                               0.5f * (vv.w + v2[k].w));
      xh[k] = act[k] ? make_float4((vv.x - mu) * rs, (vv.y - mu) * rs, (vv.z - mu) * rs, (vv.w - mu) * rs)
                     : make_float4(0.f, 0.f, 0.f, 0.f);
      aw[k].x += a.x * xh[k].x; aw[k].y += a.y * xh[k].y; aw[k].z += c.x * xh[k].z; aw[k].w += c.y * xh[k].w;
      ab[k].x += a.x; ab[k].y += a.y; ab[k].z += c.x; ab[k].w += c.y;
      dn[k] = make_float4(a.x * ww[k].x, a.y * ww[k].y, c.x * ww[k].z, c.y * ww[k].w);
      s1 += dn[k].x + dn[k].y + dn[k].z + dn[k].w;
      s2 += dn[k].x * xh[k].x + dn[k].y * xh[k].y + dn[k].z * xh[k].z + dn[k].w * xh[k].w;
    }
    s1 = warp_sum(s1);
    s2 = warp_sum(s2);
    float* red = &sred[grp][it & 1][0][0];
    if ((threadIdx.x & 31) == 0) { red[wsub * 2] = s1; red[wsub * 2 + 1] = s2; }
    asm volatile("bar.sync %0, 128;" ::"r"(grp + 1) : "memory");   // the row group's own barrier
    if (RING) {   // every thread of the group holds this row in registers: refill its slot
      if (t128 == 0 && rl.row < row_end) { issue_row(rl, slot); rl.next(sg); }
      if (++slot == ring_rows) { slot = 0; rpar ^= 1; }
    }
    s1 = (red[0] + red[2] + red[4] + red[6]) / D;
    s2 = (red[1] + red[3] + red[5] + red[7]) / D;
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      if (act[k]) {
        float4 d;
        d.x = rs * (dn[k].x - s1 - xh[k].x * s2);
        d.y = rs * (dn[k].y - s1 - xh[k].y * s2);
        d.z = rs * (dn[k].z - s1 - xh[k].z * s2);
        d.w = rs * (dn[k].w - s1 - xh[k].w * s2);
        float4* p = reinterpret_cast<float4*>(dx_out + static_cast<long long>(row) * D + colk[k]);
        if (mode == 0) {
          const float4 y = make_float4(old[k].x + d.x, old[k].y + d.y, old[k].z + d.z, old[k].w + d.w);
          *p = y;
          if (TAIL) {
            const float2 o01 = unpack_bf16x2(uo[k].x), o23 = unpack_bf16x2(uo[k].y);
            const float4 ys = make_float4(y.x * nsc[k].x, y.y * nsc[k].y, y.z * nsc[k].z, y.w * nsc[k].w);
            as2[k].x += y.x * o01.x; as2[k].y += y.y * o01.y; as2[k].z += y.z * o23.x; as2[k].w += y.w * o23.y;
            ab2[k].x += ys.x; ab2[k].y += ys.y; ab2[k].z += ys.z; ab2[k].w += ys.w;
            uint2 w2;
            w2.x = pack_bf16x2(ys.x, ys.y);
            w2.y = pack_bf16x2(ys.z, ys.w);
            *reinterpret_cast<uint2*>(nx.d_o + static_cast<long long>(row) * D + colk[k]) = w2;
          }
        } else {
          const float4 hlf = make_float4(0.5f * d.x, 0.5f * d.y, 0.5f * d.z, 0.5f * d.w);
          *p = hlf;
          *reinterpret_cast<float4*>(dx2_out + static_cast<long long>(row) * D + colk[k]) = hlf;
        }
      }
    }
  }
  // block reduction of dw / db (then, with a tail, of dscale / dbias through the same shared buffer): groups 1..
  // park their partials in shared memory, group 0 adds and publishes
  auto block_reduce_publish = [&](float4* pa, float4* pb, float* outa, float* outb) {
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      if (act[k] && grp > 0) {
        *reinterpret_cast<float4*>(&s_acc[grp - 1][0][colk[k]]) = pa[k];
        *reinterpret_cast<float4*>(&s_acc[grp - 1][1][colk[k]]) = pb[k];
      }
    }
    __syncthreads();
    if (grp == 0) {
#pragma unroll
      for (int k = 0; k < 2; ++k) {
        if (act[k]) {
          for (int g2 = 0; g2 < LNB_GROUPS - 1; ++g2) {
            const float4 qa = *reinterpret_cast<const float4*>(&s_acc[g2][0][colk[k]]);
            const float4 qb = *reinterpret_cast<const float4*>(&s_acc[g2][1][colk[k]]);
            pa[k].x += qa.x; pa[k].y += qa.y; pa[k].z += qa.z; pa[k].w += qa.w;
            pb[k].x += qb.x; pb[k].y += qb.y; pb[k].z += qb.z; pb[k].w += qb.w;
          }
          float* oa = outa + colk[k];
          float* ob = outb + colk[k];
          atomicAdd(oa, pa[k].x); atomicAdd(oa + 1, pa[k].y); atomicAdd(oa + 2, pa[k].z); atomicAdd(oa + 3, pa[k].w);
          atomicAdd(ob, pb[k].x); atomicAdd(ob + 1, pb[k].y); atomicAdd(ob + 2, pb[k].z); atomicAdd(ob + 3, pb[k].w);
        }
      }
    }
  };
  block_reduce_publish(aw, ab, dw, db);
  if (TAIL) {
    __syncthreads();
    block_reduce_publish(as2, ab2, nx.dscale, nx.dbias);
  }
}

// d_o = bf16(dy * scale); dscale += colsum(dy * o); dbias += colsum(dy * scale).   thread = 4 columns
__global__ void __launch_bounds__(256)
resid_bwd_prep_kernel(const float* __restrict__ dy, const __nv_bfloat16* __restrict__ o,
                      const float* __restrict__ scale, __nv_bfloat16* __restrict__ d_o, float* __restrict__ dscale,
                      float* __restrict__ dbias, int M, int D) {
  const int cols_per_block = blockDim.x * 4;
  const int col = (blockIdx.x * cols_per_block) % D + threadIdx.x * 4;
  const int stripes = D / cols_per_block;
  const int rsplit = blockIdx.x / stripes, nsplit = gridDim.x / stripes;
  const float4 sc = __ldg(reinterpret_cast<const float4*>(scale + col));
  float4 as = make_float4(0.f, 0.f, 0.f, 0.f), ab = make_float4(0.f, 0.f, 0.f, 0.f);
  auto one = [&](const float4& g, const uint2& u, long long off) {
    const float2 o01 = unpack_bf16x2(u.x), o23 = unpack_bf16x2(u.y);
    as.x += g.x * o01.x; as.y += g.y * o01.y; as.z += g.z * o23.x; as.w += g.w * o23.y;
    const float4 d = make_float4(g.x * sc.x, g.y * sc.y, g.z * sc.z, g.w * sc.w);
    ab.x += d.x; ab.y += d.y; ab.z += d.z; ab.w += d.w;
    uint2 w;
    w.x = pack_bf16x2(d.x, d.y);
    w.y = pack_bf16x2(d.z, d.w);
    *reinterpret_cast<uint2*>(d_o + off) = w;
  };
  int row = rsplit;
  for (; row + 3 * nsplit < M; row += 4 * nsplit) {   // four rows in flight per thread
    float4 g[4];
    uint2 u[4];
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      const long long off = static_cast<long long>(row + r * nsplit) * D + col;
      g[r] = *reinterpret_cast<const float4*>(dy + off);
      u[r] = *reinterpret_cast<const uint2*>(o + off);
    }
#pragma unroll
    for (int r = 0; r < 4; ++r) one(g[r], u[r], static_cast<long long>(row + r * nsplit) * D + col);
  }
  for (; row < M; row += nsplit) {
    const long long off = static_cast<long long>(row) * D + col;
    one(*reinterpret_cast<const float4*>(dy + off), *reinterpret_cast<const uint2*>(o + off), off);
  }
  atomicAdd(dscale + col, as.x); atomicAdd(dscale + col + 1, as.y);
  atomicAdd(dscale + col + 2, as.z); atomicAdd(dscale + col + 3, as.w);
  atomicAdd(dbias + col, ab.x); atomicAdd(dbias + col + 1, ab.y);
  atomicAdd(dbias + col + 2, ab.z); atomicAdd(dbias + col + 3, ab.w);
}

// out[col] += alpha * sum_rows x[row, col]      (bf16 input, thread = 8 columns)
__global__ void __launch_bounds__(256)
colsum_bf16_kernel(const __nv_bfloat16* __restrict__ x, long long ld, float* __restrict__ out, int M, int N,
                   float alpha) {
  const int stripes = (N + 2047) / 2048;
  const int stripe = blockIdx.x % stripes, rsplit = blockIdx.x / stripes, nsplit = gridDim.x / stripes;
  const int col = stripe * 2048 + threadIdx.x * 8;
  if (col >= N) return;
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  const bool full = col + 8 <= N;
  auto add8 = [&](const uint4& u) {
    float2 f;
    f = unpack_bf16x2(u.x); acc[0] += f.x; acc[1] += f.y;
    f = unpack_bf16x2(u.y); acc[2] += f.x; acc[3] += f.y;
    f = unpack_bf16x2(u.z); acc[4] += f.x; acc[5] += f.y;
    f = unpack_bf16x2(u.w); acc[6] += f.x; acc[7] += f.y;
  };
  int row = rsplit;
  if (full) {
    // four rows in flight per thread (one 16-byte load per thread and iteration leaves HBM half idle)
    for (; row + 3 * nsplit < M; row += 4 * nsplit) {
      uint4 u[4];
#pragma unroll
      for (int r = 0; r < 4; ++r)
        u[r] = *reinterpret_cast<const uint4*>(x + static_cast<long long>(row + r * nsplit) * ld + col);
#pragma unroll
      for (int r = 0; r < 4; ++r) add8(u[r]);
    }
  }
  for (; row < M; row += nsplit) {
    const __nv_bfloat16* p = x + static_cast<long long>(row) * ld + col;
    if (full) {
      add8(*reinterpret_cast<const uint4*>(p));
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j)
        if (col + j < N) acc[j] += __bfloat162float(p[j]);
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j)
    if (col + j < N) atomicAdd(out + col + j, alpha * acc[j]);
}

// ------------------------------------------------------------------------------------------------
// tied embedding
// ------------------------------------------------------------------------------------------------
__global__ void embed_fwd_kernel(const long long* __restrict__ ids, const float* __restrict__ W,
                                 float* __restrict__ x1, float* __restrict__ x2, int M, int D) {
  const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (row >= M) return;
  const long long id = ids[row];
  for (int c = lane * 4; c < D; c += 128) {
    const float4 v = __ldg(reinterpret_cast<const float4*>(W + id * D + c));
    *reinterpret_cast<float4*>(x1 + static_cast<long long>(row) * D + c) = v;
    if (x2) *reinterpret_cast<float4*>(x2 + static_cast<long long>(row) * D + c) = v;
  }
}
__global__ void embed_bwd_kernel(const long long* __restrict__ ids, const float* __restrict__ d1,
                                 const float* __restrict__ d2, float* __restrict__ dW, int M, int D) {
  const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (row >= M) return;
  const long long id = ids[row];
  for (int c = lane * 4; c < D; c += 128) {
    float4 v = *reinterpret_cast<const float4*>(d1 + static_cast<long long>(row) * D + c);
    if (d2) {
      const float4 u = *reinterpret_cast<const float4*>(d2 + static_cast<long long>(row) * D + c);
      v.x += u.x; v.y += u.y; v.z += u.z; v.w += u.w;
    }
    float* p = dW + id * D + c;
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
                 : "memory");
  }
}

// ------------------------------------------------------------------------------------------------
// segment cross-entropy: one block per row over n_classes logits (bf16, ld elements per row).
// loss_acc += row_weight * (lse - logit[label]);   logits <- row_weight * grad_scale * (softmax - onehot)
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
ce_fwd_bwd_kernel(__nv_bfloat16* __restrict__ logits, long long ld, const long long* __restrict__ labels,
                  int n_classes, float row_weight, const float* __restrict__ grad_scale_ptr,
                  float* __restrict__ loss_acc, int write_grad) {
  extern __shared__ float srow[];  // n_classes floats
  __shared__ float sred[8];
  __shared__ float sbc;
  __nv_bfloat16* row = logits + static_cast<long long>(blockIdx.x) * ld;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  float mx = -1e30f;
  const int n8 = n_classes / 8;
  for (int i = tid; i < n8; i += blockDim.x) {
    const uint4 u = *reinterpret_cast<const uint4*>(row + i * 8);
    float2 f;
    float* d = srow + i * 8;
    f = unpack_bf16x2(u.x); d[0] = f.x; d[1] = f.y;
    f = unpack_bf16x2(u.y); d[2] = f.x; d[3] = f.y;
    f = unpack_bf16x2(u.z); d[4] = f.x; d[5] = f.y;
    f = unpack_bf16x2(u.w); d[6] = f.x; d[7] = f.y;
#pragma unroll
    for (int j = 0; j < 8; ++j) mx = fmaxf(mx, d[j]);
  }
  for (int i = n8 * 8 + tid; i < n_classes; i += blockDim.x) {
    const float v = __bfloat162float(row[i]);
    srow[i] = v;
    mx = fmaxf(mx, v);
  }
  mx = warp_max(mx);
  if (lane == 0) sred[warp] = mx;
  __syncthreads();
  if (tid == 0) {
    float m = sred[0];
    for (int i = 1; i < (blockDim.x >> 5); ++i) m = fmaxf(m, sred[i]);
    sbc = m;
  }
  __syncthreads();
  mx = sbc;
  const long long label = labels[blockIdx.x];
  const float x_label = srow[label];
  __syncthreads();
  float s = 0.f;
  for (int i = tid; i < n_classes; i += blockDim.x) {
    const float e = __expf(srow[i] - mx);
    srow[i] = e;
    s += e;
  }
  s = warp_sum(s);
  __syncthreads();
  if (lane == 0) sred[warp] = s;
  __syncthreads();
  if (tid == 0) {
    float t = 0.f;
    for (int i = 0; i < (blockDim.x >> 5); ++i) t += sred[i];
    sbc = t;
  }
  __syncthreads();
  const float sum = sbc;
  if (tid == 0) atomicAdd(loss_acc, row_weight * (logf(sum) + mx - x_label));
  if (!write_grad) return;
  const float k = row_weight * (grad_scale_ptr ? __ldg(grad_scale_ptr) : 1.0f), inv = 1.0f / sum;
  __syncthreads();
  for (int i = tid; i < n8; i += blockDim.x) {
    float g[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int c = i * 8 + j;
      g[j] = k * (srow[c] * inv - (c == label ? 1.f : 0.f));
    }
    uint4 u;
    u.x = pack_bf16x2(g[0], g[1]); u.y = pack_bf16x2(g[2], g[3]);
    u.z = pack_bf16x2(g[4], g[5]); u.w = pack_bf16x2(g[6], g[7]);
    *reinterpret_cast<uint4*>(row + i * 8) = u;
  }
  for (int i = n8 * 8 + tid; i < n_classes; i += blockDim.x)
    row[i] = __float2bfloat16(k * (srow[i] * inv - (i == label ? 1.f : 0.f)));
}

__global__ void cast_f32_bf16_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, long long n) {
  const long long i = (blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x) * 4;
  if (i + 4 <= n) {
    const float4 v = *reinterpret_cast<const float4*>(src + i);
    uint2 u;
    u.x = pack_bf16x2(v.x, v.y);
    u.y = pack_bf16x2(v.z, v.w);
    *reinterpret_cast<uint2*>(dst + i) = u;
  } else {
    for (long long j = i; j < n; ++j) dst[j] = __float2bfloat16(src[j]);
  }
}

static inline int err() { return cudaGetLastError() == cudaSuccess ? 0 : -40; }

}  // namespace b200

using namespace b200;

extern "C" int b200_ln_shift_fwd(const float* x, const float* x2, const float* w, const float* b, void* out,
                                 float* mean, float* rstd, int M, int D, int seq, int T, int F, int shift,
                                 int row_split, float eps, void* stream) {
  if (D % 128 || D > 128 * MAXV) return -41;
  if (D % 512 && D != 256) return -41;  // 4 warps per row: D in {256, 512, 1024}
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  ShiftGeom sg{seq, T, F, shift, row_split, M / seq};
  int blocks = num_sms() * 3;   // persistent: three 8-warp blocks per SM
  if (blocks * LNF_GROUPS > M) blocks = (M + LNF_GROUPS - 1) / LNF_GROUPS;
  int ring_rows = 0;
  size_t smem = 0;
  if (!x2 && reinterpret_cast<uintptr_t>(x) % 16 == 0) {
    const size_t wsmem = static_cast<size_t>(LNW_WARPS) * LNW_RING * D * 4 + LNW_WARPS * LNW_RING * 8 + 2 * D * 4;
    static bool configured = false;
    if (!configured) {
      if (cudaFuncSetAttribute(ln_shift_fwd_warp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 112 * 1024) !=
          cudaSuccess)
        return -40;
      configured = true;
    }
    int wblocks = num_sms() * 4;   // four 8-warp blocks per SM (the next row of a warp lands while it works on this one)
    if (wblocks * LNW_WARPS > M) wblocks = (M + LNW_WARPS - 1) / LNW_WARPS;
    ln_shift_fwd_warp_kernel<<<wblocks, 32 * LNW_WARPS, wsmem, st>>>(x, w, b, reinterpret_cast<__nv_bfloat16*>(out), mean,
                                                                     rstd, M, D, eps, sg);
    return err();
  }
  ln_shift_fwd_kernel<<<blocks, 128 * LNF_GROUPS, smem, st>>>(x, x2, w, b, reinterpret_cast<__nv_bfloat16*>(out), mean,
                                                              rstd, M, D, eps, sg, ring_rows);
  return err();
}

static int ln_shift_bwd_launch(const void* dh, const float* x, const float* x2, const float* mean, const float* rstd,
                               const float* w, float* dx, float* dx2, float* dw, float* db, int M, int D, int seq,
                               int T, int F, int shift, int row_split, int mode, const LnBwdNext& nx, void* stream) {
  if (D % 128 || D > 128 * MAXV) return -41;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  ShiftGeom sg{seq, T, F, shift, row_split, M / seq};
  if (D % 512 && D != 256) return -41;  // 4 warps per row: D in {256, 512, 1024}
  if (nx.o && mode != 0) return -41;
  const __nv_bfloat16* dhp = reinterpret_cast<const __nv_bfloat16*>(dh);
  const bool ring_ok = mode == 0 && !x2 && D % 64 == 0 && reinterpret_cast<uintptr_t>(x) % 16 == 0 &&
                       reinterpret_cast<uintptr_t>(dx) % 16 == 0 && reinterpret_cast<uintptr_t>(dh) % 16 == 0 &&
                       (!nx.o || reinterpret_cast<uintptr_t>(nx.o) % 16 == 0);
  // rows in flight per 128-thread group (DALLE_B200_LNB_RING overrides: experiments)
  static const int ring_env = [] { const char* e = getenv("DALLE_B200_LNB_RING"); return e ? atoi(e) : 0; }();
  const int ring_rows = ring_env > 0 ? ring_env : 2;
  const size_t ring_smem = 2 * (static_cast<size_t>(ring_rows) * D * (nx.o ? 12 : 10) + ring_rows * 8);
  int blocks = num_sms() * 3;   // persistent: three 8-warp blocks per SM
  if (blocks * 2 > M) blocks = (M + 1) / 2;
  if (ring_ok && ring_smem <= 112 * 1024) {
    static bool configured = false;
    if (!configured) {
      cudaFuncSetAttribute(ln_shift_bwd_kernel<2, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 112 * 1024);
      cudaFuncSetAttribute(ln_shift_bwd_kernel<2, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 112 * 1024);
      configured = true;
    }
    if (nx.o && blocks > num_sms() * 2) blocks = num_sms() * 2;   // (register-limited to two blocks per SM)
    if (nx.o)
      ln_shift_bwd_kernel<2, true, true><<<blocks, 256, ring_smem, st>>>(dhp, x, x2, mean, rstd, w, dx, dx2, dw, db, M, D, sg, mode, nx, ring_rows);
    else
      ln_shift_bwd_kernel<2, false, true><<<blocks, 256, ring_smem, st>>>(dhp, x, x2, mean, rstd, w, dx, dx2, dw, db, M, D, sg, mode, nx, ring_rows);
  } else if (nx.o) {
    blocks = num_sms() * 2;   // two 8-warp blocks per SM (no register spills with the tail)
    if (blocks * 2 > M) blocks = (M + 1) / 2;
    ln_shift_bwd_kernel<2, true, false><<<blocks, 256, 0, st>>>(dhp, x, x2, mean, rstd, w, dx, dx2, dw, db, M, D, sg, mode, nx, 0);
  } else {
    ln_shift_bwd_kernel<2, false, false><<<blocks, 256, 0, st>>>(dhp, x, x2, mean, rstd, w, dx, dx2, dw, db, M, D, sg, mode, nx, 0);
  }
  return err();
}

extern "C" int b200_ln_shift_bwd(const void* dh, const float* x, const float* x2, const float* mean,
                                 const float* rstd, const float* w, float* dx, float* dx2, float* dw, float* db, int M,
                                 int D, int seq, int T, int F, int shift, int row_split, int mode, void* stream) {
  LnBwdNext nx{nullptr, nullptr, nullptr, nullptr, nullptr};
  return ln_shift_bwd_launch(dh, x, x2, mean, rstd, w, dx, dx2, dw, db, M, D, seq, T, F, shift, row_split, mode, nx,
                             stream);
}

// mode 0 with the next block's residual prologue fused in (see LnBwdNext)
extern "C" int b200_ln_shift_bwd_prep(const void* dh, const float* x, const float* mean, const float* rstd,
                                      const float* w, float* dx, float* dw, float* db, const void* o_next,
                                      const float* scale_next, void* d_o_next, float* dscale_next, float* dbias_next,
                                      int M, int D, int seq, int T, int F, int shift, int row_split, void* stream) {
  LnBwdNext nx{reinterpret_cast<const __nv_bfloat16*>(o_next), scale_next, reinterpret_cast<__nv_bfloat16*>(d_o_next),
               dscale_next, dbias_next};
  return ln_shift_bwd_launch(dh, x, nullptr, mean, rstd, w, dx, nullptr, dw, db, M, D, seq, T, F, shift, row_split, 0,
                             nx, stream);
}

extern "C" int b200_resid_bwd_prep(const float* dy, const void* o, const float* scale, void* d_o, float* dscale,
                                   float* dbias, int M, int D, void* stream) {
  int threads = 256;
  if (D % 4) return -42;
  if (D < threads * 4) threads = D / 4;
  if (D % (threads * 4)) return -42;
  const int stripes = D / (threads * 4);
  int nsplit = (num_sms() * 3) / stripes;   // few, long-lived blocks: the per-thread atomics at the end are the tail
  if (nsplit > M) nsplit = M;
  if (nsplit < 1) nsplit = 1;
  resid_bwd_prep_kernel<<<stripes * nsplit, threads, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      dy, reinterpret_cast<const __nv_bfloat16*>(o), scale, reinterpret_cast<__nv_bfloat16*>(d_o), dscale, dbias, M, D);
  return err();
}

extern "C" int b200_colsum_bf16(const void* x, long long ld, float* out, int M, int N, float alpha, void* stream) {
  const int stripes = (N + 2047) / 2048;
  int nsplit = (num_sms() * 3) / stripes;   // few, long-lived blocks: the per-thread atomics at the end are the tail
  if (nsplit > M) nsplit = M;
  if (nsplit < 1) nsplit = 1;
  colsum_bf16_kernel<<<stripes * nsplit, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const __nv_bfloat16*>(x), ld, out, M, N, alpha);
  return err();
}

extern "C" int b200_embed_fwd(const long long* ids, const float* W, float* x1, float* x2, int M, int D, void* stream) {
  if (D % 4) return -43;
  const int threads = 256, blocks = (M + 7) / 8;
  embed_fwd_kernel<<<blocks, threads, 0, reinterpret_cast<cudaStream_t>(stream)>>>(ids, W, x1, x2, M, D);
  return err();
}

extern "C" int b200_embed_bwd(const long long* ids, const float* d1, const float* d2, float* dW, int M, int D,
                              void* stream) {
  if (D % 4) return -43;
  const int threads = 256, blocks = (M + 7) / 8;
  embed_bwd_kernel<<<blocks, threads, 0, reinterpret_cast<cudaStream_t>(stream)>>>(ids, d1, d2, dW, M, D);
  return err();
}

extern "C" int b200_ce_fwd_bwd(void* logits, long long ld, const long long* labels, int rows, int n_classes,
                               float row_weight, const float* grad_scale, float* loss_acc, int write_grad,
                               void* stream) {
  if (ld % 8) return -44;
  const size_t smem = static_cast<size_t>(n_classes) * sizeof(float);
  if (smem > 200 * 1024) return -45;
  static bool configured = false;
  if (!configured) {
    if (cudaFuncSetAttribute(ce_fwd_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024) != cudaSuccess)
      return -46;
    configured = true;
  }
  ce_fwd_bwd_kernel<<<rows, 256, smem, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<__nv_bfloat16*>(logits), ld, labels, n_classes, row_weight, grad_scale, loss_acc, write_grad);
  return err();
}

extern "C" int b200_cast_f32_bf16(const float* src, void* dst, long long n, void* stream) {
  const int threads = 256;
  const long long blocks = (n / 4 + threads) / threads;
  cast_f32_bf16_kernel<<<static_cast<unsigned>(blocks), threads, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      src, reinterpret_cast<__nv_bfloat16*>(dst), n);
  return err();
}
