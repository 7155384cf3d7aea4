// Memory-bound kernels around the GEMMs: tied-embedding gather / scatter-add, LayerNorm fused with the DALL-E
// token shift (forward + backward), LayerScale/residual backward prep, column sums, masked-segment cross-entropy.
// Replaces ATen embedding / layer_norm / pad+chunk+cat / mul / cross_entropy calls of the reference model
// (SURVEY K1, K2, K3, K10, K13).  One warp per row, 128-bit accesses, fp32 math.
#include "sm100_ptx.cuh"
#include "b200_api.h"

namespace b200 {

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

// ------------------------------------------------------------------------------------------------
// LayerNorm statistics: mean / rstd per row of x (or of 0.5*(x + x2))
// ------------------------------------------------------------------------------------------------
__global__ void ln_stats_kernel(const float* __restrict__ x, const float* __restrict__ x2, float* __restrict__ mean,
                                float* __restrict__ rstd, int M, int D, float eps) {
  const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (row >= M) return;
  const int nv = D / 128;
  float4 v[MAXV];
  float s = 0.f;
#pragma unroll
  for (int k = 0; k < MAXV; ++k) {
    if (k < nv) {
      const long long off = static_cast<long long>(row) * D + k * 128 + lane * 4;
      v[k] = *reinterpret_cast<const float4*>(x + off);
      if (x2) {
        const float4 w = *reinterpret_cast<const float4*>(x2 + off);
        v[k].x = 0.5f * (v[k].x + w.x); v[k].y = 0.5f * (v[k].y + w.y);
        v[k].z = 0.5f * (v[k].z + w.z); v[k].w = 0.5f * (v[k].w + w.w);
      }
      s += v[k].x + v[k].y + v[k].z + v[k].w;
    }
  }
  const float mu = warp_sum(s) / D;
  float q = 0.f;
#pragma unroll
  for (int k = 0; k < MAXV; ++k) {
    if (k < nv) {
      const float a = v[k].x - mu, b = v[k].y - mu, c = v[k].z - mu, d = v[k].w - mu;
      q += a * a + b * b + c * c + d * d;
    }
  }
  const float var = warp_sum(q) / D;
  if (lane == 0) {
    mean[row] = mu;
    rstd[row] = rsqrtf(var + eps);
  }
}

// out[t, col] = LN(x)[src(t,col), col]  (bf16)
__global__ void ln_shift_apply_kernel(const float* __restrict__ x, const float* __restrict__ x2,
                                      const float* __restrict__ mean, const float* __restrict__ rstd,
                                      const float* __restrict__ w, const float* __restrict__ b,
                                      __nv_bfloat16* __restrict__ out, int M, int D, ShiftGeom sg) {
  const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (row >= M) return;
  const int nv = D / 128;
  const int bidx = row / sg.seq, t = row - bidx * sg.seq;
#pragma unroll
  for (int k = 0; k < MAXV; ++k) {
    if (k < nv) {
      const int col = k * 128 + lane * 4;
      const int st = shift_src(sg, t, col, D);
      uint2 o = make_uint2(0u, 0u);
      if (st >= 0) {
        const long long srow = static_cast<long long>(bidx) * sg.seq + st;
        float4 v = *reinterpret_cast<const float4*>(x + srow * D + col);
        if (x2) {
          const float4 u = *reinterpret_cast<const float4*>(x2 + srow * D + col);
          v.x = 0.5f * (v.x + u.x); v.y = 0.5f * (v.y + u.y); v.z = 0.5f * (v.z + u.z); v.w = 0.5f * (v.w + u.w);
        }
        const float mu = mean[srow], rs = rstd[srow];
        const float4 ww = __ldg(reinterpret_cast<const float4*>(w + col));
        const float4 bb = __ldg(reinterpret_cast<const float4*>(b + col));
        o.x = pack_bf16x2((v.x - mu) * rs * ww.x + bb.x, (v.y - mu) * rs * ww.y + bb.y);
        o.y = pack_bf16x2((v.z - mu) * rs * ww.z + bb.z, (v.w - mu) * rs * ww.w + bb.w);
      }
      *reinterpret_cast<uint2*>(out + permuted_row(sg, bidx, t) * D + col) = o;
    }
  }
}

// Backward of shift(LN(x)).  dh: gradient of the shifted output (bf16).  Per row t: dn = gather of dh through the
// transposed shift; dx = rstd * (w*dn - mean(w*dn) - xhat * mean(w*dn*xhat)).
//   mode 0: dx_out[t] += dx           (reversible block: accumulate into the running stream gradient)
//   mode 1: dx_out[t] = dx2_out[t] = 0.5 * dx   (final norm over the mean of the two streams)
// Row kernel (4 warps per row -> 8 elements per thread, full occupancy) + column kernel for dw, db.
__global__ void __launch_bounds__(256)
ln_shift_bwd_dx_kernel(const __nv_bfloat16* __restrict__ dh, const float* __restrict__ x, const float* __restrict__ x2,
                       const float* __restrict__ mean, const float* __restrict__ rstd, const float* __restrict__ w,
                       float* __restrict__ dx_out, float* __restrict__ dx2_out, int M, int D, ShiftGeom sg, int mode) {
  __shared__ float sred[2][4][2];
  const int sub = threadIdx.x >> 7;            // which of the block's 2 rows
  const int t128 = threadIdx.x & 127;          // thread inside the row group
  const int wsub = t128 >> 5;                  // warp inside the row group
  const int row = blockIdx.x * 2 + sub;
  const bool live = row < M;
  const int nper = D / 512;                    // float4 chunks per thread (D = 1024 -> 2, D = 512 -> 1)
  const int rrow = live ? row : 0;
  const int bidx = rrow / sg.seq, t = rrow - bidx * sg.seq;
  const float mu = mean[rrow], rs = rstd[rrow];
  float4 dn[2], xh[2];
  float s1 = 0.f, s2 = 0.f;
#pragma unroll
  for (int k = 0; k < 2; ++k) {
    const int col = k * 512 + t128 * 4;
    if (k < nper || (nper == 0 && k == 0 && col < D)) {
      const int dt = shift_dst(sg, t, col, D);
      float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
      if (dt >= 0) {
        const uint2 u = *reinterpret_cast<const uint2*>(dh + permuted_row(sg, bidx, dt) * D + col);
        const float2 a = unpack_bf16x2(u.x), c = unpack_bf16x2(u.y);
        g = make_float4(a.x, a.y, c.x, c.y);
      }
      float4 v = *reinterpret_cast<const float4*>(x + static_cast<long long>(rrow) * D + col);
      if (x2) {
        const float4 u2 = *reinterpret_cast<const float4*>(x2 + static_cast<long long>(rrow) * D + col);
        v.x = 0.5f * (v.x + u2.x); v.y = 0.5f * (v.y + u2.y); v.z = 0.5f * (v.z + u2.z); v.w = 0.5f * (v.w + u2.w);
      }
      const float4 ww = __ldg(reinterpret_cast<const float4*>(w + col));
      xh[k] = make_float4((v.x - mu) * rs, (v.y - mu) * rs, (v.z - mu) * rs, (v.w - mu) * rs);
      dn[k] = make_float4(g.x * ww.x, g.y * ww.y, g.z * ww.z, g.w * ww.w);
      s1 += dn[k].x + dn[k].y + dn[k].z + dn[k].w;
      s2 += dn[k].x * xh[k].x + dn[k].y * xh[k].y + dn[k].z * xh[k].z + dn[k].w * xh[k].w;
    } else {
      dn[k] = xh[k] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
  }
  s1 = warp_sum(s1);
  s2 = warp_sum(s2);
  if ((threadIdx.x & 31) == 0) { sred[sub][wsub][0] = s1; sred[sub][wsub][1] = s2; }
  __syncthreads();
  s1 = (sred[sub][0][0] + sred[sub][1][0] + sred[sub][2][0] + sred[sub][3][0]) / D;
  s2 = (sred[sub][0][1] + sred[sub][1][1] + sred[sub][2][1] + sred[sub][3][1]) / D;
  if (!live) return;
#pragma unroll
  for (int k = 0; k < 2; ++k) {
    const int col = k * 512 + t128 * 4;
    if (k < nper || (nper == 0 && k == 0 && col < D)) {
      float4 d;
      d.x = rs * (dn[k].x - s1 - xh[k].x * s2);
      d.y = rs * (dn[k].y - s1 - xh[k].y * s2);
      d.z = rs * (dn[k].z - s1 - xh[k].z * s2);
      d.w = rs * (dn[k].w - s1 - xh[k].w * s2);
      float4* p = reinterpret_cast<float4*>(dx_out + static_cast<long long>(row) * D + col);
      if (mode == 0) {
        const float4 o = *p;
        *p = make_float4(o.x + d.x, o.y + d.y, o.z + d.z, o.w + d.w);
      } else {
        const float4 hlf = make_float4(0.5f * d.x, 0.5f * d.y, 0.5f * d.z, 0.5f * d.w);
        *p = hlf;
        *reinterpret_cast<float4*>(dx2_out + static_cast<long long>(row) * D + col) = hlf;
      }
    }
  }
}

// dw[col] += sum_rows dn*xhat ; db[col] += sum_rows dn.  thread = 4 columns, block = one row slice; 4 rows in flight.
__global__ void __launch_bounds__(256)
ln_shift_bwd_dwdb_kernel(const __nv_bfloat16* __restrict__ dh, const float* __restrict__ x,
                         const float* __restrict__ x2, const float* __restrict__ mean, const float* __restrict__ rstd,
                         float* __restrict__ dw, float* __restrict__ db, int M, int D, ShiftGeom sg) {
  const int cols_per_block = blockDim.x * 4;
  const int stripes = D / cols_per_block;
  const int col = (blockIdx.x % stripes) * cols_per_block + threadIdx.x * 4;
  const int rsplit = blockIdx.x / stripes, nsplit = gridDim.x / stripes;
  const int rows_per = (M + nsplit - 1) / nsplit;
  const int r0 = rsplit * rows_per, r1 = min(M, r0 + rows_per);
  float4 aw = make_float4(0.f, 0.f, 0.f, 0.f), ab = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int base = r0; base < r1; base += 4) {
    uint2 u[4];
    float4 v[4];
    float mu[4], rs[4];
    bool ok[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int row = base + j;
      ok[j] = false;
      u[j] = make_uint2(0u, 0u);
      v[j] = make_float4(0.f, 0.f, 0.f, 0.f);
      mu[j] = 0.f; rs[j] = 0.f;
      if (row < r1) {
        const int bidx = row / sg.seq, t = row - bidx * sg.seq;
        const int dt = shift_dst(sg, t, col, D);
        if (dt >= 0) {
          ok[j] = true;
          u[j] = *reinterpret_cast<const uint2*>(dh + permuted_row(sg, bidx, dt) * D + col);
          v[j] = *reinterpret_cast<const float4*>(x + static_cast<long long>(row) * D + col);
          if (x2) {
            const float4 u2 = *reinterpret_cast<const float4*>(x2 + static_cast<long long>(row) * D + col);
            v[j] = make_float4(0.5f * (v[j].x + u2.x), 0.5f * (v[j].y + u2.y), 0.5f * (v[j].z + u2.z), 0.5f * (v[j].w + u2.w));
          }
          mu[j] = mean[row];
          rs[j] = rstd[row];
        }
      }
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (ok[j]) {
        const float2 a = unpack_bf16x2(u[j].x), c = unpack_bf16x2(u[j].y);
        aw.x += a.x * (v[j].x - mu[j]) * rs[j]; aw.y += a.y * (v[j].y - mu[j]) * rs[j];
        aw.z += c.x * (v[j].z - mu[j]) * rs[j]; aw.w += c.y * (v[j].w - mu[j]) * rs[j];
        ab.x += a.x; ab.y += a.y; ab.z += c.x; ab.w += c.y;
      }
    }
  }
  atomicAdd(dw + col, aw.x); atomicAdd(dw + col + 1, aw.y); atomicAdd(dw + col + 2, aw.z); atomicAdd(dw + col + 3, aw.w);
  atomicAdd(db + col, ab.x); atomicAdd(db + col + 1, ab.y); atomicAdd(db + col + 2, ab.z); atomicAdd(db + col + 3, ab.w);
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
  for (int row = rsplit; row < M; row += nsplit) {
    const long long off = static_cast<long long>(row) * D + col;
    const float4 g = *reinterpret_cast<const float4*>(dy + off);
    const uint2 u = *reinterpret_cast<const uint2*>(o + off);
    const float2 o01 = unpack_bf16x2(u.x), o23 = unpack_bf16x2(u.y);
    as.x += g.x * o01.x; as.y += g.y * o01.y; as.z += g.z * o23.x; as.w += g.w * o23.y;
    const float4 d = make_float4(g.x * sc.x, g.y * sc.y, g.z * sc.z, g.w * sc.w);
    ab.x += d.x; ab.y += d.y; ab.z += d.z; ab.w += d.w;
    uint2 w;
    w.x = pack_bf16x2(d.x, d.y);
    w.y = pack_bf16x2(d.z, d.w);
    *reinterpret_cast<uint2*>(d_o + off) = w;
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
  for (int row = rsplit; row < M; row += nsplit) {
    const __nv_bfloat16* p = x + static_cast<long long>(row) * ld + col;
    if (full) {
      const uint4 u = *reinterpret_cast<const uint4*>(p);
      float2 f;
      f = unpack_bf16x2(u.x); acc[0] += f.x; acc[1] += f.y;
      f = unpack_bf16x2(u.y); acc[2] += f.x; acc[3] += f.y;
      f = unpack_bf16x2(u.z); acc[4] += f.x; acc[5] += f.y;
      f = unpack_bf16x2(u.w); acc[6] += f.x; acc[7] += f.y;
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
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int threads = 256, rows_per_block = threads / 32;
  const int blocks = (M + rows_per_block - 1) / rows_per_block;
  ln_stats_kernel<<<blocks, threads, 0, st>>>(x, x2, mean, rstd, M, D, eps);
  ShiftGeom sg{seq, T, F, shift, row_split, M / seq};
  ln_shift_apply_kernel<<<blocks, threads, 0, st>>>(x, x2, mean, rstd, w, b, reinterpret_cast<__nv_bfloat16*>(out), M,
                                                    D, sg);
  return err();
}

extern "C" int b200_ln_shift_bwd(const void* dh, const float* x, const float* x2, const float* mean,
                                 const float* rstd, const float* w, float* dx, float* dx2, float* dw, float* db, int M,
                                 int D, int seq, int T, int F, int shift, int row_split, int mode, void* stream) {
  if (D % 128 || D > 128 * MAXV) return -41;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  ShiftGeom sg{seq, T, F, shift, row_split, M / seq};
  if (D % 512 && D != 256) return -41;  // 4 warps per row: D in {256, 512, 1024}
  const int threads = 256;
  const int blocks = (M + 1) / 2;
  ln_shift_bwd_dx_kernel<<<blocks, threads, 0, st>>>(reinterpret_cast<const __nv_bfloat16*>(dh), x, x2, mean, rstd, w,
                                                     dx, dx2, M, D, sg, mode);
  int cthreads = 256;
  if (D < cthreads * 4) cthreads = D / 4;
  const int stripes = D / (cthreads * 4);
  int nsplit = (148 * 8) / stripes;
  if (nsplit > M) nsplit = M;
  ln_shift_bwd_dwdb_kernel<<<stripes * nsplit, cthreads, 0, st>>>(reinterpret_cast<const __nv_bfloat16*>(dh), x, x2,
                                                                  mean, rstd, dw, db, M, D, sg);
  return err();
}

extern "C" int b200_resid_bwd_prep(const float* dy, const void* o, const float* scale, void* d_o, float* dscale,
                                   float* dbias, int M, int D, void* stream) {
  int threads = 256;
  if (D % 4) return -42;
  if (D < threads * 4) threads = D / 4;
  if (D % (threads * 4)) return -42;
  const int stripes = D / (threads * 4);
  int nsplit = (148 * 8) / stripes;
  if (nsplit > M) nsplit = M;
  if (nsplit < 1) nsplit = 1;
  resid_bwd_prep_kernel<<<stripes * nsplit, threads, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      dy, reinterpret_cast<const __nv_bfloat16*>(o), scale, reinterpret_cast<__nv_bfloat16*>(d_o), dscale, dbias, M, D);
  return err();
}

extern "C" int b200_colsum_bf16(const void* x, long long ld, float* out, int M, int N, float alpha, void* stream) {
  const int stripes = (N + 2047) / 2048;
  int nsplit = (148 * 8) / stripes;
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
