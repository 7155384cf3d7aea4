// Host-side numeric kernels (C++/OpenMP): the CPU twins of the device code, used by the CPU plumbing configuration,
// by the inter-box TCP data plane and as oracles.
//   * block-wise 8-bit quantisation with a 256-entry code (replaces bitsandbytes' cquantize_blockwise_cpu_fp32 /
//     cdequantize_blockwise_cpu_fp32 that reference lib/training/lamb_8bit.py:181-192 reaches)
//   * fp16 and uniform-8-bit averaging codecs + weighted accumulation (hivemind compression / TensorPartReducer,
//     SURVEY H7, H8, K18, K19; Appendix D.2)
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>

#include "b200_host_api.h"

static inline int nearest_code(const float* code, float x) {
  int lo = 0, hi = 255;
  for (int it = 0; it < 8; ++it) {
    const int mid = (lo + hi + 1) >> 1;
    if (code[mid] <= x) lo = mid; else hi = mid - 1;
  }
  int idx = lo;
  if (idx < 255) {
    const float dl = std::fabs(x - code[idx]), dr = std::fabs(x - code[idx + 1]);
    if (dr < dl) idx += 1;
  }
  return idx;
}

extern "C" void b200h_quantize_blockwise(const float* code, const float* x, float* absmax, uint8_t* out, long long n,
                                         int blocksize) {
  const long long nb = (n + blocksize - 1) / blocksize;
#pragma omp parallel for schedule(static)
  for (long long b = 0; b < nb; ++b) {
    const long long s = b * blocksize, e = std::min<long long>(n, s + blocksize);
    float mx = 0.f;
    for (long long i = s; i < e; ++i) mx = std::max(mx, std::fabs(x[i]));
    absmax[b] = mx;
    const float inv = mx > 0.f ? 1.0f / mx : 0.f;
    for (long long i = s; i < e; ++i) out[i] = static_cast<uint8_t>(nearest_code(code, x[i] * inv));
  }
}

extern "C" void b200h_dequantize_blockwise(const float* code, const uint8_t* q, const float* absmax, float* out,
                                           long long n, int blocksize) {
  const long long nb = (n + blocksize - 1) / blocksize;
#pragma omp parallel for schedule(static)
  for (long long b = 0; b < nb; ++b) {
    const long long s = b * blocksize, e = std::min<long long>(n, s + blocksize);
    const float a = absmax[b];
    for (long long i = s; i < e; ++i) out[i] = code[q[i]] * a;
  }
}

// ---- uniform 8-bit (hivemind Uniform8BitQuantization): mean / sigma -> 256 uniform buckets over 6 sigma, the
// codebook entry of a bucket is the mean of the ORIGINAL values that fell into it ----
extern "C" void b200h_uniform8_quantize(const float* x, long long n, uint8_t* q, float* lookup) {
  double sum = 0.0;
#pragma omp parallel for reduction(+ : sum)
  for (long long i = 0; i < n; ++i) sum += x[i];
  const float mu = static_cast<float>(sum / std::max<long long>(n, 1));
  double ss = 0.0;
#pragma omp parallel for reduction(+ : ss)
  for (long long i = 0; i < n; ++i) { const double c = x[i] - mu; ss += c * c; }
  const float sigma = static_cast<float>(std::sqrt(ss) / std::sqrt(static_cast<double>(std::max<long long>(n - 1, 1))));
  float scale = 6.0f * sigma / 256.0f;
  if (!(scale > 0.f)) scale = 1.0f;
  double sums[256];
  long long counts[256];
  for (int k = 0; k < 256; ++k) { sums[k] = 0.0; counts[k] = 0; }
#pragma omp parallel
  {
    double ls[256];
    long long lc[256];
    for (int k = 0; k < 256; ++k) { ls[k] = 0.0; lc[k] = 0; }
#pragma omp for schedule(static) nowait
    for (long long i = 0; i < n; ++i) {
      // round half to even, like torch.round
      const float r = std::nearbyint((x[i] - mu) / scale) + 128.0f;
      const int k = static_cast<int>(std::min(255.0f, std::max(0.0f, r)));
      q[i] = static_cast<uint8_t>(k);
      ls[k] += x[i];
      lc[k] += 1;
    }
#pragma omp critical
    for (int k = 0; k < 256; ++k) { sums[k] += ls[k]; counts[k] += lc[k]; }
  }
  for (int k = 0; k < 256; ++k) lookup[k] = static_cast<float>(sums[k] / static_cast<double>(std::max<long long>(counts[k], 1)));
}

extern "C" void b200h_uniform8_dequantize(const uint8_t* q, const float* lookup, float* out, long long n) {
#pragma omp parallel for schedule(static)
  for (long long i = 0; i < n; ++i) out[i] = lookup[q[i]];
}

// acc += w * x   (owner-side accumulation of one tensor part)
extern "C" void b200h_weighted_add(float* acc, const float* x, float w, long long n) {
#pragma omp parallel for simd schedule(static)
  for (long long i = 0; i < n; ++i) acc[i] += w * x[i];
}

// out = acc / denom - mine   (the delta sent back to a contributor)
extern "C" void b200h_delta(const float* acc, float inv_denom, const float* mine, float* out, long long n) {
#pragma omp parallel for simd schedule(static)
  for (long long i = 0; i < n; ++i) out[i] = acc[i] * inv_denom - mine[i];
}
