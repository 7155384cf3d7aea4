// Shared definitions of the sparse-attention kernels (forward: attn_fwd_sm100.cu, backward: attn_bwd_sm100.cu).
//
// The three DALL-E sparsity patterns (dalle-pytorch SparseAxialCausalAttention /
// SparseConvCausalAttention as configured at reference task.py:63-64) are evaluated as dense 128-query blocks on
// the tensor cores with an in-register mask:
//   TEXT : 128 text queries x text keys (causal)
//   ROW  : 128 image queries (consecutive tokens = 128/F image rows)  x (all text keys + the same 128 tokens)
//   COL  : 128 image queries enumerated column-major (128/F columns)  x (all text keys + the same 128 tokens)
//   CONV : 128 image queries (consecutive tokens) x (all text keys + tokens [q0-2F, q0+128))
// q, k and v arrive already rotated (and q pre-scaled) from the QKV GEMM epilogue.
#pragma once
#include <cstdlib>
#include "sm100_ptx.cuh"

namespace b200 {

enum { ATT_TEXT = 0, ATT_ROW = 1, ATT_COL = 2, ATT_CONV = 3 };

struct AttnGeom {
  int B, H;        // batch, heads (dim_head is fixed at 64)
  int seq;         // tokens per sample in the activations (1280)
  int T;           // text length incl. <bos> (257)
  int T_pad;       // T rounded up to 16
  int F;           // image feature-map side (32)
  int P;           // F*F image positions (the last one is padding: it has no token)
  int n_text_blocks, n_img_blocks;
  int conv_half;   // kernel_size / 2 (2)
};

// token index (0..seq) of query i of image block `blk`
template <int MODE>
__device__ __forceinline__ int img_query_pos(const AttnGeom& g, int blk, int i) {
  const int pc = blk * 128 + i;
  if (MODE == ATT_COL) {
    const int c = pc / g.F, r = pc - c * g.F;
    return r * g.F + c;
  }
  return pc;
}

// 16-byte chunk address inside a [rows x 64 bf16] 128B-swizzled tile (row pitch 128 B, base 1024-aligned)
__device__ __forceinline__ uint32_t sw128_offset(int row, int chunk16) {
  return static_cast<uint32_t>(row) * 128u + (static_cast<uint32_t>(chunk16 ^ (row & 7)) << 4);
}

constexpr float kLog2e = 1.4426950408889634f;

// single MUFU.EX2 (exp2f() adds range handling that the softmax does not need)
__device__ __forceinline__ float fast_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}


// Host side: run two independent launches of one attention call CONCURRENTLY (the small text-block kernel beside the
// image-block kernel): the caller's stream forks into a per-device side stream and joins it again, so stream order
// towards earlier and later work is unchanged and the pattern is capturable in a CUDA graph.  The text kernels are
// a few per cent of the work but took a quarter of the time when run alone (three partly filled waves); next to
// the image kernel their CTAs fill its gaps (-8 us forward, -11 us backward per layer).
// DALLE_B200_ATTN_FORK=0 keeps the launches in one stream (2 = fork the forward only, 3 = the backward only).
// (Run-to-run differences of ~1e-3 in some gradients seen while this was introduced were traced to the
// convolution-like layer's backward, whose shared-key gradients meet through fp32 red.add before the bf16 store: the
// order of the atomics flips single bf16 ulps, with or without the fork - tools/exp/flake_attn_bwd.py.)
struct AttnFork {
  cudaStream_t side = nullptr;
  cudaEvent_t fork = nullptr, join = nullptr;
  bool ok = false;
};
inline AttnFork* attn_fork_state(int which = 1) {   // which: 2 forward, 3 backward (DALLE_B200_ATTN_FORK=2|3 selects one)
  static thread_local AttnFork per_dev[64];
  static const int mode = [] { const char* e = getenv("DALLE_B200_ATTN_FORK"); return e ? atoi(e) : 1; }();
  if (mode == 0 || (mode > 1 && mode != which)) return nullptr;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
  AttnFork& f = per_dev[dev];
  if (!f.ok) {
    if (cudaStreamCreateWithFlags(&f.side, cudaStreamNonBlocking) != cudaSuccess) return nullptr;
    if (cudaEventCreateWithFlags(&f.fork, cudaEventDisableTiming) != cudaSuccess) return nullptr;
    if (cudaEventCreateWithFlags(&f.join, cudaEventDisableTiming) != cudaSuccess) return nullptr;
    f.ok = true;
  }
  return &f;
}
// side = stream for the forked launch (== st when forking is off); call attn_join(f, st) after both launches
inline cudaStream_t attn_fork(AttnFork* f, cudaStream_t st) {
  if (!f) return st;
  if (cudaEventRecord(f->fork, st) != cudaSuccess || cudaStreamWaitEvent(f->side, f->fork, 0) != cudaSuccess) return st;
  return f->side;
}
inline void attn_join(AttnFork* f, cudaStream_t side, cudaStream_t st) {
  if (!f || side == st) return;
  cudaEventRecord(f->join, side);
  cudaStreamWaitEvent(st, f->join, 0);

}

}  // namespace b200
