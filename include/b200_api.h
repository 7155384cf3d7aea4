/* C ABI of libdalle_b200.so - the sm_100a kernel library behind dalle_b200.ops.
 *
 * Every entry point takes raw device pointers plus a CUDA stream handle (cudaStream_t passed as void*),
 * launches asynchronously on that stream and returns 0 on success or a negative error code.
 * Unless a comment says otherwise an entry point allocates nothing: the caller owns inputs, outputs and
 * any scratch. Matrices are row-major; "ld" is the leading dimension in ELEMENTS.
 */
#ifndef B200_API_H_
#define B200_API_H_

#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------------------------------------
 * bf16 tensor-core GEMM (tcgen05 + TMEM + TMA).  Replaces the cuBLAS calls under the reference's
 * nn.Linear layers (dalle-pytorch Attention.to_qkv/to_out, FeedForward.net, DALLE.to_logits; configured at
 * reference task.py:62-83) and their autograd twins.
 * ---------------------------------------------------------------------------------------------- */
enum {
  B200_GEMM_NT = 0, /* C[M,N] = A[M,K] * B[N,K]^T   forward  (both operands K-major)            */
  B200_GEMM_NN = 1, /* C[M,N] = A[M,K] * B[K,N]     dgrad    (B read MN-major)                  */
  B200_GEMM_TN = 2  /* C[M,N] = A[K,M]^T * B[K,N]   wgrad    (both MN-major, split-K + red.add) */
};

enum {
  B200_EPI_BF16 = 0,       /* out(bf16) = alpha * (acc + bias)                                             */
  B200_EPI_F32 = 1,        /* out(f32)  = alpha * acc                                                      */
  B200_EPI_ATOMIC_F32 = 2, /* out(f32) += alpha * acc   (red.global.add; the only epilogue allowing splits) */
  B200_EPI_ROTARY = 3,     /* out(bf16) = q_scale? * rotary(acc + bias): QKV projection, SURVEY K4+K5       */
  B200_EPI_RESID = 4,      /* out(f32) = res + alpha*scale[col]*(acc+bias[col]); out2(bf16, opt) = acc+bias */
  B200_EPI_GEGLU = 5,      /* out(bf16)[.,j] = (acc[j]+b[j]) * gelu(acc[n_half+j]+b[n_half+j]); out2 = pre */
  B200_EPI_GEGLU_BWD = 6   /* acc = d(out); aux = pre-activations; out(bf16) = [d value | d gate]           */
};

typedef struct B200Gemm {
  int mode, epi;
  int M, N, K;
  const void* A; long long lda; /* bf16 */
  const void* B; long long ldb; /* bf16 */
  void* out;  long long ldo;
  void* out2; long long ldo2;
  const float* bias;   /* [N] (GEGLU: [2*n_half]) or NULL */
  const float* scale;  /* [N] LayerScale (RESID)           */
  const float* res; long long ldres; /* fp32 residual (RESID) */
  const void* rot;     /* float2[32][rot_len] = (cos, sin) per interleaved pair, pair-major (ROTARY) */
  int seq_len; int rot_len; int q_cols; float q_scale;
  const void* pos_ptr; /* optional device int: one position for all rows (decode) */
  const void* aux; long long ldaux; /* bf16 [M, 2*n_half] (GEGLU_BWD) */
  int n_half;          /* GEGLU: N counts value columns only; B has 2*n_half rows */
  float alpha;
  int splits;          /* 0 = auto (only ATOMIC_F32 may split) */
  float* a_colsum;     /* TN + ATOMIC_F32 only, or NULL: a_colsum[m] += sum_k A[k, m] (the bias gradient that goes
                          with a weight gradient), summed from the A tiles while they sit in shared memory */
} B200Gemm;

int b200_gemm_bf16(const B200Gemm* desc, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Sparse DALL-E attention (tcgen05).  qkv: bf16 [B*seq (+1 padding row), 3*H*64] = [q | k | v], already rotated and
 * q pre-scaled (B200_EPI_ROTARY).  mode: 1 = axial_row, 2 = axial_col, 3 = conv_like (text->text causal attention
 * is always included).  out: bf16 [B*seq, H*64]; lse: fp32 [B, H, seq] (log2 domain, consumed by the backward).
 * Replaces dalle-pytorch SparseAxialCausalAttention / SparseConvCausalAttention (reference task.py:63-64).
 * ---------------------------------------------------------------------------------------------- */
int b200_attn_fwd(const void* qkv, void* out, float* lse, int B, int H, int seq, int T, int F, int mode,
                  int conv_kernel, void* stream);

/* Backward of b200_attn_fwd.  d_out: bf16 [B*seq (+1 padding row), H*64].  Writes dqkv bf16 [B*seq, 3*H*64] =
 * gradient w.r.t. the RAW projection output (inverse rotary and the q scale are applied inside).
 * Scratch: delta (unused since delta = rowsum(dO*O) is formed inside the kernel; kept for ABI stability); kv_acc fp32 [B*seq, 2*H*64] which must be all-zero on entry and is returned
 * all-zero (gradients of keys shared between CTAs are reduced there with red.add before conversion). */
int b200_attn_bwd(const void* qkv, const void* out, const void* d_out, const float* lse, float* delta, void* dqkv,
                  float* kv_acc, const void* rot /* [seq+1][32] */, const void* rot_t /* [32][seq+1] */, float q_scale, int B, int H, int seq, int T, int F, int mode,
                  int conv_kernel, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Memory-bound kernels (elementwise.cu).  fp32 residual streams, bf16 GEMM operands.
 * ---------------------------------------------------------------------------------------------- */
/* row_split > 0 stores the bf16 side as [first row_split tokens of every sample | the remaining tokens] so that the
 * text / image vocabulary segments of the logits GEMM see contiguous rows.
 * out(bf16)[t] = token_shift(LayerNorm(x (or 0.5*(x+x2)))) ; also writes per-row mean / rstd (fp32 [M]).
 * Replaces PreNorm + PreShiftToken of dalle-pytorch. shift=0 gives a plain LayerNorm. */
int b200_ln_shift_fwd(const float* x, const float* x2, const float* w, const float* b, void* out, float* mean,
                      float* rstd, int M, int D, int seq, int T, int F, int shift, int row_split, float eps,
                      void* stream);
/* Backward of the above. mode 0: dx += grad ; mode 1: dx = dx2 = 0.5*grad (final norm over the stream mean).
 * dw/db (fp32 [D]) are accumulated into. */
int b200_ln_shift_bwd(const void* dh, const float* x, const float* x2, const float* mean, const float* rstd,
                      const float* w, float* dx, float* dx2, float* dw, float* db, int M, int D, int seq, int T, int F,
                      int shift, int row_split, int mode, void* stream);
/* b200_ln_shift_bwd (mode 0) that also runs the LayerScale/residual prologue of the block the backward walk visits
 * next on the stream gradient it has just completed: d_o_next = bf16(dx * scale_next),
 * dscale_next += colsum(dx * o_next), dbias_next += colsum(dx * scale_next)  (what b200_resid_bwd_prep would do in a
 * second pass over dx).  Needs o_next kept from the forward, i.e. a layer that stores its activations. */
int b200_ln_shift_bwd_prep(const void* dh, const float* x, const float* mean, const float* rstd, const float* w,
                           float* dx, float* dw, float* db, const void* o_next, const float* scale_next,
                           void* d_o_next, float* dscale_next, float* dbias_next, int M, int D, int seq, int T, int F,
                           int shift, int row_split, void* stream);
/* LayerScale/residual backward prep: d_o(bf16) = dy*scale ; dscale += colsum(dy*o) ; dbias += colsum(dy*scale). */
int b200_resid_bwd_prep(const float* dy, const void* o, const float* scale, void* d_o, float* dscale, float* dbias,
                        int M, int D, void* stream);
/* out[col] += alpha * sum over rows of x (bf16 [M, N], leading dimension ld). */
int b200_colsum_bf16(const void* x, long long ld, float* out, int M, int N, float alpha, void* stream);
/* Tied embedding (dalle-pytorch SharedEmbedding): gather rows of W (fp32) into one or two fp32 streams /
 * scatter-add (d1 + d2) back into dW. */
int b200_embed_fwd(const long long* ids, const float* W, float* x1, float* x2, int M, int D, void* stream);
int b200_embed_bwd(const long long* ids, const float* d1, const float* d2, float* dW, int M, int D, void* stream);
/* Cross-entropy over one vocabulary segment, in place: logits(bf16 [rows, ld]) -> dlogits ; loss_acc += weighted
 * NLL.  The logits mask of the reference (text positions predict text ids only, image positions image ids only,
 * SURVEY M1) is realised by running the segment GEMMs separately, so only n_classes columns exist. */
int b200_ce_fwd_bwd(void* logits, long long ld, const long long* labels, int rows, int n_classes, float row_weight,
                    const float* grad_scale /* device scalar or NULL = 1 */, float* loss_acc, int write_grad,
                    void* stream);
int b200_cast_f32_bf16(const float* src, void* dst, long long n, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Fused LAMB with block-wise 8-bit moments over one flat fp32 parameter buffer (lamb.cu).
 * Replaces reference lib/training/lamb_8bit.py (CPULAMB8Bit.step) + bitsandbytes quantize/dequantize_blockwise.
 * blocks: one 24-byte record {int tensor, int valid, long long state_off, int quantized, int absmax_idx} per
 * 4096-element block.  tensor_sums [2*n_tensors], trust_out [3*n_tensors] (weight_norm, step_norm, trust),
 * stats [2] = {sum g^2, non-finite flag}: all device scratch owned by the caller.
 * If a non-finite gradient is seen the parameters are left untouched and stats[1] = 1.
 * ---------------------------------------------------------------------------------------------- */
int b200_lamb_step(float* grad, float* param, void* param_bf16, long long n_padded, void* q1, void* q2, float* f1,
                   float* f2, float* absmax1, float* absmax2, const float* code1, const float* code2,
                   const void* blocks, int n_blocks, const float* wd_per_tensor, float* tensor_sums, int n_tensors,
                   float* trust_out, float* stats, float beta1, float beta2, float eps, float max_grad_norm,
                   float clamp_value, float lr, float grad_scale, int zero_grad, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Fused intra-box step (group_avg.cu): gradient reduce-scatter over NVLink peer memory -> LAMB-8bit on this rank's
 * shard -> parameter all-gather with P2P stores, as ONE cooperative kernel per rank; no NCCL call on the path.
 * Replaces nn.DataParallel's reduce/broadcast inside a box (reference lib/training/hf_trainer.py:35-38 via HF
 * Trainer) and the CPU optimizer step after averaging (reference lib/training/lamb_8bit.py:83-158).
 * All *_ptrs[r] are rank r's copies of symmetric (peer-mapped) allocations; signal_ptrs[r] holds >= world uint32
 * slots that start at zero; xchg_ptrs[r] holds world * xstride floats; epoch must grow by 4 with every call.
 * Every rank of the box must launch the kernel for the same step (the ranks wait for each other in device code).
 * ---------------------------------------------------------------------------------------------- */
typedef struct B200Group {
  int rank, world;
  void* grad_ptrs[8]; void* param_ptrs[8]; void* shadow_ptrs[8]; void* payload_ptrs[8];
  void* signal_ptrs[8]; void* xchg_ptrs[8];
  int xstride, n_tensors, n_blocks, blk0, blk1;
  unsigned int epoch;
  int payload_mode;   /* 0: fp32 gradients read over NVLink; 1: bf16 payload (half the bytes); 2: absmax int8 per
                         4096-element block; 3: hivemind's codecs per chunk of 32 blocks (chunk_codec[]: 3 = fp16,
                         4 = uniform 8-bit with a 256-entry bucket-mean codebook - reference task.py:124-126) */
  int zero_grad;
  int reduce_only;    /* 1: only reduce; every rank ends with grad_scale * sum of all ranks' gradients */
  void* q1; void* q2; float* f1; float* f2; float* absmax1; float* absmax2;
  const float* code1; const float* code2; const void* blocks; const float* wd_per_tensor;
  float* tensor_sums; float* stats /* [4] */; float* trust_out;
  float beta1, beta2, eps, max_grad_norm, clamp_value, lr, grad_scale;
  const void* chunk_codec;   /* uint8 [ceil(n_blocks / 32)] on the device; payload_mode 3 only */
} B200Group;
int b200_group_lamb_step(const B200Group* desc, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Single-token decode kernels (decode.cu) for KV-cached image sampling; every kernel reads the current position
 * from device memory (pos_ptr) so a whole decode step can be captured in a CUDA graph.  Replaces the cached
 * forward + sampling loop of dalle-pytorch generate_images (reference inference/run_inference.py:87-91).
 * ---------------------------------------------------------------------------------------------- */
int b200_decode_embed(const long long* ids, int ids_stride, const float* W, float* x1, float* x2, const int* pos_ptr,
                      int B, int D, void* stream);
int b200_decode_ln_shift(const float* x, const float* x2, const float* w, const float* bias, void* cache, void* out,
                         const int* pos_ptr, int B, int D, int seq, int T, int F, float eps, int shift, void* stream);
int b200_decode_attn(const void* qkv, void* kcache, void* vcache, void* out, const int* pos_ptr, int B, int H, int seq,
                     int T, int F, int mode, int conv_kernel, void* stream);
int b200_decode_sample(const void* logits, int ld, int n, float temperature, int top_k, float top_p,
                       const unsigned long long* seed /* device */, long long* ids, int ids_stride, long long id_offset,
                       long long* codes,
                       int codes_stride, const int* pos_ptr, int T, int B, void* stream);
int b200_decode_advance(int* pos_ptr, void* stream);

/* Second-generation decode step (decode2.cu): skinny tcgen05 GEMMs with the weights on the M side and the batch
 * (<= 64 rows, buffers allocated with nb = 32 or 64 rows) on an N = 32 tile, K split over a thread-block cluster and reduced through
 * distributed shared memory; layer norm + token shift as the prologue of the QKV / FF1 GEMMs (row statistics left
 * behind by the producer of the residual stream); attention over a sparsity-aware cache of `slots` keys per layer
 * (text + one image row | transposed image | conv_half + 1 rows).  All launched with programmatic dependent launch.
 * kind: 0 QKV (LN prologue, rotary), 1 residual (bf16 operand via TMA), 2 GEGLU (LN prologue; W = [2N, K], bias [2N]),
 * 3 logits (bf16 operand via TMA). */
int b200_decode2_gemm(int kind, const void* W, int N, int K, const void* xin_bf16, const float* x,
                      const float* stats_in, const float* ln_w, const float* ln_b, void* ring, const int* pos_ptr, int T,
                      int F, float eps, const float* bias, const float* scale, float* xout, float* stats_out, void* out,
                      int ldo, const void* rot, int q_cols, float q_scale, int nb /* batch tile: 32 or 64 rows */,
                      void* stream);
int b200_decode2_attn(const void* qkv, void* kcache, void* vcache, void* out, const int* pos_ptr, int B, int H, int T,
                      int F, int mode, int conv_kernel, int slots, void* stream);
int b200_decode2_embed(const long long* ids, int ids_stride, const float* W, float* x1, float* x2, float* stats,
                       int n_sites, const int* pos_ptr, int B, int D, int nb, void* stream);
int b200_decode2_set_pdl(int on);

/* b200_lamb_step for callers that cannot provide scratch memory.  No counterpart in the reference: it exists for
 * embedders of this library that may not allocate device memory once they are running (all of it belongs to their
 * own pool, or an allocation inside the optimizer step could fail or stall the device).  The two work arrays of
 * b200_lamb_step (stats, tensor_sums) live in a fixed __device__ array inside the library instead
 * (B200_LAMB_NOALLOC_SCRATCH_FLOATS floats, about 33 KB, held by every GPU that loads the library until exit), so
 * at most B200_LAMB_NOALLOC_MAX_TENSORS parameter tensors are accepted; more returns
 * B200_ERR_NOALLOC_TOO_MANY_TENSORS with a message on stderr and launches nothing.
 *   - inputs and outputs are the caller's as in b200_lamb_step; trust_out [3*n_tensors] and stats_out [2] are
 *     optional outputs (NULL to skip);
 *   - calls on one device share the array: it is cleared where needed at the start of each call, and calls that can
 *     overlap on the GPU race with each other - order them on one stream or with events;
 *   - the first call on a device looks up the array's device address (cudaGetSymbolAddress); make that call before
 *     any stream capture begins;
 *   - calls no cudaMalloc / cudaFree / allocating library routine. */
#define B200_LAMB_NOALLOC_MAX_TENSORS 4096
#define B200_LAMB_NOALLOC_SCRATCH_FLOATS (64 + 2 * B200_LAMB_NOALLOC_MAX_TENSORS)
#define B200_ERR_NOALLOC_TOO_MANY_TENSORS (-53)
#define B200_ERR_NOALLOC_SETUP (-52)
int b200_lamb_step_noalloc(float* grad, float* param, void* param_bf16, long long n_padded, void* q1, void* q2,
                           float* f1, float* f2, float* absmax1, float* absmax2, const float* code1,
                           const float* code2, const void* blocks, int n_blocks, const float* wd_per_tensor,
                           int n_tensors, float* trust_out, float* stats_out, float beta1, float beta2, float eps,
                           float max_grad_norm, float clamp_value, float lr, float grad_scale, int zero_grad,
                           void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200_API_H_ */
