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
  const void* rot;     /* float2[seq_len][32] = (cos, sin) per interleaved pair (ROTARY) */
  int seq_len; int q_cols; float q_scale;
  const void* aux; long long ldaux; /* bf16 [M, 2*n_half] (GEGLU_BWD) */
  int n_half;          /* GEGLU: N counts value columns only; B has 2*n_half rows */
  float alpha;
  int splits;          /* 0 = auto (only ATOMIC_F32 may split) */
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
 * Scratch: delta fp32 [B,H,seq]; kv_acc fp32 [B*seq, 2*H*64] which must be all-zero on entry and is returned
 * all-zero (gradients of keys shared between CTAs are reduced there with red.add before conversion). */
int b200_attn_bwd(const void* qkv, const void* out, const void* d_out, const float* lse, float* delta, void* dqkv,
                  float* kv_acc, const void* rot, float q_scale, int B, int H, int seq, int T, int F, int mode,
                  int conv_kernel, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200_API_H_ */
