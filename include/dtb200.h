/* dtb200 — C ABI of the B200-native Decision-Transformer hot path (libdtb200.so).
 *
 * The reference (hopl1t/multimodal-transformer-rl) is pure Python over PyTorch library ops and has
 * no FFI of its own.  This header is the boundary a maintainer binds with
 * ctypes; every entry point names the reference-side computation it replaces.
 * "HF:" line numbers refer to transformers/models/decision_transformer/modeling_decision_transformer.py
 * (the in-container restatement of the un-vendored kzl decision-transformer submodule),
 * "REF:" to files under the reference tree.
 *
 * Conventions
 *   - every pointer is a raw CUDA device pointer owned by the caller (PyTorch); nothing is retained
 *   - `stream` is a cudaStream_t passed as void*; all functions enqueue and return (graph-capturable)
 *   - return value: 0 = OK, negative = error (dtb_last_error() gives the message); never aborts
 *   - bf16 buffers are passed as void*; fp32 as float*; int64 as long long*
 */
#ifndef DTB200_H
#define DTB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DTB_ABI_VERSION 1

#define DTB_OK 0
#define DTB_ERR_BAD_ARG (-1)
#define DTB_ERR_UNSUPPORTED (-2)
#define DTB_ERR_CUDA (-3)
#define DTB_ERR_ARCH (-4)
#define DTB_ERR_WORKSPACE (-5)

int dtb_abi_version(void);
const char* dtb_last_error(void);
/* 0 iff the current CUDA device is sm_100 (no fallback path exists on other devices). */
int dtb_check_device(void);

/* ---- tensor-core GEMM (tcgen05 + TMEM + TMA) --------------------------------------------------
 * D[M,N] = epilogue(A x B^T-style product), bf16 operands, fp32 accumulation.
 * Replaces torch addmm behind Conv1D / nn.Linear — HF:179 (c_attn), HF:217 (c_proj), HF:234-236 (MLP),
 * REF src/agents.py:40-42 (tower FC) — and their autograd dgrad / wgrad.
 *   a_mn_major = 0: A stored [M][K] (lda elements);  1: A stored [K][M]
 *   b_mn_major = 0: B stored [N][K] (ldb elements);  1: B stored [K][N]
 *   act : 0 none, 1 relu, 2 gelu_new.   dact: multiply by act'(aux) (1 relu: aux = saved output; 2 gelu: aux = pre-activation)
 *   outputs: any of out_bf16 / out_f32 / out_preact may be NULL.
 *   split_k > 1 needs `partial` (fp32 [split_k][M][N]); the call then also reduces into out_f32
 *   (accumulate != 0: out_f32 += result).
 */
typedef struct {
  const float* bias;
  const float* residual;
  int ldr;
  const void* aux;
  int ldaux;
  int act;
  int dact;
  void* out_bf16;
  int ldo_b;
  float* out_f32;
  int ldo_f;
  void* out_preact;
  float* partial;
  float alpha;
  int accumulate;
} dtb_gemm_epilogue;

int dtb_gemm_bf16(void* stream, const void* A, int lda, int a_mn_major, const void* B, int ldb, int b_mn_major, int M, int N, int K,
                  const dtb_gemm_epilogue* ep, int split_k, int bn_override);

/* ---- LayerNorm over the fp32 residual stream (HF:272 ln_1, HF:303 ln_2, HF:450 ln_f, HF:628 embed_ln) -------------
 * fwd: x fp32 [M,H] -> y bf16 [M,H]; mean/rstd fp32 [M] saved for backward (may be NULL).
 * bwd: dx = dres + LN'(dy): dx_f32 and/or dx_bf16 outputs; dgamma/dbeta fp32 [H] overwritten.
 *      scratch: fp32, dtb_layernorm_bwd_scratch_floats(H) elements. */
int dtb_layernorm_fwd(void* stream, const float* x, const float* gamma, const float* beta, int M, int H, float eps, void* y_bf16,
                      float* mean, float* rstd);
size_t dtb_layernorm_bwd_scratch_floats(int H);
int dtb_layernorm_bwd(void* stream, const void* dy_bf16, const float* x, const float* mean, const float* rstd, const float* gamma,
                      const float* dres, int M, int H, float* dx_f32, void* dx_bf16, float* dgamma, float* dbeta, float* scratch);

/* ---- fused causal attention with the [B,K] padding mask expanded by index (HF:47-65, 408-414, 631-635) -------------
 * qkv bf16 [B,T,3H] (q|k|v, head h at h*d), mask fp32 [B,T/3] or NULL; allowed(i,j) = j<=i && mask[b,j/3]>0.
 * fwd writes out bf16 [B,T,H] and lse fp32 [B,nh,T] (base-2 log-sum-exp of the scaled scores).
 * bwd writes dqkv bf16 [B,T,3H]; delta fp32 [B,nh,T] is scratch. head_dim d in {32,64,128}. */
int dtb_attention_fwd(void* stream, const void* qkv, const float* mask, int B, int T, int n_head, int head_dim, void* out, float* lse);
int dtb_attention_bwd(void* stream, const void* qkv, const float* mask, const void* out, const void* dout, const float* lse, int B,
                      int T, int n_head, int head_dim, void* dqkv, float* delta);

/* ---- weight staging + bias gradients -----------------------------------------------------------------------------
 * cast_transpose: fp32 [R,C] -> bf16 [R,C] (dst, may be NULL) and bf16 [C,R] (dst_t, may be NULL).
 * colsum: out[N] (+)= column sums of bf16 x [M,N] (row stride ld); scratch fp32 dtb_colsum_scratch_floats(N). */
int dtb_cast_transpose(void* stream, const float* src, int R, int C, void* dst, void* dst_t);
size_t dtb_colsum_scratch_floats(int N);
int dtb_colsum_bf16(void* stream, const void* x, int M, int N, int ld, float* out, int accumulate, float* scratch);

#ifdef __cplusplus
}
#endif
#endif /* DTB200_H */
