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
/* Number of kernels this library has launched in this process (the bench's gpu_launches evidence). */
long long dtb_launch_count(void);
/* Live per-kernel-family timing for bench.py's roofline: when enabled (and the stream is not being captured) every launch of a
 * family is bracketed by CUDA events on its own stream.  family: 0 gemm (work = flops), 1 conv1 fwd (bytes), 2 conv1 wgrad (bytes),
 * 3 attention fwd (flops), 4 attention bwd (flops), 5 LN fwd (bytes), 6 LN bwd (bytes), 7 embed fwd (bytes), 8 colsum (bytes),
 * 9 im2col (bytes), 10 col2im (bytes).  report() synchronises the events and returns summed milliseconds, algorithmic work, launches. */
void dtb_profile_enable(int on);
/* machine balance (flop per byte: sustained bf16 peak / copy bandwidth of MEASURED_PEAKS.json) that splits GEMM launches into the
 * tensor-bound family 0 and the HBM-bound family 11 */
void dtb_profile_set_balance(double flop_per_byte);
/* Runtime switches between kernel forms (A/B measurements and tests inside one process); unknown names are an error.
 *   "splitk_inkernel" (default 0): split-K slices reduced by the last-arriving CTA inside the GEMM instead of a second launch
 *   "ln_fused"        (default 1): LayerNorm in the c_proj GEMM epilogue when hidden <= 256
 *   "conv_implicit"   (default 1): implicit-GEMM conv2 / conv3 (read when an encoder handle is created) */
int dtb_set_option(const char* name, int value);
int dtb_profile_report(int family, double* total_ms, double* work, long long* launches);
/* timeline of the last profiled step: number of recorded launches, and launch i's family + [start, end] in ms from the first one */
int dtb_profile_spans(void);
int dtb_profile_span(int i, int* family, double* start_ms, double* end_ms);

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
/* GEMM with LayerNorm fused behind the epilogue (replaces c_proj + residual add + the next nn.LayerNorm: HF:217-218 + :303, HF:236-237 +
 * :272 / :450): out_f32 (if given) receives the fp32 rows v = epilogue(acc); ln_out bf16 [M, ld_ln] = (v - mean) * rstd * gamma + beta;
 * mean / rstd fp32 [M] (may be NULL).  Needs N <= 256, N % 64 == 0 (one CTA owns whole rows), no split-K. */
int dtb_gemm_bf16_ln(void* stream, const void* A, int lda, int a_mn_major, const void* B, int ldb, int b_mn_major, int M, int N, int K,
                     const dtb_gemm_epilogue* ep, const float* ln_gamma, const float* ln_beta, float ln_eps, void* ln_out_bf16, int ld_ln,
                     float* ln_mean, float* ln_rstd);

/* ---- LayerNorm over the fp32 residual stream (HF:272 ln_1, HF:303 ln_2, HF:450 ln_f, HF:628 embed_ln) -------------
 * fwd: x fp32 [M,H] -> y bf16 [M,H]; mean/rstd fp32 [M] saved for backward (may be NULL).
 * bwd: dx = dres + LN'(dy): dx_f32 and/or dx_bf16 outputs; dgamma/dbeta fp32 [H] overwritten.
 *      scratch: fp32, dtb_layernorm_bwd_scratch_floats(H) elements. */
int dtb_layernorm_fwd(void* stream, const float* x, const float* gamma, const float* beta, int M, int H, float eps, void* y_bf16,
                      float* mean, float* rstd);
size_t dtb_layernorm_bwd_scratch_floats(int H);
int dtb_layernorm_bwd(void* stream, const void* dy_bf16, const float* x, const float* mean, const float* rstd, const float* gamma,
                      const float* dres, int M, int H, float* dx_f32, void* dx_bf16, float* dgamma, float* dbeta, float* scratch);
/* Same, plus dcolsum fp32 [H] = column sums of the dx values handed on as dx_bf16 (before bf16 rounding): the bias gradient of the
 * Linear / Conv1D whose output gradient dx is (HF:217 attn.c_proj, HF:236 mlp.c_proj), reduced inside the same launch. */
int dtb_layernorm_bwd_colsum(void* stream, const void* dy_bf16, const float* x, const float* mean, const float* rstd, const float* gamma,
                             const float* dres, int M, int H, float* dx_f32, void* dx_bf16, float* dgamma, float* dbeta, float* dcolsum,
                             float* scratch);

/* ---- fused causal attention with the [B,K] padding mask expanded by index (HF:47-65, 408-414, 631-635) -------------
 * qkv bf16 [B,T,3H] (q|k|v, head h at h*d), mask fp32 [B,T/3] or NULL; allowed(i,j) = j<=i && mask[b,j/3]>0.
 * fwd writes out bf16 [B,T,H] and lse fp32 [B,nh,T] (base-2 log-sum-exp of the scaled scores).
 * bwd writes dqkv bf16 [B,T,3H]; delta fp32 [B,nh,T] is scratch. head_dim d in {32,64,128}. */
int dtb_attention_fwd(void* stream, const void* qkv, const float* mask, int B, int T, int n_head, int head_dim, void* out, float* lse);
int dtb_attention_bwd(void* stream, const void* qkv, const float* mask, const void* out, const void* dout, const float* lse, int B,
                      int T, int n_head, int head_dim, void* dqkv, float* delta);
/* The same kernels with attn_dropout on the softmax weights (HF eager attention :60-62): keep bits of `site` drawn from the device
 * generator state (see dtb_dt_set_dropout); the backward re-draws the bits of the forward.  p = 0 or drop_state NULL = no dropout. */
int dtb_attention_fwd_dropout(void* stream, const void* qkv, const float* mask, int B, int T, int n_head, int head_dim, void* out,
                              float* lse, const void* drop_state, unsigned site, float p);
int dtb_attention_bwd_dropout(void* stream, const void* qkv, const float* mask, const void* out, const void* dout, const float* lse, int B,
                              int T, int n_head, int head_dim, void* dqkv, float* delta, const void* drop_state, unsigned site, float p);

/* ---- weight staging + bias gradients -----------------------------------------------------------------------------
 * cast_transpose: fp32 [R,C] -> bf16 [R,C] (dst, may be NULL) and bf16 [C,R] (dst_t, may be NULL).
 * colsum: out[N] (+)= column sums of bf16 x [M,N] (row stride ld; N, ld multiples of 8); scratch fp32 dtb_colsum_scratch_floats(N). */
int dtb_cast_transpose(void* stream, const float* src, int R, int C, void* dst, void* dst_t);
/* get_action history window (upstream DecisionTransformer.get_action): for each of nseg per-timestep tensors
 * (row_bytes[s] bytes per timestep, multiples of 4; srcs[s] holds the t most recent timesteps) write K rows to dsts[s]: K - t zero
 * rows, then the t rows; mask_out[K] = [0]*(K-t) + [1]*t (may be NULL).  One launch; bit-exact copies. */
int dtb_pad_left(void* stream, int nseg, const void* const* srcs, void* const* dsts, const int* row_bytes, int t, int K, float* mask_out);
size_t dtb_colsum_scratch_floats(int N);
int dtb_colsum_bf16(void* stream, const void* x, int M, int N, int ld, float* out, int accumulate, float* scratch);
/* Allocation-free variant of dtb_colsum_bf16.  NO counterpart in the reference: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same kernels,
 * same slab count and therefore bit-identical results; the slab partials live in a 256 KB file-scope `static __device__` array inside
 * the library (held by every GPU that loads libdtb200.so until exit) instead of caller scratch.  Limit: N <= DTB_COLSUM_STATIC_MAX_N
 * columns (larger N -> DTB_ERR_BAD_ARG with a message; use dtb_colsum_bf16).  Calls no cudaMalloc / cudaFree / allocating routine.
 * Every call on a device shares the one array: calls that overlap on the GPU (different streams without ordering) race; calls on
 * one stream, or ordered by events, are safe.  Nothing in the array has to start at zero (stage 1 overwrites what stage 2 reads). */
#define DTB_COLSUM_STATIC_MAX_N 1024
int dtb_colsum_bf16_static(void* stream, const void* x, int M, int N, int ld, float* out, int accumulate);

/* ---- fused token embedding (HF:611-628): timestep gather + RTG/state/action projections + interleave + embed_ln -----
 * x fp32 [B*K*3, H] in token order R_k, s_k, a_k.  Exactly one of states (fp32 [BK, sd], sd <= 64) / state_proj
 * (fp32 [BK, H], the tensor-core projection of a wide multimodal state, no bias) is given.  pre / mean / rstd / U are the
 * optional training saves (U = bf16 [BK, P8] packed [rtg | actions | states]). */
int dtb_embed_fwd(void* stream, const float* states, const float* state_proj, const float* actions, const float* rtg,
                  const long long* timesteps, const float* table, const float* wR, const float* bR, const float* Ws, const float* bs,
                  const float* Wa, const float* ba, const float* gamma, const float* beta, int BK, int sd, int ad, int H, int E,
                  float eps, float* x, float* pre, float* mean, float* rstd, void* U, int P8, int u_states);

/* ---- prediction heads + masked action MSE (HF:650-655; upstream SequenceTrainer.train_step loss) -------------------- */
int dtb_heads_fwd(void* stream, const void* y_bf16, const float* Wa, const float* ba, const float* Ws, const float* bs,
                  const float* Wr, const float* br, int BK, int H, int sd, int ad, int action_tanh, float* action_preds,
                  float* state_preds, float* return_preds);
size_t dtb_mse_scratch_floats(void);
/* loss_out[0] = mean((pred-target)^2) over rows with mask > 0, loss_out[1] = #valid rows; d_pred (nullable) = grad_scale * dloss/dpred */
int dtb_masked_mse(void* stream, const float* pred, const float* target, const float* mask, int BK, int ad, float grad_scale,
                   float* loss_out, float* d_pred, float* scratch);

/* ---- whole-model Decision Transformer (A1-A10) -----------------------------------------------------------------------
 * Mirrors DecisionTransformer(state_dim, act_dim, hidden_size, max_length, max_ep_len, action_tanh, n_layer, n_head,
 * n_inner, activation_function) — HF:510-533 / upstream decision_transformer.py.  Parameters live in ONE flat fp32
 * buffer owned by the caller (layout from dtb_dt_param_layout, each tensor padded to 8 elements), with a bf16 shadow
 * of the same layout refreshed by dtb_dt_stage_params after every optimizer step; gradients go to a flat fp32 buffer
 * of the same layout.  Activations saved for backward live in a caller-owned workspace (dtb_dt_workspace_bytes). */
typedef struct {
  int state_dim, act_dim, hidden, n_layer, n_head, n_inner, max_ep_len;
  int activation;  /* 1 relu, 2 gelu_new */
  int action_tanh; /* HF:527-529 */
  float ln_eps;
} dtb_dt_config;

typedef struct {
  int B, K;
  const float* states;          /* fp32 [B,K,sd], or NULL when states_bf16 is given */
  const void* states_bf16;      /* bf16 [B,K,sd] (sd % 8 == 0): wide multimodal state straight from the encoder towers */
  const float* actions;         /* fp32 [B,K,ad] */
  const float* returns_to_go;   /* fp32 [B,K] */
  const long long* timesteps;   /* int64 [B,K] */
  const float* attention_mask;  /* fp32 [B,K] (> 0 = valid) or NULL = all valid (HF:606-608) */
  int states_relu;              /* states_bf16 is a ReLU output (encoder towers): d_states is returned already masked by (state > 0) */
} dtb_dt_batch;

typedef struct dtb_dt dtb_dt;

int dtb_dt_create(const dtb_dt_config* cfg, dtb_dt** out);
int dtb_dt_destroy(dtb_dt* h);
/* ---- fused gradient clipping + AdamW (trainer tail) -------------------------------------------------
 * Replaces, after loss.backward() in the reference's SequenceTrainer.train_step():
 *     torch.nn.utils.clip_grad_norm_(model.parameters(), .25); optimizer.step()        (torch.optim.AdamW(lr, weight_decay))
 * over the flat fp32 parameter / gradient buffers of up to 8 modules (one global norm across all of them, as clip_grad_norm_ over
 * model.parameters()).  Three launches, no host synchronisation: the step count (and optionally the learning rate, for schedulers
 * inside a CUDA graph) live in device memory.  When `params_bf16` is given the bf16 shadow the GEMMs read is rewritten in the same pass.
 * scratch: dtb_adamw_scratch_floats() floats; after the call scratch[3] = the global gradient norm before clipping, scratch[0] = the
 * clip coefficient applied. */
typedef struct {
  float* params;       /* fp32 master parameters [n] */
  float* grads;        /* fp32 gradients [n] (scaled in place by the clip coefficient when write_clipped_grads != 0) */
  float* exp_avg;      /* Adam first moment [n] */
  float* exp_avg_sq;   /* Adam second moment [n] */
  void* params_bf16;   /* optional bf16 shadow [n] */
  long long n;
} dtb_adamw_segment;
typedef struct {
  double lr, beta1, beta2, eps, weight_decay; /* doubles: torch forms 1 - beta, lr / (1 - beta1^t), ... in Python doubles */
  double max_norm;            /* <= 0: no clipping */
  const float* lr_dev;        /* optional device scalar overriding lr */
  int write_clipped_grads;
  int warmup_steps;           /* > 0: linear warm-up, lr * min(t / warmup_steps, 1) at the t-th step (upstream's LambdaLR), on the device */
} dtb_adamw_hparams;
size_t dtb_adamw_scratch_floats(void);
int dtb_clip_adamw_step(void* stream, int nseg, const dtb_adamw_segment* segs, const dtb_adamw_hparams* hp, int* step_dev, float* scratch);

/* ---- GPU-side get_batch ------------------------------------------------------------------------------
 * Replaces upstream experiment.py get_batch() after the (trajectory, start) pairs are sampled on the host: window slice of at most K
 * steps, LEFT padding to K (states 0 — then normalised like every other row —, actions -10, rewards 0, dones 2, timesteps 0, mask 0),
 * `(s - mean) / std` and `rtg / scale` evaluated in double and rounded to fp32 exactly as numpy's promotions do, timesteps clamped to
 * max_ep_len - 1, returns-to-go with K + 1 entries (0 past the end of the trajectory).  The store is flat per-timestep rows in device
 * memory; `rtg_suffix` holds upstream's discount_cumsum(rewards, gamma = 1) per trajectory (computed once by the host loader).
 * Raw fields (frames, spectrograms: any 16-byte-multiple row) are copied with zero padding. */
typedef struct {
  const float* obs; int state_dim; const double* state_mean; const double* state_std; /* obs may be NULL when only raw fields are used */
  const float* act; int act_dim;
  const float* rew; const float* done; const double* rtg_suffix;                    /* [N_total] each; suffix sums in double (exact for fp32 data) */
  const long long* traj_offset;                                                      /* [n_traj + 1] first timestep of each trajectory */
  int n_raw; const void* raw_src[4]; int raw_row_bytes[4];
} dtb_traj_store;
typedef struct {
  float* states;         /* [B, K, state_dim] or NULL */
  float* actions;        /* [B, K, act_dim] or NULL */
  float* rewards;        /* [B, K] or NULL */
  long long* dones;      /* [B, K] or NULL */
  float* rtg;            /* [B, K + 1] or NULL */
  long long* timesteps;  /* [B, K] or NULL */
  float* mask;           /* [B, K] or NULL */
  void* raw_dst[4];      /* [B, K, row_bytes] per raw field */
} dtb_batch_out;
int dtb_get_batch(void* stream, const dtb_traj_store* store, const long long* traj_idx, const long long* start, int B, int K,
                  int max_ep_len, double scale, const dtb_batch_out* out);

/* Dropout (nn.Dropout at the three GPT-2 sites of the reference's trunk; HF modeling_gpt2 GPT2Model.drop (embd_pdrop),
 * GPT2Attention.attn_dropout (attn_pdrop, :60-62 of modeling_decision_transformer eager attention) and resid_dropout / GPT2MLP.dropout
 * (resid_pdrop, :217-218, :236-237); kzl-style DecisionTransformer passes resid_pdrop = attn_pdrop = dropout, default 0.1).
 * The masks are counter-based and stateless: keep(site, step, idx) = mix32(key(seed, step, site [, b*nh+h]) + idx) >= p * 2^32, drawn
 * inside the fused kernels (embed LN epilogue, c_proj GEMM epilogues before the residual add, flash-attention P tiles) and recomputed
 * by their backward kernels; nothing is stored.  `state` = 3 x uint32 in DEVICE memory {seed_lo, seed_hi, step}, caller-owned and
 * read at kernel run time, so a captured CUDA graph draws fresh masks on every replay: enqueue dtb_dropout_advance() once per
 * training step (it increments `step`).  All probabilities 0 (or state NULL) = evaluation mode.  Host-side setting: takes effect for
 * launches enqueued afterwards (re-capture graphs when it changes). */
int dtb_dt_set_dropout(dtb_dt* h, float embd_pdrop, float attn_pdrop, float resid_pdrop, const void* state);
int dtb_dropout_advance(void* stream, void* state);
/* Test / oracle hook: the keep bits of one site exactly as the kernels draw them at the state's current step.
 * site: 0 embd, 1 + 3l attention of layer l, 2 + 3l / 3 + 3l the residual dropouts after attn / mlp c_proj of layer l.
 * groups = 0: elementwise sites, out[i] for i < n (i = row * H + col).  groups = B * n_head: attention, out[g * n + i * T + j], n = T*T. */
int dtb_dropout_mask(void* stream, const void* state, unsigned site, float p, int groups, long long n, unsigned char* out);
/* Number of parameter tensors (9 + 12 * n_layer + 8) in flat order:
 *  embed_timestep.weight, embed_return.{weight,bias}, embed_state.{weight,bias}, embed_action.{weight,bias}, embed_ln.{weight,bias},
 *  per layer: ln_1.{w,b}, attn.c_attn.{w,b}, attn.c_proj.{w,b}, ln_2.{w,b}, mlp.c_fc.{w,b}, mlp.c_proj.{w,b},
 *  ln_f.{w,b}, predict_state.weight, predict_return.weight, predict_action.weight, predict_state.bias, predict_return.bias,
 *  predict_action.bias */
int dtb_dt_param_entries(const dtb_dt* h);
/* Fills offsets[i], numels[i] (element counts, fp32) and returns the padded total element count. */
long long dtb_dt_param_layout(const dtb_dt* h, long long* offsets, long long* numels);
size_t dtb_dt_workspace_bytes(const dtb_dt* h, int B, int K);
/* fp32 master -> bf16 shadow (one pass over the flat buffer). */
int dtb_dt_stage_params(const dtb_dt* h, void* stream, const float* params, void* params_bf16);
/* forward (HF:536-657).  state_preds / return_preds may be NULL (the loss never reads them, A8). */
int dtb_dt_forward(const dtb_dt* h, void* stream, const float* params, const void* params_bf16, const dtb_dt_batch* batch,
                   void* workspace, size_t workspace_bytes, float* action_preds, float* state_preds, float* return_preds);
/* masked action MSE on the action_preds of the last forward (kept in the workspace); writes loss_out[2] and the loss
 * gradient (scaled by grad_scale, e.g. 1/world_size) into the workspace for dtb_dt_backward(d_action_preds = NULL). */
int dtb_dt_loss(const dtb_dt* h, void* stream, const dtb_dt_batch* batch, const float* action_target, float grad_scale, void* workspace,
                size_t workspace_bytes, float* loss_out);
/* backward.  d_action_preds NULL = use the gradient dtb_dt_loss left in the workspace; d_state_preds / d_return_preds may be
 * NULL (zero).  Overwrites the gradients of the stages it runs.  Stages: 0 = ln_f + heads, 1..L = blocks L-1..0, L+1 = embedding;
 * [stage_begin, stage_end] inclusive lets a data-parallel host all-reduce finished stages while earlier layers still run.
 * d_states_bf16 (nullable): gradient w.r.t. the states — bf16 [B,K,sd] for a wide state (states_bf16 path, for the encoder towers), fp32
 * [B,K,sd] for vector states (state_dim <= 64). */
int dtb_dt_backward(const dtb_dt* h, void* stream, const float* params, const void* params_bf16, const dtb_dt_batch* batch,
                    void* workspace, size_t workspace_bytes, const float* d_action_preds, const float* d_state_preds,
                    const float* d_return_preds, float* grads, void* d_states_bf16, int stage_begin, int stage_end);
/* [begin, end) element range of the flat gradient buffer written by a backward stage. */
/* ---- KV-cached incremental get_action: one kernel launch (a cluster of 8 CTAs) per rollout step -------------------
 * Exact while the whole history fits the window (t <= K: no positional embedding, causal attention => the hidden states of earlier
 * tokens never change).  `cache` (dtb_dt_decode_cache_bytes(h, K) bytes, caller-owned, device) holds the per-layer keys / values of the
 * tokens of the first t_cached timesteps, written by earlier calls with the same history prefix; the call processes timesteps
 * t_cached .. t-1 (normally t = t_cached + 1), (re)computing the action token of timestep t_cached - 1 whose action was a placeholder
 * when it was cached, and writes action_out[act_dim] = prediction of the last state token.  History pointers: one row per timestep,
 * t rows (states fp32 [t, sd] for state_dim <= 64, else states_bf16 [t, sd]).  t_cached = 0 starts a new episode. */
int dtb_dt_decode_supported(const dtb_dt* h);
size_t dtb_dt_decode_cache_bytes(const dtb_dt* h, int K);
/* The kernel streams the weights by TMA bulk copies from a decode layout kept inside `cache` (per layer, matrix and CTA one contiguous
 * [K][N / C] block): build it once with dtb_dt_decode_stage, and again after every parameter update (the cached keys / values are then
 * stale too: restart with t_cached = 0). */
int dtb_dt_decode_stage(const dtb_dt* h, void* stream, const void* params_bf16, int K, void* cache, size_t cache_bytes);
int dtb_dt_decode(const dtb_dt* h, void* stream, const float* params, const void* params_bf16, const float* states, const void* states_bf16,
                  const float* actions, const float* returns_to_go, const long long* timesteps, int t, int t_cached, int K, void* cache,
                  size_t cache_bytes, float* action_out);
/* One rollout step of the same kernel in a form a CUDA graph can replay: the step count t (1-based, the history length of THIS call) is read
 * from device memory (*t_dev) and advanced to t + 1 by the kernel; the per-timestep inputs are 2-row windows holding timesteps t-2 and t-1
 * (row 0 unused when t = 1; dtb_pad_left with K = 2 builds them): states2 fp32 [2, state_dim] (state_dim <= 64) or, for wide states, the
 * newest timestep's features state_new_bf16 [state_dim]; actions2 [2, act_dim], returns_to_go2 [2], timesteps2 [2].  The cache must hold
 * the keys / values of timesteps 0 .. t-2 (append-only use after dtb_dt_decode); the caller keeps t <= K. */
int dtb_dt_decode_step(const dtb_dt* h, void* stream, const float* params, const void* params_bf16, const float* states2, const void* state_new_bf16,
                       const float* actions2, const float* returns_to_go2, const long long* timesteps2, int* t_dev, int K, void* cache,
                       size_t cache_bytes, float* action_out);
int dtb_dt_stage_grad_range(const dtb_dt* h, int stage, long long* begin, long long* end);

/* ---- multimodal state encoder (REF src/agents.py:30-55 conv_factory, :118-119 towers, :133-134 /255, :137-143 fusion) ------
 * frames fp32 [N, Cv, H, W], audio fp32 [N, Ca, H, W] (N = B*K timesteps, NCHW as the reference's obs tensors) ->
 * states bf16 [N, state_dim] = [relu(video_fc) | relu(audio_fc)] (fusion 0, 'concat', per-sample dim=1 semantics) or their
 * sum (fusion 1).  Convolutions = patch gather + tcgen05 GEMM with bias+ReLU epilogue, NHWC bf16 activations.
 * Parameter ENTRY order (PyTorch layouts, nn.Sequential indices of conv_factory): per tower (video, audio):
 *   big  : 0.weight [32,C,8,8], 0.bias, 2.weight [64,32,4,4], 2.bias, 4.weight [64,64,3,3], 4.bias, 7.weight [512, 64*h*w], 7.bias
 *   small: 0.weight [16,C,8,8], 0.bias, 2.weight [32,16,4,4], 2.bias, 5.weight [256, 32*h*w], 5.bias
 * The OFFSETS dtb_encoder_param_layout reports put both towers' conv tensors first and both FC tensors last (see
 * dtb_encoder_fc_grad_range); always address a tensor through its reported offset. */
typedef struct {
  int in_channels, in_h, in_w;
  int big; /* 1 = conv_factory('big'), 0 = 'small' */
} dtb_tower_config;
typedef struct {
  dtb_tower_config video, audio;
  int fusion;        /* 0 concat, 1 sum */
  float input_scale; /* 1/255 in the reference (agents.py:133-134) */
} dtb_encoder_config;
typedef struct dtb_encoder dtb_encoder;

int dtb_encoder_create(const dtb_encoder_config* cfg, dtb_encoder** out);
int dtb_encoder_destroy(dtb_encoder* e);
int dtb_encoder_state_dim(const dtb_encoder* e);
int dtb_encoder_param_entries(const dtb_encoder* e);
long long dtb_encoder_param_layout(const dtb_encoder* e, long long* offsets, long long* numels);
size_t dtb_encoder_workspace_bytes(const dtb_encoder* e, int N);
/* fp32 master (PyTorch layouts) -> bf16 staged copy in tensor-core K order (same offsets), followed by the data-gradient weight
 * matrices of the implicit-GEMM convolutions; the staged buffer has dtb_encoder_staged_elems() bf16 elements. */
long long dtb_encoder_staged_elems(const dtb_encoder* e);
int dtb_encoder_stage_params(const dtb_encoder* e, void* stream, const float* params, void* staged_bf16);
int dtb_encoder_forward(const dtb_encoder* e, void* stream, const float* params, const void* staged_bf16, int N, const float* frames,
                        const float* audio, void* workspace, size_t workspace_bytes, void* states_bf16);
/* d_states_bf16 [N, state_dim]; d_states_masked != 0 when it is already multiplied by (states > 0) (dtb_dt_backward does this when
 * dtb_dt_batch.states_relu is set); otherwise states_bf16 must be given and d_states_bf16 is masked IN PLACE.  frames / audio are the forward inputs again (the
 * implicit-GEMM conv1 weight gradient re-gathers its patches from them instead of saving them).  Overwrites grads. */
int dtb_encoder_backward(const dtb_encoder* e, void* stream, const float* params, const void* staged_bf16, int N, const float* frames,
                         const float* audio, const void* states_bf16,
                         const void* d_states_bf16, int d_states_masked, void* workspace, size_t workspace_bytes, float* grads);

/* Data-parallel overlap (REF cleanrl/ppo_atari_multigpu.py:345-359 exchanges all gradients after backward): the two FC layers hold
 * 7/8 of the towers' gradient bytes and are finished first in backward.  grads[begin, end) of dtb_encoder_fc_grad_range is that
 * contiguous FC range ([0, begin) = the conv layers); dtb_encoder_wait_fc_grads makes `stream` wait (cudaStreamWaitEvent, capturable)
 * until the FC gradients of the most recent dtb_encoder_backward call are final, so a communication stream can all-reduce them while
 * the conv layers' backward is still running. */
int dtb_encoder_fc_grad_range(const dtb_encoder* e, long long* begin, long long* end);
int dtb_encoder_wait_fc_grads(const dtb_encoder* e, void* stream);

/* Same, with the video frames as uint8 pixels [N, Cv, H, W]: the reference's pixel decode `img * 2. / 255. - 1` (fp32,
 * REF environments/Minecraft/Minecraft.py:333-335) runs inside the implicit-GEMM conv1 gather, bit-identical to passing the decoded
 * fp32 frames to dtb_encoder_forward / _backward; 4x fewer bytes cross PCIe and HBM.  Needs in_w % 4 == 0 and in_channels <= 4. */
int dtb_encoder_forward_u8(const dtb_encoder* e, void* stream, const float* params, const void* staged_bf16, int N,
                           const unsigned char* frames_u8, const float* audio, void* workspace, size_t workspace_bytes, void* states_bf16);
int dtb_encoder_backward_u8(const dtb_encoder* e, void* stream, const float* params, const void* staged_bf16, int N,
                            const unsigned char* frames_u8, const float* audio, const void* states_bf16, const void* d_states_bf16,
                            int d_states_masked, void* workspace, size_t workspace_bytes, float* grads);

/* ---- modality attention over the two tower features (REF src/agents.py:58-75 CaslAttention, :78-90 NewAttention; the call site
 * is agents.py:135-136 between the towers and the fusion) ---------------------------------------------------------------------
 *   type 0 'casl': z = tanh(video_fc(v) + audio_fc(a) + state_fc(h)) [32];   type 1 'new': z = tanh(fc([v | a | h])) [128]
 *   w = softmax(attention(z)) [2];  v' = w[0] v, a' = w[1] a;  states = [v' | a'] (fusion 0) or v' + a' (fusion 1, agents.py:137-143)
 * feat_bf16 [N, 2*feat] = [relu(video_fc) | relu(audio_fc)] as dtb_encoder_forward (fusion 0) writes it; hidden fp32 [N, 128] = the
 * recurrent state the reference's LSTM agent feeds in (lstm_state[0]); NULL = zeros (the Decision Transformer has no such state).
 * Stateless: the config is passed to every call; the workspace of the forward must be handed unchanged to the backward.
 * Flat fp32 parameter order (PyTorch layouts, module order of the reference classes):
 *   casl: audio_fc.{weight [32,feat], bias}, video_fc.{weight [32,feat], bias}, state_fc.{weight [32,128], bias}, attention.{weight [2,32], bias}
 *   new : fc.{weight [128, 2*feat+128], bias}, attention.{weight [2,128], bias} */
typedef struct {
  int type;   /* 0 casl, 1 new */
  int feat;   /* per-tower feature size: 512 ('big') or 256 ('small') */
  int fusion; /* 0 concat, 1 sum */
} dtb_modal_attn_config;
int dtb_modal_attn_param_entries(const dtb_modal_attn_config* cfg);
long long dtb_modal_attn_param_layout(const dtb_modal_attn_config* cfg, long long* offsets, long long* numels);
int dtb_modal_attn_state_dim(const dtb_modal_attn_config* cfg);
size_t dtb_modal_attn_workspace_bytes(const dtb_modal_attn_config* cfg, int N);
/* states_bf16 [N, state_dim]; attn_weights fp32 [N, 2] (nullable). */
int dtb_modal_attn_forward(const dtb_modal_attn_config* cfg, void* stream, const float* params, int N, const void* feat_bf16,
                           const float* hidden, void* workspace, size_t workspace_bytes, void* states_bf16, float* attn_weights);
/* d_states_bf16 [N, state_dim] -> d_feat_bf16 [N, 2*feat], already multiplied by the towers' ReLU mask (feat > 0): hand it to
 * dtb_encoder_backward with d_states_masked = 1.  Overwrites grads (flat, same layout as params); has_hidden = whether the forward
 * was given a hidden state (its weight columns get a zero gradient otherwise).  No gradient flows into hidden (the reference detaches
 * the LSTM state between steps). */
int dtb_modal_attn_backward(const dtb_modal_attn_config* cfg, void* stream, const float* params, int N, const void* feat_bf16,
                            int has_hidden, const void* d_states_bf16, void* workspace, size_t workspace_bytes, void* d_feat_bf16,
                            float* grads);

/* ---- audio front-end: wav -> MFCC image (REF environments/Minecraft/Minecraft.py:226-245 `_get_mfcc` / `_wav_to_mfcc`; the same code in
 * environments/Doorpuzzle/Doorpuzzle.py:264-282) -------------------------------------------------------------------------------------
 * cepstra = python_speech_features.mfcc(signal, samplerate) with its release-0.6 defaults (25 ms frames every 10 ms, rectangular
 * window, pre-emphasis 0.97, FFT = next power of two, 26 mel filters, log, DCT-II ortho -> 13, lifter 22, c0 = log frame energy),
 * computed in double like the reference; image = Pillow `Image.fromarray(cepstra.T).resize((out_w, out_h), BICUBIC)` (float32, bit-exact)
 * followed by the reference's float32 min-max normalisation to [-1, 1].
 * signal: int16 PCM (is_i16 = 1, as scipy.io.wavfile returns it) or float32, n_clips rows of n_samples (row stride clip_stride).  The
 * reference hands scipy's [n, channels] array of a stereo file straight to python_speech_features, whose pre-emphasis then acts per
 * channel and whose numpy.append flattens the result: pass the interleaved samples (n_samples = n * channels) and `channels` (the
 * pre-emphasis lag); framing then runs over the interleaved stream exactly as in the reference. */
long long dtb_mfcc_num_frames(long long n_samples, int samplerate);
/* nfft and the 28 filter-bank edge bins the kernel uses for this sample rate (test hook). */
int dtb_mfcc_filterbank_bins(int samplerate, int* nfft, int* bins);
size_t dtb_mfcc_workspace_bytes(int n_clips, long long n_samples, int samplerate, int out_w);
/* tables: dtb_mfcc_table_doubles(samplerate) doubles in device memory holding the FFT twiddles (per radix-4 pass), the DCT x lifter
 * matrix and the mel weights of every bin for that sample rate; built once by dtb_mfcc_build_tables and reusable by every later call. */
size_t dtb_mfcc_table_doubles(int samplerate);
int dtb_mfcc_build_tables(void* stream, int samplerate, double* tables);
/* cepstra: double [n_clips, num_frames, 13] */
int dtb_mfcc_cepstra(void* stream, const void* signal, int is_i16, int channels, int n_clips, long long n_samples, long long clip_stride, int samplerate,
                     const double* tables, double* cepstra);
/* scratch: float [n_clips, 13, out_w]; image: float [n_clips, out_h, out_w] */
int dtb_mfcc_resize_normalize(void* stream, const double* cepstra, int n_clips, long long numframes, int out_w, int out_h, float* scratch,
                              float* image);
int dtb_wav_to_mfcc_image(void* stream, const void* signal, int is_i16, int channels, int n_clips, long long n_samples, long long clip_stride,
                          int samplerate, int out_w, int out_h, void* workspace, size_t workspace_bytes, const double* tables /* NULL: built into the workspace */,
                          float* image);

#ifdef __cplusplus
}
#endif
#endif /* DTB200_H */
