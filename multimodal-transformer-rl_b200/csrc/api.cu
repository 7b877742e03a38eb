// extern "C" boundary of libdtb200.so (declared in include/dtb200.h).  No torch types, no exceptions.
#include "../../include/dtb200.h"
#include "kernels.h"

#include <string.h>

namespace dtb {
int prof_span(int i, int* family, double* start_ms, double* end_ms);
int prof_count();
void prof_enable(bool on);
void prof_reset();
int prof_report(int family, double* total_ms, double* work, long long* launches);
const char* last_error();
int require_sm100();
}  // namespace dtb

using namespace dtb;

extern "C" {

int dtb_abi_version(void) { return DTB_ABI_VERSION; }
const char* dtb_last_error(void) { return dtb::last_error(); }
int dtb_check_device(void) { return dtb::require_sm100(); }
long long dtb_launch_count(void) { return dtb::launch_count(); }
int dtb_set_option(const char* name, int value) {
  DTB_REQUIRE(name != nullptr, DTB_ERR_BAD_ARG, "dtb_set_option: null name");
  return dtb::opt_set(name, value);
}
void dtb_profile_set_balance(double flop_per_byte) { dtb::opt_set_balance(flop_per_byte); }
void dtb_profile_enable(int on) { dtb::prof_enable(on != 0); if (on) dtb::prof_reset(); }
int dtb_profile_spans(void) { return dtb::prof_count(); }
int dtb_profile_span(int i, int* family, double* start_ms, double* end_ms) { return dtb::prof_span(i, family, start_ms, end_ms); }
int dtb_profile_report(int family, double* total_ms, double* work, long long* launches) {
  return dtb::prof_report(family, total_ms, work, launches);
}

int dtb_gemm_bf16(void* stream, const void* A, int lda, int a_mn_major, const void* B, int ldb, int b_mn_major, int M, int N, int K,
                  const dtb_gemm_epilogue* e, int split_k, int bn_override) {
  DTB_REQUIRE(A && B && e, DTB_ERR_BAD_ARG, "dtb_gemm_bf16: null operand");
  DTB_TRY(require_sm100());
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  GemmEpilogue ep;
  memset(&ep, 0, sizeof(ep));
  ep.alpha = e->alpha == 0.f ? 1.f : e->alpha;
  const bool split = split_k > 1;
  if (split) {
    DTB_REQUIRE(e->partial && e->out_f32, DTB_ERR_BAD_ARG, "dtb_gemm_bf16: split-K needs partial and out_f32");
    ep.partial = e->partial;
    if (!e->accumulate && ep.alpha == 1.f) {  // plain store: the last-arriving CTA of every tile reduces the slices inside the GEMM
      ep.sk_out = e->out_f32;
      ep.sk_ldo = e->ldo_f;
    }
    int rc = gemm_bf16(st, (const bf16*)A, lda, a_mn_major, (const bf16*)B, ldb, b_mn_major, M, N, K, ep, split_k, bn_override);
    if (rc <= 0) return rc;
    return splitk_reduce(st, e->partial, rc, M, N, e->out_f32, e->ldo_f, e->accumulate, ep.alpha);
  }
  ep.bias = e->bias;
  ep.residual = e->residual;
  ep.ldr = e->ldr;
  ep.aux = (const bf16*)e->aux;
  ep.ldaux = e->ldaux;
  ep.act = e->act;
  ep.dact = e->dact;
  ep.out_bf16 = (bf16*)e->out_bf16;
  ep.ldo_b = e->ldo_b;
  ep.out_f32 = e->out_f32;
  ep.ldo_f = e->ldo_f;
  ep.out_preact = (bf16*)e->out_preact;
  int rc = gemm_bf16(st, (const bf16*)A, lda, a_mn_major, (const bf16*)B, ldb, b_mn_major, M, N, K, ep, 1, bn_override);
  return rc < 0 ? rc : DTB_OK;
}

int dtb_gemm_bf16_ln(void* stream, const void* A, int lda, int a_mn_major, const void* B, int ldb, int b_mn_major, int M, int N, int K,
                     const dtb_gemm_epilogue* e, const float* ln_gamma, const float* ln_beta, float ln_eps, void* ln_out_bf16, int ld_ln,
                     float* ln_mean, float* ln_rstd) {
  DTB_REQUIRE(A && B && e && ln_gamma && ln_beta && ln_out_bf16, DTB_ERR_BAD_ARG, "dtb_gemm_bf16_ln: null operand");
  DTB_REQUIRE(N <= 256 && N % 64 == 0, DTB_ERR_UNSUPPORTED, "dtb_gemm_bf16_ln: N = %d unsupported (multiple of 64, <= 256)", N);
  DTB_TRY(require_sm100());
  GemmEpilogue ep;
  memset(&ep, 0, sizeof(ep));
  ep.alpha = e->alpha == 0.f ? 1.f : e->alpha;
  ep.bias = e->bias;
  ep.residual = e->residual;
  ep.ldr = e->ldr;
  ep.aux = (const bf16*)e->aux;
  ep.ldaux = e->ldaux;
  ep.act = e->act;
  ep.dact = e->dact;
  ep.out_f32 = e->out_f32;
  ep.ldo_f = e->ldo_f;
  ep.ln_gamma = ln_gamma; ep.ln_beta = ln_beta; ep.ln_eps = ln_eps;
  ep.ln_out = (bf16*)ln_out_bf16; ep.ld_ln = ld_ln; ep.ln_mean = ln_mean; ep.ln_rstd = ln_rstd;
  int rc = gemm_bf16(static_cast<cudaStream_t>(stream), (const bf16*)A, lda, a_mn_major, (const bf16*)B, ldb, b_mn_major, M, N, K, ep, 1, N);
  return rc < 0 ? rc : DTB_OK;
}

int dtb_layernorm_fwd(void* stream, const float* x, const float* gamma, const float* beta, int M, int H, float eps, void* y, float* mean,
                      float* rstd) {
  DTB_REQUIRE(x && gamma && beta && y, DTB_ERR_BAD_ARG, "dtb_layernorm_fwd: null pointer");
  return ln_fwd((cudaStream_t)stream, x, gamma, beta, M, H, eps, (bf16*)y, mean, rstd);
}
size_t dtb_layernorm_bwd_scratch_floats(int H) { return ln_bwd_scratch_floats(H); }
int dtb_layernorm_bwd(void* stream, const void* dy, const float* x, const float* mean, const float* rstd, const float* gamma,
                      const float* dres, int M, int H, float* dx_f32, void* dx_bf16, float* dgamma, float* dbeta, float* scratch) {
  DTB_REQUIRE(dy && x && mean && rstd && gamma && dgamma && dbeta && scratch, DTB_ERR_BAD_ARG, "dtb_layernorm_bwd: null pointer");
  return ln_bwd((cudaStream_t)stream, (const bf16*)dy, x, mean, rstd, gamma, dres, M, H, dx_f32, (bf16*)dx_bf16, dgamma, dbeta, scratch);
}
int dtb_layernorm_bwd_colsum(void* stream, const void* dy, const float* x, const float* mean, const float* rstd, const float* gamma,
                             const float* dres, int M, int H, float* dx_f32, void* dx_bf16, float* dgamma, float* dbeta, float* dcolsum,
                             float* scratch) {
  DTB_REQUIRE(dy && x && mean && rstd && gamma && dgamma && dbeta && dcolsum && scratch, DTB_ERR_BAD_ARG, "dtb_layernorm_bwd_colsum: null pointer");
  return ln_bwd((cudaStream_t)stream, (const bf16*)dy, x, mean, rstd, gamma, dres, M, H, dx_f32, (bf16*)dx_bf16, dgamma, dbeta, scratch,
                drop_none(), dcolsum);
}
int dtb_attention_fwd(void* stream, const void* qkv, const float* mask, int B, int T, int n_head, int head_dim, void* out, float* lse) {
  DTB_REQUIRE(qkv && out && lse, DTB_ERR_BAD_ARG, "dtb_attention_fwd: null pointer");
  return attn_fwd((cudaStream_t)stream, (const bf16*)qkv, mask, B, T, n_head, head_dim, (bf16*)out, lse);
}
int dtb_attention_bwd(void* stream, const void* qkv, const float* mask, const void* out, const void* dout, const float* lse, int B,
                      int T, int n_head, int head_dim, void* dqkv, float* delta) {
  DTB_REQUIRE(qkv && out && dout && lse && dqkv && delta, DTB_ERR_BAD_ARG, "dtb_attention_bwd: null pointer");
  return attn_bwd((cudaStream_t)stream, (const bf16*)qkv, mask, (const bf16*)out, (const bf16*)dout, lse, B, T, n_head, head_dim,
                  (bf16*)dqkv, delta);
}
int dtb_attention_fwd_dropout(void* stream, const void* qkv, const float* mask, int B, int T, int n_head, int head_dim, void* out,
                              float* lse, const void* drop_state, unsigned site, float p) {
  DTB_REQUIRE(qkv && out && lse, DTB_ERR_BAD_ARG, "dtb_attention_fwd_dropout: null pointer");
  DTB_REQUIRE(p >= 0.f && p < 1.f, DTB_ERR_BAD_ARG, "dtb_attention_fwd_dropout: p must be in [0, 1)");
  return attn_fwd((cudaStream_t)stream, (const bf16*)qkv, mask, B, T, n_head, head_dim, (bf16*)out, lse, drop_make(drop_state, site, p));
}
int dtb_attention_bwd_dropout(void* stream, const void* qkv, const float* mask, const void* out, const void* dout, const float* lse, int B,
                              int T, int n_head, int head_dim, void* dqkv, float* delta, const void* drop_state, unsigned site, float p) {
  DTB_REQUIRE(qkv && out && dout && lse && dqkv && delta, DTB_ERR_BAD_ARG, "dtb_attention_bwd_dropout: null pointer");
  DTB_REQUIRE(p >= 0.f && p < 1.f, DTB_ERR_BAD_ARG, "dtb_attention_bwd_dropout: p must be in [0, 1)");
  return attn_bwd((cudaStream_t)stream, (const bf16*)qkv, mask, (const bf16*)out, (const bf16*)dout, lse, B, T, n_head, head_dim,
                  (bf16*)dqkv, delta, drop_make(drop_state, site, p));
}
int dtb_get_batch(void* stream, const dtb_traj_store* store, const long long* traj_idx, const long long* start, int B, int K, int max_ep_len,
                  double scale, const dtb_batch_out* out) {
  DTB_TRY(require_sm100());
  return get_batch((cudaStream_t)stream, store, traj_idx, start, B, K, max_ep_len, scale, out);
}
size_t dtb_adamw_scratch_floats(void) { return adamw_scratch_floats(); }
int dtb_clip_adamw_step(void* stream, int nseg, const dtb_adamw_segment* segs, const dtb_adamw_hparams* hp, int* step_dev, float* scratch) {
  DTB_TRY(require_sm100());
  return clip_adamw_step((cudaStream_t)stream, nseg, segs, hp, step_dev, scratch);
}
int dtb_cast_transpose(void* stream, const float* src, int R, int C, void* dst, void* dst_t) {
  DTB_REQUIRE(src, DTB_ERR_BAD_ARG, "dtb_cast_transpose: null pointer");
  return cast_transpose((cudaStream_t)stream, src, R, C, (bf16*)dst, (bf16*)dst_t);
}
int dtb_pad_left(void* stream, int nseg, const void* const* srcs, void* const* dsts, const int* row_bytes, int t, int K, float* mask_out) {
  DTB_REQUIRE(srcs && dsts && row_bytes, DTB_ERR_BAD_ARG, "dtb_pad_left: null argument");
  return pad_left((cudaStream_t)stream, nseg, srcs, dsts, row_bytes, t, K, mask_out);
}
size_t dtb_colsum_scratch_floats(int N) { return colsum_scratch_floats(N); }
int dtb_colsum_bf16(void* stream, const void* x, int M, int N, int ld, float* out, int accumulate, float* scratch) {
  DTB_REQUIRE(x && out && scratch, DTB_ERR_BAD_ARG, "dtb_colsum_bf16: null pointer");
  return colsum_bf16((cudaStream_t)stream, (const bf16*)x, M, N, ld, out, accumulate, scratch);
}
int dtb_colsum_bf16_static(void* stream, const void* x, int M, int N, int ld, float* out, int accumulate) {
  DTB_REQUIRE(x && out, DTB_ERR_BAD_ARG, "dtb_colsum_bf16_static: null pointer");
  return colsum_bf16_static((cudaStream_t)stream, (const bf16*)x, M, N, ld, out, accumulate);
}

int dtb_embed_fwd(void* stream, const float* states, const float* state_proj, const float* actions, const float* rtg,
                  const long long* timesteps, const float* table, const float* wR, const float* bR, const float* Ws, const float* bs,
                  const float* Wa, const float* ba, const float* gamma, const float* beta, int BK, int sd, int ad, int H, int E,
                  float eps, float* x, float* pre, float* mean, float* rstd, void* U, int P8, int u_states) {
  DTB_REQUIRE(actions && rtg && timesteps && table && wR && bR && Ws && bs && Wa && ba && gamma && beta && x, DTB_ERR_BAD_ARG,
              "dtb_embed_fwd: null pointer");
  return embed_fwd((cudaStream_t)stream, states, state_proj, actions, rtg, timesteps, table, wR, bR, Ws, bs, Wa, ba, gamma, beta, BK, sd,
                   ad, H, E, eps, x, pre, mean, rstd, (bf16*)U, P8, u_states);
}
int dtb_heads_fwd(void* stream, const void* y, const float* Wa, const float* ba, const float* Ws, const float* bs, const float* Wr,
                  const float* br, int BK, int H, int sd, int ad, int action_tanh, float* action_preds, float* state_preds,
                  float* return_preds) {
  DTB_REQUIRE(y && Wa && ba && action_preds, DTB_ERR_BAD_ARG, "dtb_heads_fwd: null pointer");
  DTB_REQUIRE((!state_preds || (Ws && bs)) && (!return_preds || (Wr && br)), DTB_ERR_BAD_ARG, "dtb_heads_fwd: missing head weights");
  return heads_fwd((cudaStream_t)stream, (const bf16*)y, Wa, ba, Ws, bs, Wr, br, BK, H, sd, ad, action_tanh, action_preds, state_preds,
                   return_preds);
}
size_t dtb_mse_scratch_floats(void) { return mse_scratch_floats(); }
int dtb_masked_mse(void* stream, const float* pred, const float* target, const float* mask, int BK, int ad, float grad_scale,
                   float* loss_out, float* d_pred, float* scratch) {
  DTB_REQUIRE(pred && target && loss_out && scratch, DTB_ERR_BAD_ARG, "dtb_masked_mse: null pointer");
  return mse_loss((cudaStream_t)stream, pred, target, mask, BK, ad, grad_scale, loss_out, d_pred, scratch);
}

}  // extern "C"
