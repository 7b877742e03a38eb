// Whole-model Decision Transformer forward / backward over the sm_100a kernels.
// Host-side orchestration only: every buffer (parameters, gradients, saved activations, scratch) is caller-owned;
// this file carves the workspace, picks GEMM operand majors so that weights are read in place
// (Conv1D W[in,out] = MN-major B forward, K-major B in dgrad; nn.Linear W[out,in] the other way round) and enqueues.
//   forward : HF modeling_decision_transformer.py:604-657 (embed, interleave, embed_ln, GPT-2 blocks :261-308, ln_f, heads)
//   backward: the autograd of the same graph, in reverse, split into stages for data-parallel overlap.
#include "../../include/dtb200.h"
#include "kernels.h"
#include "side.h"
#include "dt_handle.h"

#include <new>
#include <stdlib.h>
#include <string.h>
#include <vector>

namespace dtb {
int require_sm100();

static inline long long pad8(long long n) { return (n + 7) & ~7LL; }
static inline size_t align256(size_t n) { return (n + 255) & ~size_t(255); }

struct Workspace {
  // saved activations
  std::vector<float*> xs, xmid, mean1, rstd1, mean2, rstd2, lse;
  std::vector<bf16*> h1, h2, qkv, o, fc, pre;
  float *pre0, *mean0, *rstd0, *state_proj, *meanf, *rstdf, *ap, *d_ap;
  bf16 *U, *states_b, *yf;
  // backward temporaries
  float *dX, *delta, *G, *bsum, *partial, *cs_scratch, *ln_scratch, *mse_scratch, *tmpf;
  bf16 *dXb, *dXb2, *tmp_big, *tmp_big2, *tmp_h, *dzA, *dzS;
  size_t partial_floats;
  size_t total;
};
}  // namespace dtb

using namespace dtb;

namespace dtb {

static void build_layout(dtb_dt* h) {
  const dtb_dt_config& c = h->c;
  const long long H = c.hidden, I = c.n_inner, sd = c.state_dim, ad = c.act_dim, E = c.max_ep_len;
  std::vector<long long> n;
  n.insert(n.end(), {E * H, H, H, H * sd, H, H * ad, H, H, H});
  for (int l = 0; l < c.n_layer; ++l) n.insert(n.end(), {H, H, H * 3 * H, 3 * H, H * H, H, H, H, H * I, I, I * H, H});
  n.insert(n.end(), {H, H, sd * H, H, ad * H, sd, 1, ad});
  h->n_entries = (int)n.size();
  h->num = n;
  h->off.resize(n.size());
  long long o = 0;
  for (size_t i = 0; i < n.size(); ++i) {
    h->off[i] = o;
    o += pad8(n[i]);
  }
  h->total = o;
}

template <typename T>
static T* carve(uint8_t*& p, size_t count) {
  T* r = reinterpret_cast<T*>(p);
  p += align256(count * sizeof(T));
  return r;
}

static void carve_workspace(const dtb_dt* h, int B, int K, uint8_t* base, Workspace& w) {
  const dtb_dt_config& c = h->c;
  const size_t L = c.n_layer, H = c.hidden, I = c.n_inner, sd = c.state_dim, ad = c.act_dim;
  const size_t BK = (size_t)B * K, M = BK * 3, T = (size_t)K * 3;
  uint8_t* p = base;
  w.xs.resize(L + 1); w.xmid.resize(L); w.mean1.resize(L); w.rstd1.resize(L); w.mean2.resize(L); w.rstd2.resize(L); w.lse.resize(L);
  w.h1.resize(L); w.h2.resize(L); w.qkv.resize(L); w.o.resize(L); w.fc.resize(L); w.pre.resize(L);
  for (size_t l = 0; l <= L; ++l) w.xs[l] = carve<float>(p, M * H);
  for (size_t l = 0; l < L; ++l) {
    w.xmid[l] = carve<float>(p, M * H);
    w.mean1[l] = carve<float>(p, M); w.rstd1[l] = carve<float>(p, M);
    w.mean2[l] = carve<float>(p, M); w.rstd2[l] = carve<float>(p, M);
    w.lse[l] = carve<float>(p, (size_t)B * c.n_head * T);
    w.h1[l] = carve<bf16>(p, M * H); w.h2[l] = carve<bf16>(p, M * H);
    w.qkv[l] = carve<bf16>(p, M * 3 * H); w.o[l] = carve<bf16>(p, M * H);
    w.fc[l] = carve<bf16>(p, M * I);
    w.pre[l] = c.activation == 2 ? carve<bf16>(p, M * I) : nullptr;
  }
  w.pre0 = carve<float>(p, M * H); w.mean0 = carve<float>(p, M); w.rstd0 = carve<float>(p, M);
  w.U = carve<bf16>(p, BK * h->P8);
  w.state_proj = h->sd_small ? nullptr : carve<float>(p, BK * H);
  w.states_b = h->sd_small ? nullptr : carve<bf16>(p, BK * sd);
  w.yf = carve<bf16>(p, M * H); w.meanf = carve<float>(p, M); w.rstdf = carve<float>(p, M);
  w.ap = carve<float>(p, BK * ad); w.d_ap = carve<float>(p, BK * ad);
  // backward temporaries
  w.dX = carve<float>(p, M * H); w.dXb = carve<bf16>(p, M * H); w.dXb2 = carve<bf16>(p, M * H);
  const size_t big = I > 3 * H ? I : 3 * H;
  w.tmp_big = carve<bf16>(p, M * big); w.tmp_big2 = carve<bf16>(p, M * 3 * H); w.tmp_h = carve<bf16>(p, M * H);
  w.delta = carve<float>(p, (size_t)B * c.n_head * T);
  w.dzA = carve<bf16>(p, BK * h->ZA8); w.dzS = carve<bf16>(p, BK * h->ZS8);
  w.G = carve<float>(p, 3 * H * h->P8); w.bsum = carve<float>(p, 3 * H);
  w.tmpf = carve<float>(p, 64 + h->ZS8);
  // split-K partials: the largest split * Mout * Nout over the weight-gradient GEMMs
  size_t pf = 0;
  auto need = [&](size_t Mo, size_t No, size_t Kr) {
    const size_t s = (size_t)gemm_wgrad_split((int)Mo, (int)No, (int)Kr, (size_t)-1);
    if (s > 1 && s * Mo * No > pf) pf = s * Mo * No;
  };
  need(H, 3 * H, M); need(H, H, M); need(H, I, M); need(I, H, M);
  need(3 * H, h->P8, BK); need(ad, H, BK); need(sd + 1, H, BK);
  if (!h->sd_small) need(H, sd, BK);
  w.partial_floats = pf;
  w.partial = carve<float>(p, pf);
  size_t maxn = 3 * H > I ? 3 * H : I;
  if ((size_t)h->P8 > maxn) maxn = h->P8;
  w.cs_scratch = carve<float>(p, colsum_scratch_floats((int)maxn));
  w.ln_scratch = carve<float>(p, ln_bwd_scratch_floats((int)H));
  w.mse_scratch = carve<float>(p, mse_scratch_floats());
  w.total = (size_t)(p - base);
}

static int check_batch(const dtb_dt* h, const dtb_dt_batch* b) {
  DTB_REQUIRE(h && b, DTB_ERR_BAD_ARG, "dt: null handle / batch");
  DTB_REQUIRE(b->B >= 0 && b->K >= 1, DTB_ERR_BAD_ARG, "dt: bad batch shape B=%d K=%d", b->B, b->K);
  DTB_REQUIRE(b->actions && b->returns_to_go && b->timesteps, DTB_ERR_BAD_ARG, "dt: actions / returns_to_go / timesteps are required");
  if (h->sd_small) DTB_REQUIRE(b->states != nullptr, DTB_ERR_BAD_ARG, "dt: fp32 states required when state_dim <= 64");
  else DTB_REQUIRE(b->states || b->states_bf16, DTB_ERR_BAD_ARG, "dt: states (fp32) or states_bf16 required");
  return DTB_OK;
}

static inline const bf16* bfp(const void* params_bf16, long long off) { return reinterpret_cast<const bf16*>(params_bf16) + off; }

// D = A[M,K] x W, epilogue options; W in place.  conv1d: W is [K_in][N_out] (MN-major B); linear: W is [N_out][K_in] (K-major B).
static int gemm_fwd(cudaStream_t st, const bf16* A, int lda, const bf16* W, int ldw, bool w_mn_major, int M, int N, int K,
                    const GemmEpilogue& ep, int bn_override = 0) {
  const int rc = gemm_bf16(st, A, lda, 0, W, ldw, w_mn_major ? 1 : 0, M, N, K, ep, 1, bn_override);
  return rc < 0 ? rc : DTB_OK;
}
// LayerNorm in the c_proj epilogue: the CTA must own whole rows (BN == H).  dtb_set_option("ln_fused", 0) / DTB_LN_SEPARATE=1 keeps the
// standalone ln_fwd launches.
static bool dt_ln_fused(int H, int M) {
  // BN == H must also be a legal MN-major B tile (multiple of 64).  Below ~512 rows (get_action: M = 3K) the GEMM is pure latency and
  // one CTA walking whole rows through three TMEM passes is slower than four N-tile CTAs + a 3 us ln_fwd launch (measured: +20 us per call).
  return opt_ln_fused() && H <= 256 && H % 64 == 0 && M >= 512;
}
static void ep_set_ln(GemmEpilogue& ep, const float* gamma, const float* beta, bf16* out, int ld, float* mean, float* rstd, float eps) {
  ep.ln_gamma = gamma; ep.ln_beta = beta; ep.ln_out = out; ep.ld_ln = ld; ep.ln_mean = mean; ep.ln_rstd = rstd; ep.ln_eps = eps;
}
static GemmEpilogue ep_zero() {
  GemmEpilogue ep;
  memset(&ep, 0, sizeof(ep));
  ep.alpha = 1.f;
  return ep;
}

static int dt_forward(const dtb_dt* h, cudaStream_t st, const float* P, const void* Pb, const dtb_dt_batch* b, Workspace& w,
                      float* action_preds, float* state_preds, float* return_preds) {
  const dtb_dt_config& c = h->c;
  const int L = c.n_layer, H = c.hidden, I = c.n_inner, sd = c.state_dim, ad = c.act_dim, nh = c.n_head, d = H / nh;
  const int B = b->B, K = b->K, BK = B * K, M = BK * 3, T = 3 * K;
  if (BK == 0) return DTB_OK;
  // ---- embedding (A2) ----
  const float* states_f = nullptr;
  const float* sproj = nullptr;
  if (h->sd_small) {
    states_f = b->states;
  } else {
    const bf16* sb = reinterpret_cast<const bf16*>(b->states_bf16);
    if (!sb) {
      DTB_TRY(cast_bf16(st, b->states, (size_t)BK * sd, w.states_b));
      sb = w.states_b;
    }
    GemmEpilogue ep = ep_zero();
    ep.out_f32 = w.state_proj;
    ep.ldo_f = H;
    DTB_TRY(gemm_fwd(st, sb, sd, bfp(Pb, h->e(E_WS)), sd, false, BK, H, sd, ep));
    sproj = w.state_proj;
  }
  DTB_TRY(embed_fwd(st, states_f, sproj, b->actions, b->returns_to_go, b->timesteps, P + h->e(E_TABLE), P + h->e(E_WR), P + h->e(E_BR),
                    P + h->e(E_WS), P + h->e(E_BS), P + h->e(E_WA), P + h->e(E_BA), P + h->e(E_LNW), P + h->e(E_LNB), BK, sd, ad, H,
                    c.max_ep_len, c.ln_eps, w.xs[0], w.pre0, w.mean0, w.rstd0, w.U, h->P8, h->sd_small, h->drop_embd(),
                    P + h->l(0, L_LN1W), P + h->l(0, L_LN1B), w.h1[0], w.mean1[0], w.rstd1[0]));  // + ln_1 of block 0 in the same pass
  // Every other LayerNorm of the forward (ln_2, the next block's ln_1, ln_f) runs in the epilogue of the c_proj GEMM that produces its
  // input row (north star (2): "LayerNorm/GELU/residual fused into GEMM epilogues") when a CTA can own whole rows: H <= 256, H % 32 == 0.
  const bool ln_fused = dt_ln_fused(H, M);
  // ---- GPT-2 blocks (A4-A6) ----
  for (int l = 0; l < L; ++l) {
    {
      GemmEpilogue ep = ep_zero();
      ep.bias = P + h->l(l, L_BQKV);
      ep.out_bf16 = w.qkv[l];
      ep.ldo_b = 3 * H;
      DTB_TRY(gemm_fwd(st, w.h1[l], H, bfp(Pb, h->l(l, L_WQKV)), 3 * H, true, M, 3 * H, H, ep));
    }
    DTB_TRY(attn_fwd(st, w.qkv[l], b->attention_mask, B, T, nh, d, w.o[l], w.lse[l], h->drop_attn(l)));
    {
      GemmEpilogue ep = ep_zero();
      ep.bias = P + h->l(l, L_BO);
      ep.residual = w.xs[l];
      ep.ldr = H;
      ep.out_f32 = w.xmid[l];
      ep.ldo_f = H;
      ep.drop = h->drop_resid(l, 0);
      if (ln_fused) ep_set_ln(ep, P + h->l(l, L_LN2W), P + h->l(l, L_LN2B), w.h2[l], H, w.mean2[l], w.rstd2[l], c.ln_eps);
      DTB_TRY(gemm_fwd(st, w.o[l], H, bfp(Pb, h->l(l, L_WO)), H, true, M, H, H, ep, ln_fused ? H : 0));
    }
    if (!ln_fused) DTB_TRY(ln_fwd(st, w.xmid[l], P + h->l(l, L_LN2W), P + h->l(l, L_LN2B), M, H, c.ln_eps, w.h2[l], w.mean2[l], w.rstd2[l]));
    {
      GemmEpilogue ep = ep_zero();
      ep.bias = P + h->l(l, L_B1);
      ep.act = c.activation;
      ep.out_bf16 = w.fc[l];
      ep.ldo_b = I;
      ep.out_preact = w.pre[l];
      DTB_TRY(gemm_fwd(st, w.h2[l], H, bfp(Pb, h->l(l, L_W1)), I, true, M, I, H, ep));
    }
    {
      GemmEpilogue ep = ep_zero();
      ep.bias = P + h->l(l, L_B2);
      ep.residual = w.xmid[l];
      ep.ldr = H;
      ep.out_f32 = w.xs[l + 1];
      ep.ldo_f = H;
      ep.drop = h->drop_resid(l, 1);
      // the LayerNorm that reads this block's output: the next block's ln_1, or ln_f after the last block
      const float* g = l + 1 < L ? P + h->l(l + 1, L_LN1W) : P + h->t(T_LNFW);
      const float* be = l + 1 < L ? P + h->l(l + 1, L_LN1B) : P + h->t(T_LNFB);
      if (ln_fused) ep_set_ln(ep, g, be, l + 1 < L ? w.h1[l + 1] : w.yf, H, l + 1 < L ? w.mean1[l + 1] : w.meanf, l + 1 < L ? w.rstd1[l + 1] : w.rstdf, c.ln_eps);
      DTB_TRY(gemm_fwd(st, w.fc[l], I, bfp(Pb, h->l(l, L_W2)), H, true, M, H, I, ep, ln_fused ? H : 0));
      if (!ln_fused) DTB_TRY(ln_fwd(st, w.xs[l + 1], g, be, M, H, c.ln_eps, l + 1 < L ? w.h1[l + 1] : w.yf, l + 1 < L ? w.mean1[l + 1] : w.meanf,
                                    l + 1 < L ? w.rstd1[l + 1] : w.rstdf));
    }
  }
  // ---- heads (A7; ln_f ran in the last c_proj epilogue / just above) ----
  DTB_TRY(heads_fwd(st, w.yf, P + h->t(T_WPA), P + h->t(T_BPA), P + h->t(T_WPS), P + h->t(T_BPS), P + h->t(T_WPR), P + h->t(T_BPR), BK, H,
                    sd, ad, c.action_tanh, w.ap, state_preds, return_preds));
  if (action_preds) DTB_CUDA(cudaMemcpyAsync(action_preds, w.ap, (size_t)BK * ad * sizeof(float), cudaMemcpyDeviceToDevice, st));
  return DTB_OK;
}

static int bwd_heads(const dtb_dt* h, cudaStream_t st, const float* P, const dtb_dt_batch* b, Workspace& w, const float* d_action,
                     const float* d_state, const float* d_return, float* G) {
  SideStream& sd_ = h->side;
  const dtb_dt_config& c = h->c;
  const int L = c.n_layer, H = c.hidden, sd = c.state_dim, ad = c.act_dim;
  const int BK = b->B * b->K, M = BK * 3;
  DTB_TRY(heads_bwd(st, d_action, d_state, d_return, w.ap, P + h->t(T_WPA), P + h->t(T_WPS), P + h->t(T_WPR), BK, H, sd, ad,
                    c.action_tanh, h->ZA8, h->ZS8, w.tmp_h, w.dzA, w.dzS));
  DTB_TRY(sd_.fork(st));
  cudaStream_t ss = sd_.pick(st);  // head weight gradients: side branch
  const bf16* y_s = w.yf + H;      // state-token rows, stride 3H
  const bf16* y_a = w.yf + 2 * H;  // action-token rows
  if (d_action) {
    DTB_TRY(gemm_wgrad(ss, w.dzA, h->ZA8, y_s, 3 * H, ad, H, BK, G + h->t(T_WPA), H, w.partial, w.partial_floats));
    DTB_TRY(colsum_bf16(ss, w.dzA, BK, h->ZA8, h->ZA8, G + h->t(T_BPA), 0, w.cs_scratch));
  } else {
    DTB_CUDA(cudaMemsetAsync(G + h->t(T_WPA), 0, pad8((long long)ad * H) * sizeof(float), ss));
    DTB_CUDA(cudaMemsetAsync(G + h->t(T_BPA), 0, pad8(ad) * sizeof(float), ss));
  }
  if (d_state || d_return) {
    // predict_state.weight [sd,H] and predict_return.weight [1,H] are adjacent in the flat layout: one [sd+1, H] product
    DTB_TRY(gemm_wgrad(ss, w.dzS, h->ZS8, y_a, 3 * H, sd + 1, H, BK, G + h->t(T_WPS), H, w.partial, w.partial_floats));
    DTB_TRY(colsum_bf16(ss, w.dzS, BK, h->ZS8, h->ZS8, w.tmpf, 0, w.cs_scratch));
    DTB_TRY(copy2d(ss, w.tmpf, h->ZS8, 1, sd, G + h->t(T_BPS), sd));
    DTB_TRY(copy2d(ss, w.tmpf + sd, h->ZS8, 1, 1, G + h->t(T_BPR), 1));
  } else {
    DTB_CUDA(cudaMemsetAsync(G + h->t(T_WPS), 0, (pad8((long long)sd * H) + pad8(H)) * sizeof(float), ss));
    DTB_CUDA(cudaMemsetAsync(G + h->t(T_BPS), 0, (pad8(sd) + pad8(1)) * sizeof(float), ss));
  }
  // the bf16 copy of every residual-stream gradient carries the dropout mask of the c_proj site that consumes it next; its column sums
  // are that c_proj's bias gradient (here: mlp.c_proj.bias of the last block), reduced inside the same kernel
  DTB_TRY(ln_bwd(st, w.tmp_h, w.xs[L], w.meanf, w.rstdf, P + h->t(T_LNFW), nullptr, M, H, w.dX, w.dXb, G + h->t(T_LNFW), G + h->t(T_LNFB),
                 w.ln_scratch, h->drop_resid(L - 1, 1), G + h->l(L - 1, L_B2)));
  return DTB_OK;
}

static int bwd_layer(const dtb_dt* h, cudaStream_t st, const float* P, const void* Pb, const dtb_dt_batch* b, Workspace& w, int l,
                     float* G, bool& attn_marked) {
  const dtb_dt_config& c = h->c;
  const int H = c.hidden, I = c.n_inner, nh = c.n_head, d = H / nh;
  const int B = b->B, BK = B * b->K, M = BK * 3, T = 3 * b->K;
  SideStream& sd = h->side;
  cudaStream_t ss = sd.pick(st);
  // Main stream: the data-gradient chain.  Side stream: weight / bias gradients (they only read).  X / Y ping-pong the bf16 gradient
  // stream and T1 / T2 hold dfc / dqkv, so the chain never overwrites a buffer the side branch may still be reading without a wait.
  bf16 *X = w.dXb, *Y = w.dXb2, *T1 = w.tmp_big, *T2 = w.tmp_big2;
  // ---- MLP ----
  {  // dfc = (dX W2^T) * act'(.)
    GemmEpilogue ep = ep_zero();
    ep.dact = c.activation;
    ep.aux = c.activation == 2 ? w.pre[l] : w.fc[l];
    ep.ldaux = I;
    ep.out_bf16 = T1;
    ep.ldo_b = I;
    DTB_TRY(gemm_fwd(st, X, H, bfp(Pb, h->l(l, L_W2)), H, false, M, I, H, ep));
  }
  DTB_TRY(sd.fork(st));
  DTB_TRY(gemm_wgrad(ss, w.fc[l], I, X, H, I, H, M, G + h->l(l, L_W2), H, w.partial, w.partial_floats));  // (c_proj.bias: by the ln_bwd that produced X)
  DTB_TRY(gemm_wgrad(ss, w.h2[l], H, T1, I, H, I, M, G + h->l(l, L_W1), I, w.partial, w.partial_floats));
  DTB_TRY(colsum_bf16(ss, T1, M, I, I, G + h->l(l, L_B1), 0, w.cs_scratch));
  DTB_TRY(sd.mark(sd.ev_a));
  {  // dh2 = dfc W1^T
    GemmEpilogue ep = ep_zero();
    ep.out_bf16 = w.tmp_h;
    ep.ldo_b = H;
    DTB_TRY(gemm_fwd(st, T1, I, bfp(Pb, h->l(l, L_W1)), I, false, M, H, I, ep));
  }
  if (attn_marked) DTB_TRY(sd.wait(st, sd.ev_b));  // the previous block's attention weight gradients still read Y / T2
  DTB_TRY(ln_bwd(st, w.tmp_h, w.xmid[l], w.mean2[l], w.rstd2[l], P + h->l(l, L_LN2W), w.dX, M, H, w.dX, Y, G + h->l(l, L_LN2W),
                 G + h->l(l, L_LN2B), w.ln_scratch, h->drop_resid(l, 0), G + h->l(l, L_BO)));  // + attn.c_proj.bias gradient = colsum(Y)
  // ---- attention ----
  {  // do = dX Wo^T
    GemmEpilogue ep = ep_zero();
    ep.out_bf16 = w.tmp_h;
    ep.ldo_b = H;
    DTB_TRY(gemm_fwd(st, Y, H, bfp(Pb, h->l(l, L_WO)), H, false, M, H, H, ep));
  }
  DTB_TRY(attn_bwd(st, w.qkv[l], b->attention_mask, w.o[l], w.tmp_h, w.lse[l], B, T, nh, d, T2, w.delta, h->drop_attn(l)));
  DTB_TRY(sd.fork(st));
  DTB_TRY(gemm_wgrad(ss, w.o[l], H, Y, H, H, H, M, G + h->l(l, L_WO), H, w.partial, w.partial_floats));
  DTB_TRY(gemm_wgrad(ss, w.h1[l], H, T2, 3 * H, H, 3 * H, M, G + h->l(l, L_WQKV), 3 * H, w.partial, w.partial_floats));
  DTB_TRY(colsum_bf16(ss, T2, M, 3 * H, 3 * H, G + h->l(l, L_BQKV), 0, w.cs_scratch));
  DTB_TRY(sd.mark(sd.ev_b));
  attn_marked = sd.on();
  {  // dh1 = dqkv Wqkv^T
    GemmEpilogue ep = ep_zero();
    ep.out_bf16 = w.tmp_h;
    ep.ldo_b = H;
    DTB_TRY(gemm_fwd(st, T2, 3 * H, bfp(Pb, h->l(l, L_WQKV)), 3 * H, false, M, H, 3 * H, ep));
  }
  DTB_TRY(sd.wait(st, sd.ev_a));  // the MLP weight gradients read X and T1, which the next two kernels / the next block overwrite
  DTB_TRY(ln_bwd(st, w.tmp_h, w.xs[l], w.mean1[l], w.rstd1[l], P + h->l(l, L_LN1W), w.dX, M, H, w.dX, X, G + h->l(l, L_LN1W),
                 G + h->l(l, L_LN1B), w.ln_scratch, l > 0 ? h->drop_resid(l - 1, 1) : h->drop_embd(),
                 l > 0 ? G + h->l(l - 1, L_B2) : nullptr));  // + the previous block's mlp.c_proj.bias gradient = colsum(X)
  return DTB_OK;
}

static int bwd_embed(const dtb_dt* h, cudaStream_t st, const float* P, const void* Pb, const dtb_dt_batch* b, Workspace& w, float* G,
                     void* d_states_out) {
  bf16* d_states = h->sd_small ? nullptr : reinterpret_cast<bf16*>(d_states_out);
  const dtb_dt_config& c = h->c;
  const int H = c.hidden, sd = c.state_dim, ad = c.act_dim, E = c.max_ep_len;
  const int BK = b->B * b->K, M = BK * 3;
  // embed_ln backward: d_pre (bf16, [BK, 3H] view of [M, H]) -> tmp_h
  DTB_TRY(ln_bwd(st, w.dXb, w.pre0, w.mean0, w.rstd0, P + h->e(E_LNW), nullptr, M, H, nullptr, w.tmp_h, G + h->e(E_LNW), G + h->e(E_LNB),
                 w.ln_scratch));
  SideStream& sside = h->side;
  DTB_TRY(sside.fork(st));
  cudaStream_t ss = sside.pick(st);  // parameter gradients on the side branch; d_states (towers' input) stays on the main chain
  DTB_TRY(embed_te_scatter(ss, w.tmp_h, b->timesteps, b->attention_mask, BK, H, E, G + h->e(E_TABLE)));
  DTB_TRY(gemm_wgrad(ss, w.tmp_h, 3 * H, w.U, h->P8, 3 * H, h->P8, BK, w.G, h->P8, w.partial, w.partial_floats));
  DTB_TRY(colsum_bf16(ss, w.tmp_h, BK, 3 * H, 3 * H, w.bsum, 0, w.cs_scratch));
  DTB_TRY(embed_unpack(ss, w.G, w.bsum, H, sd, ad, h->P8, h->sd_small, G + h->e(E_WR), G + h->e(E_BR), G + h->e(E_WS), G + h->e(E_BS),
                       G + h->e(E_WA), G + h->e(E_BA)));
  if (h->sd_small && d_states_out)  // gradient w.r.t. the fp32 vector states (autograd bridge with states.requires_grad)
    DTB_TRY(embed_dstates(st, w.tmp_h, P + h->e(E_WS), BK, H, sd, reinterpret_cast<float*>(d_states_out)));
  if (!h->sd_small) {
    const bf16* dpre_s = w.tmp_h + H;  // state-token rows of d_pre, stride 3H
    const bf16* sb = b->states_bf16 ? reinterpret_cast<const bf16*>(b->states_bf16) : w.states_b;
    DTB_TRY(gemm_wgrad(ss, dpre_s, 3 * H, sb, sd, H, sd, BK, G + h->e(E_WS), sd, w.partial, w.partial_floats));
    if (d_states) {  // d_states = d_pre_s Ws   (Ws [H, sd] read as MN-major B)
      GemmEpilogue ep = ep_zero();
      ep.out_bf16 = d_states;
      ep.ldo_b = sd;
      if (b->states_relu) {  // the wide state is relu(features): hand the towers dZ directly
        ep.dact = 1;
        ep.aux = sb;
        ep.ldaux = sd;
      }
      DTB_TRY(gemm_fwd(st, dpre_s, 3 * H, bfp(Pb, h->e(E_WS)), sd, true, BK, sd, H, ep));
    }
  }
  return DTB_OK;
}

}  // namespace dtb

extern "C" {

int dtb_dt_create(const dtb_dt_config* cfg, dtb_dt** out) {
  DTB_REQUIRE(cfg && out, DTB_ERR_BAD_ARG, "dtb_dt_create: null argument");
  const dtb_dt_config& c = *cfg;
  DTB_REQUIRE(c.state_dim >= 1 && c.act_dim >= 1 && c.act_dim <= 64 && c.hidden >= 8 && c.n_layer >= 1 && c.n_head >= 1 && c.n_inner >= 8 &&
                  c.max_ep_len >= 1,
              DTB_ERR_BAD_ARG, "dtb_dt_create: bad config");
  DTB_REQUIRE(c.hidden % 8 == 0 && c.hidden <= 1024, DTB_ERR_UNSUPPORTED, "hidden_size %d unsupported (multiple of 8, <= 1024)", c.hidden);
  DTB_REQUIRE(c.n_inner % 8 == 0, DTB_ERR_UNSUPPORTED, "n_inner %d must be a multiple of 8", c.n_inner);
  DTB_REQUIRE(c.hidden % c.n_head == 0, DTB_ERR_BAD_ARG, "hidden_size %d not divisible by n_head %d", c.hidden, c.n_head);
  const int d = c.hidden / c.n_head;
  DTB_REQUIRE(d == 32 || d == 64 || d == 128, DTB_ERR_UNSUPPORTED, "head_dim %d unsupported (32, 64, 128)", d);
  DTB_REQUIRE(c.activation == 1 || c.activation == 2, DTB_ERR_UNSUPPORTED, "activation must be relu (1) or gelu_new (2)");
  DTB_REQUIRE(c.state_dim <= 64 || c.state_dim % 8 == 0, DTB_ERR_UNSUPPORTED, "state_dim %d > 64 must be a multiple of 8", c.state_dim);
  dtb_dt* h = new (std::nothrow) dtb_dt();
  DTB_REQUIRE(h != nullptr, DTB_ERR_BAD_ARG, "dtb_dt_create: out of memory");
  h->c = c;
  h->sd_small = c.state_dim <= 64;
  h->P8 = (int)pad8(1 + c.act_dim + (h->sd_small ? c.state_dim : 0));
  h->ZA8 = (int)pad8(c.act_dim);
  h->ZS8 = (int)pad8(c.state_dim + 1);
  build_layout(h);
  *out = h;
  return DTB_OK;
}

int dtb_dt_destroy(dtb_dt* h) {
  if (h) h->side.destroy();
  delete h;
  return DTB_OK;
}

int dtb_dt_set_dropout(dtb_dt* h, float embd_pdrop, float attn_pdrop, float resid_pdrop, const void* state) {
  DTB_REQUIRE(h != nullptr, DTB_ERR_BAD_ARG, "dtb_dt_set_dropout: null handle");
  DTB_REQUIRE(embd_pdrop >= 0.f && embd_pdrop < 1.f && attn_pdrop >= 0.f && attn_pdrop < 1.f && resid_pdrop >= 0.f && resid_pdrop < 1.f,
              DTB_ERR_BAD_ARG, "dtb_dt_set_dropout: probabilities must be in [0, 1)");
  const bool any = embd_pdrop > 0.f || attn_pdrop > 0.f || resid_pdrop > 0.f;
  DTB_REQUIRE(!any || state != nullptr, DTB_ERR_BAD_ARG, "dtb_dt_set_dropout: a device state buffer is required when a probability is > 0");
  h->p_embd = embd_pdrop; h->p_attn = attn_pdrop; h->p_resid = resid_pdrop;
  h->drop_state = any ? state : nullptr;
  return DTB_OK;
}

int dtb_dropout_advance(void* stream, void* state) { return dropout_advance((cudaStream_t)stream, state); }

int dtb_dropout_mask(void* stream, const void* state, unsigned site, float p, int groups, long long n, unsigned char* out) {
  DTB_REQUIRE(state && p > 0.f && p < 1.f, DTB_ERR_BAD_ARG, "dtb_dropout_mask: state and 0 < p < 1 required");
  return dropout_mask((cudaStream_t)stream, drop_make(state, site, p), groups, n, out);
}

int dtb_dt_param_entries(const dtb_dt* h) { return h ? h->n_entries : DTB_ERR_BAD_ARG; }

long long dtb_dt_param_layout(const dtb_dt* h, long long* offsets, long long* numels) {
  if (!h) return DTB_ERR_BAD_ARG;
  for (int i = 0; i < h->n_entries; ++i) {
    if (offsets) offsets[i] = h->off[i];
    if (numels) numels[i] = h->num[i];
  }
  return h->total;
}

size_t dtb_dt_workspace_bytes(const dtb_dt* h, int B, int K) {
  if (!h || B < 0 || K < 1) return 0;
  Workspace w;
  carve_workspace(h, B < 1 ? 1 : B, K, nullptr, w);
  return w.total + 256;
}

int dtb_dt_stage_params(const dtb_dt* h, void* stream, const float* params, void* params_bf16) {
  DTB_REQUIRE(h && params && params_bf16, DTB_ERR_BAD_ARG, "dtb_dt_stage_params: null argument");
  return cast_bf16((cudaStream_t)stream, params, (size_t)h->total, (bf16*)params_bf16);
}

static int get_ws(const dtb_dt* h, const dtb_dt_batch* b, void* workspace, size_t bytes, Workspace& w) {
  DTB_REQUIRE(workspace != nullptr, DTB_ERR_WORKSPACE, "dt: null workspace");
  uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(workspace) + 255) & ~uintptr_t(255));
  carve_workspace(h, b->B < 1 ? 1 : b->B, b->K, base, w);
  DTB_REQUIRE((size_t)(base - reinterpret_cast<uint8_t*>(workspace)) + w.total <= bytes, DTB_ERR_WORKSPACE,
              "dt: workspace too small (%zu bytes given, %zu needed)", bytes, w.total + 256);
  return DTB_OK;
}

int dtb_dt_forward(const dtb_dt* h, void* stream, const float* params, const void* params_bf16, const dtb_dt_batch* batch, void* workspace,
                   size_t workspace_bytes, float* action_preds, float* state_preds, float* return_preds) {
  DTB_TRY(check_batch(h, batch));
  DTB_REQUIRE(params && params_bf16, DTB_ERR_BAD_ARG, "dtb_dt_forward: null parameters");
  DTB_TRY(require_sm100());
  Workspace w;
  DTB_TRY(get_ws(h, batch, workspace, workspace_bytes, w));
  return dt_forward(h, (cudaStream_t)stream, params, params_bf16, batch, w, action_preds, state_preds, return_preds);
}

int dtb_dt_loss(const dtb_dt* h, void* stream, const dtb_dt_batch* batch, const float* action_target, float grad_scale, void* workspace,
                size_t workspace_bytes, float* loss_out) {
  DTB_TRY(check_batch(h, batch));
  DTB_REQUIRE(action_target && loss_out, DTB_ERR_BAD_ARG, "dtb_dt_loss: null argument");
  Workspace w;
  DTB_TRY(get_ws(h, batch, workspace, workspace_bytes, w));
  return mse_loss((cudaStream_t)stream, w.ap, action_target, batch->attention_mask, batch->B * batch->K, h->c.act_dim, grad_scale, loss_out,
                  w.d_ap, w.mse_scratch);
}

int dtb_dt_backward(const dtb_dt* h, void* stream, const float* params, const void* params_bf16, const dtb_dt_batch* batch, void* workspace,
                    size_t workspace_bytes, const float* d_action_preds, const float* d_state_preds, const float* d_return_preds,
                    float* grads, void* d_states_bf16, int stage_begin, int stage_end) {
  DTB_TRY(check_batch(h, batch));
  DTB_REQUIRE(params && params_bf16 && grads, DTB_ERR_BAD_ARG, "dtb_dt_backward: null argument");
  const int L = h->c.n_layer;
  DTB_REQUIRE(stage_begin >= 0 && stage_end <= L + 1 && stage_begin <= stage_end, DTB_ERR_BAD_ARG, "dtb_dt_backward: bad stage range [%d, %d]",
              stage_begin, stage_end);
  cudaStream_t st = (cudaStream_t)stream;
  Workspace w;
  DTB_TRY(get_ws(h, batch, workspace, workspace_bytes, w));
  if (batch->B * batch->K == 0) {
    long long b0, e0, b1, e1;
    dtb_dt_stage_grad_range(h, stage_begin, &b0, &e0);
    dtb_dt_stage_grad_range(h, stage_end, &b1, &e1);
    const long long lo = b0 < b1 ? b0 : b1, hi = e0 > e1 ? e0 : e1;
    DTB_CUDA(cudaMemsetAsync(grads + lo, 0, (size_t)(hi - lo) * sizeof(float), st));
    return DTB_OK;
  }
  DTB_TRY(h->side.init());
  bool attn_marked = false;
  for (int s = stage_begin; s <= stage_end; ++s) {
    if (s == 0) DTB_TRY(bwd_heads(h, st, params, batch, w, d_action_preds ? d_action_preds : w.d_ap, d_state_preds, d_return_preds, grads));
    else if (s <= L) DTB_TRY(bwd_layer(h, st, params, params_bf16, batch, w, L - s, grads, attn_marked));
    else DTB_TRY(bwd_embed(h, st, params, params_bf16, batch, w, grads, d_states_bf16));
  }
  return h->side.join(st);
}

int dtb_dt_stage_grad_range(const dtb_dt* h, int stage, long long* begin, long long* end) {
  DTB_REQUIRE(h && begin && end, DTB_ERR_BAD_ARG, "dtb_dt_stage_grad_range: null argument");
  const int L = h->c.n_layer;
  DTB_REQUIRE(stage >= 0 && stage <= L + 1, DTB_ERR_BAD_ARG, "dtb_dt_stage_grad_range: bad stage %d", stage);
  if (stage == 0) { *begin = h->t(0); *end = h->total; }
  else if (stage <= L) { const int l = L - stage; *begin = h->l(l, 0); *end = l + 1 < L ? h->l(l + 1, 0) : h->t(0); }
  else { *begin = 0; *end = h->l(0, 0); }
  return DTB_OK;
}

}  // extern "C"
