// Multimodal state encoder: video + audio Nature-CNN towers and fusion (REF src/agents.py:30-55 conv_factory,
// :118-119 two towers, :133-134 channel select + /255, :137-143 fusion; A11, A12) on the tensor-core GEMM.
//   tower('big')  : Conv(C,32,8,s4) ReLU Conv(32,64,4,s2) ReLU Conv(64,64,3,s1) ReLU Flatten Linear(64*h3*w3,512) ReLU
//   tower('small'): Conv(C,16,8,s4) ReLU Conv(16,32,4,s2) ReLU Flatten Linear(32*h2*w2,256) ReLU
// Activations are channels-last bf16 (the GEMM output [N*ho*wo, Cout] IS the NHWC tensor); weights are staged once per
// optimizer step into bf16 in the matching K order; the FC epilogue writes straight into its half of the fused state
// buffer [N, state_dim] ('concat' costs nothing).  Parameters / gradients use one flat fp32 buffer in PyTorch layouts
// (names video_net.{0,2,4,7}.*, audio_net.* as in the reference's nn.Sequential), so state_dicts round-trip.
#include "../../include/dtb200.h"
#include "kernels.h"
#include "side.h"

#include <new>
#include <string.h>
#include <vector>

namespace dtb {
int require_sm100();

struct ConvSpec { int cin, cout, k, stride, hin, win, hout, wout; };

struct Tower {
  int C, H, W, big;
  int nconv;
  ConvSpec conv[3];
  int flat, feat;
  // flat-parameter offsets: conv w/b ..., fc w/b
  long long w_off[4], b_off[4];
  // conv layers >= 1 that run as implicit GEMMs (convn.cu): offset of their data-gradient weights in the staged bf16 buffer
  bool implicit[3];
  long long dg_off[3];
};

static inline long long pad8l(long long n) { return (n + 7) & ~7LL; }
static inline size_t al256(size_t n) { return (n + 255) & ~size_t(255); }

struct TowerWs {
  bf16* P[3];   // saved patches per conv
  bf16* Y[3];   // saved activations (NHWC)
  bf16* dZ[3];  // gradient w.r.t. pre-activation outputs
  bf16* dP;     // patch-gradient scratch
  float* gstage;  // staged-order weight gradient scratch
  float* c1_partial;  // implicit conv1 wgrad per-CTA partials
  float* cn_partial;  // implicit conv2 / conv3 wgrad per-CTA partials
};

struct EncWs {
  TowerWs t[2];
  bf16* feat[2];   // 'sum' fusion: per-tower features
  bf16* dfeat;     // 'sum' fusion: masked feature gradient
  float* partial[2];     // split-K partials, one buffer per tower (their weight-gradient branches run concurrently)
  size_t partial_floats;
  float* cs_scratch[2];  // column-sum scratch, one per tower
  size_t total;
};
}  // namespace dtb

using namespace dtb;

struct dtb_encoder {
  dtb_encoder_config c;
  Tower t[2];
  int state_dim;
  long long total;
  long long fc_begin;  // flat offset where the two FC layers' tensors start (everything before it belongs to the conv layers)
  long long staged_total;  // elements of the staged bf16 buffer: the flat layout + the data-gradient weights of the implicit convs
  int n_entries;
  std::vector<long long> off, num;
  mutable dtb::SideStream side;   // weight-gradient branch of the backward (and the audio tower in forward)
  mutable dtb::SideStream side2;  // the audio tower's data-gradient chain in backward, concurrent with the video tower's
  mutable dtb::SideStream side3;  // the audio tower's weight-gradient branch (the video tower's uses `side`)
  mutable cudaEvent_t fc_ev[2] = {nullptr, nullptr};  // recorded once a tower's FC weight + bias gradients are enqueued (backward)
};

namespace dtb {

static int build_tower(Tower& t, const dtb_tower_config& c) {
  t.C = c.in_channels; t.H = c.in_h; t.W = c.in_w; t.big = c.big;
  DTB_REQUIRE(t.C >= 1 && t.C <= 16, DTB_ERR_UNSUPPORTED, "encoder: in_channels %d unsupported (1..16)", t.C);
  DTB_REQUIRE(t.H >= 36 && t.W >= 36 && t.W % 4 == 0, DTB_ERR_UNSUPPORTED, "encoder: input %dx%d unsupported (>= 36, W %% 4 == 0)", t.H, t.W);
  const int spec_big[3][3] = {{32, 8, 4}, {64, 4, 2}, {64, 3, 1}};
  const int spec_small[2][3] = {{16, 8, 4}, {32, 4, 2}};
  t.nconv = t.big ? 3 : 2;
  int cin = t.C, h = t.H, w = t.W;
  for (int i = 0; i < t.nconv; ++i) {
    const int* s = t.big ? spec_big[i] : spec_small[i];
    ConvSpec& cs = t.conv[i];
    cs.cin = cin; cs.cout = s[0]; cs.k = s[1]; cs.stride = s[2]; cs.hin = h; cs.win = w;
    cs.hout = (h - cs.k) / cs.stride + 1; cs.wout = (w - cs.k) / cs.stride + 1;
    DTB_REQUIRE(cs.hout >= 1 && cs.wout >= 1, DTB_ERR_UNSUPPORTED, "encoder: input too small for the tower");
    cin = cs.cout; h = cs.hout; w = cs.wout;
  }
  t.flat = cin * h * w;
  t.feat = t.big ? 512 : 256;
  return DTB_OK;
}

static void layout(dtb_encoder* e) {
  // Offsets: the conv tensors of both towers first, then both FC tensors, so that the FC gradients (the bulk of the bytes, finished
  // first in backward) are one contiguous range and the conv gradients (finished last) another: a data-parallel host exchanges the
  // FC range while the conv layers are still running (dtb_encoder_fc_grad_range / dtb_encoder_wait_fc_grads).  The ENTRY order
  // reported to the host stays the reference's module order (video_net.{0,2,4,7}, audio_net.{0,2,4,7}).
  long long o = 0;
  e->off.clear(); e->num.clear();
  for (int ti = 0; ti < 2; ++ti) {
    Tower& t = e->t[ti];
    for (int i = 0; i < t.nconv; ++i) {
      const ConvSpec& cs = t.conv[i];
      t.w_off[i] = o; o += pad8l((long long)cs.cout * cs.cin * cs.k * cs.k);
      t.b_off[i] = o; o += pad8l(cs.cout);
    }
  }
  e->fc_begin = o;
  for (int ti = 0; ti < 2; ++ti) {
    Tower& t = e->t[ti];
    t.w_off[3] = o; o += pad8l((long long)t.feat * t.flat);
    t.b_off[3] = o; o += pad8l(t.feat);
  }
  for (int ti = 0; ti < 2; ++ti) {
    const Tower& t = e->t[ti];
    for (int i = 0; i < t.nconv; ++i) {
      const ConvSpec& cs = t.conv[i];
      e->off.push_back(t.w_off[i]); e->num.push_back((long long)cs.cout * cs.cin * cs.k * cs.k);
      e->off.push_back(t.b_off[i]); e->num.push_back(cs.cout);
    }
    e->off.push_back(t.w_off[3]); e->num.push_back((long long)t.feat * t.flat);
    e->off.push_back(t.b_off[3]); e->num.push_back(t.feat);
  }
  e->total = o;
  e->n_entries = (int)e->off.size();
  for (int ti = 0; ti < 2; ++ti) {
    Tower& t = e->t[ti];
    for (int i = 0; i < 3; ++i) { t.implicit[i] = false; t.dg_off[i] = 0; }
    for (int i = 1; i < t.nconv; ++i) {
      const ConvSpec& cs = t.conv[i];
      if (!convn_supported(cs.cin, cs.cout, cs.k, cs.stride)) continue;
      t.implicit[i] = true;
      t.dg_off[i] = o;
      o += pad8l((long long)convn_dgrad_weight_elems(cs.cin, cs.cout, cs.k));
    }
  }
  e->staged_total = o;
}

template <typename T>
static T* carve(uint8_t*& p, size_t count) {
  T* r = reinterpret_cast<T*>(p);
  p += al256(count * sizeof(T));
  return r;
}

static void carve_ws(const dtb_encoder* e, int N, uint8_t* base, EncWs& w) {
  uint8_t* p = base;
  size_t pf = 0, maxn = 8;
  auto need = [&](size_t Mo, size_t No, size_t Kr) {
    const size_t s = (size_t)gemm_wgrad_split((int)Mo, (int)No, (int)Kr, (size_t)-1);
    if (s > 1 && s * Mo * No > pf) pf = s * Mo * No;
  };
  for (int ti = 0; ti < 2; ++ti) {
    const Tower& t = e->t[ti];
    TowerWs& tw = w.t[ti];
    size_t max_dp = 0, max_g = (size_t)t.feat * t.flat;
    for (int i = 0; i < 3; ++i) tw.P[i] = tw.Y[i] = tw.dZ[i] = nullptr;
    size_t cn_pf = 0;
    for (int i = 0; i < t.nconv; ++i) {
      const ConvSpec& cs = t.conv[i];
      const size_t rows = (size_t)N * cs.hout * cs.wout, kc = (size_t)cs.cin * cs.k * cs.k;
      const bool implicit = i == 0 ? conv1_implicit_supported(cs.cin, cs.win, cs.stride, cs.cout) : t.implicit[i];
      tw.P[i] = implicit ? nullptr : carve<bf16>(p, rows * kc);  // patch matrices exist only on the explicit (im2col) path
      if (i == 0) tw.c1_partial = implicit ? carve<float>(p, conv1_wgrad_partial_floats(cs.cin, cs.cout)) : nullptr;
      tw.Y[i] = carve<bf16>(p, rows * cs.cout);
      tw.dZ[i] = carve<bf16>(p, rows * cs.cout);
      if (i > 0 && !implicit && rows * kc > max_dp) max_dp = rows * kc;
      if ((size_t)cs.cout * kc > max_g) max_g = (size_t)cs.cout * kc;
      if (!implicit) need(cs.cout, kc, rows);
      if (i > 0 && implicit && convn_wgrad_partial_floats(cs.cin, cs.k) > cn_pf) cn_pf = convn_wgrad_partial_floats(cs.cin, cs.k);
      if ((size_t)cs.cout > maxn) maxn = cs.cout;
    }
    need(t.feat, t.flat, N);
    if ((size_t)t.feat > maxn) maxn = t.feat;
    tw.dP = carve<bf16>(p, max_dp);
    tw.gstage = carve<float>(p, max_g);
    tw.cn_partial = carve<float>(p, cn_pf);
  }
  const bool sum = e->c.fusion == 1;
  w.feat[0] = sum ? carve<bf16>(p, (size_t)N * e->t[0].feat) : nullptr;
  w.feat[1] = sum ? carve<bf16>(p, (size_t)N * e->t[1].feat) : nullptr;
  w.dfeat = sum ? carve<bf16>(p, (size_t)N * e->t[0].feat) : nullptr;
  w.partial_floats = pf;
  for (int ti = 0; ti < 2; ++ti) {
    w.partial[ti] = carve<float>(p, pf);
    w.cs_scratch[ti] = carve<float>(p, colsum_scratch_floats((int)maxn));
  }
  w.total = (size_t)(p - base);
}

static GemmEpilogue epz() {
  GemmEpilogue ep;
  memset(&ep, 0, sizeof(ep));
  ep.alpha = 1.f;
  return ep;
}
static int gemm_kk(cudaStream_t st, const bf16* A, int lda, const bf16* B, int ldb, int b_mn, int M, int N, int K, const GemmEpilogue& ep) {
  const int rc = gemm_bf16(st, A, lda, 0, B, ldb, b_mn, M, N, K, ep, 1, 0);
  return rc < 0 ? rc : DTB_OK;
}

static int tower_forward(const dtb_encoder* e, int ti, cudaStream_t st, const float* P, const bf16* Ps, int N, const void* x, int x_u8,
                         TowerWs& tw, bf16* out, int ldo) {
  const Tower& t = e->t[ti];
  for (int i = 0; i < t.nconv; ++i) {
    const ConvSpec& cs = t.conv[i];
    const int rows = N * cs.hout * cs.wout, kc = cs.cin * cs.k * cs.k;
    if (i == 0 && tw.P[0] == nullptr) {  // implicit GEMM: patches gathered inside the kernel, never written to HBM
      DTB_TRY(conv1_fwd(st, x, x_u8, N, cs.cin, cs.hin, cs.win, cs.stride, e->c.input_scale, Ps + t.w_off[0], P + t.b_off[0], cs.cout, tw.Y[0]));
      continue;
    }
    if (i > 0 && t.implicit[i]) {  // implicit GEMM over the previous layer's NHWC activation
      DTB_TRY(convn_fwd(st, tw.Y[i - 1], N, cs.hin, cs.win, cs.cin, cs.k, cs.stride, Ps + t.w_off[i], P + t.b_off[i], cs.cout, tw.Y[i]));
      continue;
    }
    if (i == 0) DTB_REQUIRE(!x_u8, DTB_ERR_UNSUPPORTED, "uint8 pixel input needs the implicit-GEMM conv1 (W %% 4 == 0, stride 4, C <= 4)");
    if (i == 0) DTB_TRY(im2col_nchw_k8(st, reinterpret_cast<const float*>(x), N, cs.cin, cs.hin, cs.win, cs.stride, e->c.input_scale, tw.P[0]));
    else DTB_TRY(im2col_nhwc(st, tw.Y[i - 1], N, cs.hin, cs.win, cs.cin, cs.k, cs.stride, tw.P[i]));
    GemmEpilogue ep = epz();
    ep.bias = P + t.b_off[i];
    ep.act = 1;
    ep.out_bf16 = tw.Y[i];
    ep.ldo_b = cs.cout;
    DTB_TRY(gemm_kk(st, tw.P[i], kc, Ps + t.w_off[i], kc, 0, rows, cs.cout, kc, ep));
  }
  GemmEpilogue ep = epz();
  ep.bias = P + t.b_off[3];
  ep.act = 1;
  ep.out_bf16 = out;
  ep.ldo_b = ldo;
  return gemm_kk(st, tw.Y[t.nconv - 1], t.flat, Ps + t.w_off[3], t.flat, 0, N, t.feat, t.flat, ep);
}

// dF: gradient w.r.t. the tower's feature output, already multiplied by the ReLU mask; bf16 [N, feat] with row stride ldf.
static int tower_backward(const dtb_encoder* e, int ti, cudaStream_t st, const bf16* Ps, int N, const void* x, int x_u8, const bf16* dF,
                          int ldf, TowerWs& tw, EncWs& w, float* G, SideStream& wside) {
  const Tower& t = e->t[ti];
  const int last = t.nconv - 1;
  const ConvSpec& cl = t.conv[last];
  const int hw = cl.hout * cl.wout;
  SideStream& sd = wside;
  cudaStream_t ss = sd.pick(st);  // weight / bias gradients: side branch; data gradients stay on the caller's stream
  float* const partial = w.partial[ti];
  float* const cs_scratch = w.cs_scratch[ti];
  // ---- FC ----
  DTB_TRY(sd.fork(st));
  DTB_TRY(gemm_wgrad(ss, dF, ldf, tw.Y[last], t.flat, t.feat, t.flat, N, tw.gstage, t.flat, partial, w.partial_floats));
  DTB_TRY(permute_unstage(ss, tw.gstage, t.feat, cl.cout, hw, G + t.w_off[3]));  // staged (h,w,c) -> PyTorch flatten (c,h,w)
  DTB_TRY(colsum_bf16(ss, dF, N, t.feat, ldf, G + t.b_off[3], 0, cs_scratch));
  DTB_CUDA(cudaEventRecord(e->fc_ev[ti], ss));  // this tower's FC gradients are final from here on (dtb_encoder_wait_fc_grads)
  {
    GemmEpilogue ep = epz();
    ep.dact = 1;
    ep.aux = tw.Y[last];
    ep.ldaux = t.flat;
    ep.out_bf16 = tw.dZ[last];
    ep.ldo_b = t.flat;
    DTB_TRY(gemm_kk(st, dF, ldf, Ps + t.w_off[3], t.flat, 1, N, t.flat, t.feat, ep));
  }
  // ---- convs, last to first ----
  for (int i = last; i >= 0; --i) {
    const ConvSpec& cs = t.conv[i];
    const int rows = N * cs.hout * cs.wout, kc = cs.cin * cs.k * cs.k;
    DTB_TRY(sd.fork(st));  // dZ[i] is complete on the main stream
    if (i > 0 && t.implicit[i]) {
      // weight + bias gradient (side branch): patches gathered again from Y[i-1]; written straight in PyTorch order
      DTB_TRY(convn_wgrad(ss, tw.Y[i - 1], N, cs.hin, cs.win, cs.cin, cs.k, cs.stride, tw.dZ[i], cs.cout, tw.cn_partial, G + t.w_off[i],
                          G + t.b_off[i]));
      // data gradient (main chain): implicit transposed convolution with the ReLU mask of Y[i-1] in the epilogue
      DTB_TRY(convn_dgrad(st, tw.dZ[i], N, cs.hin, cs.win, cs.cin, cs.k, cs.stride, Ps + t.dg_off[i], cs.cout, tw.Y[i - 1], tw.dZ[i - 1]));
      continue;
    }
    if (i == 0 && tw.P[0] == nullptr) {
      DTB_REQUIRE(x != nullptr, DTB_ERR_BAD_ARG, "encoder backward: frames / audio are needed again for the implicit conv1 weight gradient");
      DTB_TRY(conv1_wgrad(ss, x, x_u8, N, cs.cin, cs.hin, cs.win, cs.stride, e->c.input_scale, tw.dZ[0], cs.cout, tw.c1_partial, G + t.w_off[0],
                          G + t.b_off[0]));  // weight and bias gradient in one pass
      continue;
    } else if (i == 0) {
      DTB_TRY(gemm_wgrad(ss, tw.dZ[0], cs.cout, tw.P[0], kc, cs.cout, kc, rows, G + t.w_off[0], kc, partial, w.partial_floats));
    } else {
      DTB_TRY(gemm_wgrad(ss, tw.dZ[i], cs.cout, tw.P[i], kc, cs.cout, kc, rows, tw.gstage, kc, partial, w.partial_floats));
      DTB_TRY(permute_unstage(ss, tw.gstage, cs.cout, cs.cin, cs.k * cs.k, G + t.w_off[i]));  // (kh,kw,c) -> (c,kh,kw)
    }
    DTB_TRY(colsum_bf16(ss, tw.dZ[i], rows, cs.cout, cs.cout, G + t.b_off[i], 0, cs_scratch));
    if (i > 0) {
      GemmEpilogue ep = epz();
      ep.out_bf16 = tw.dP;
      ep.ldo_b = kc;
      DTB_TRY(gemm_kk(st, tw.dZ[i], cs.cout, Ps + t.w_off[i], kc, 1, rows, kc, cs.cout, ep));  // dP = dZ W (W [Cout][kc] as MN-major B)
      DTB_TRY(col2im_nhwc_relu(st, tw.dP, tw.Y[i - 1], N, cs.hin, cs.win, cs.cin, cs.k, cs.stride, tw.dZ[i - 1]));
    }
  }
  return DTB_OK;
}

}  // namespace dtb

extern "C" {

int dtb_encoder_create(const dtb_encoder_config* cfg, dtb_encoder** out) {
  DTB_REQUIRE(cfg && out, DTB_ERR_BAD_ARG, "dtb_encoder_create: null argument");
  DTB_REQUIRE(cfg->fusion == 0 || cfg->fusion == 1, DTB_ERR_BAD_ARG, "encoder: fusion must be 0 (concat) or 1 (sum)");
  DTB_REQUIRE(cfg->video.big == cfg->audio.big, DTB_ERR_UNSUPPORTED, "encoder: both towers use the same conv_type");
  dtb_encoder* e = new (std::nothrow) dtb_encoder();
  DTB_REQUIRE(e != nullptr, DTB_ERR_BAD_ARG, "dtb_encoder_create: out of memory");
  e->c = *cfg;
  int rc = build_tower(e->t[0], cfg->video);
  if (rc == DTB_OK) rc = build_tower(e->t[1], cfg->audio);
  if (rc != DTB_OK) { delete e; return rc; }
  e->state_dim = cfg->fusion == 0 ? e->t[0].feat + e->t[1].feat : e->t[0].feat;
  layout(e);
  *out = e;
  return DTB_OK;
}
int dtb_encoder_destroy(dtb_encoder* e) {
  if (e) {
    e->side.destroy(); e->side2.destroy(); e->side3.destroy();
    for (int ti = 0; ti < 2; ++ti)
      if (e->fc_ev[ti]) cudaEventDestroy(e->fc_ev[ti]);
  }
  delete e;
  return DTB_OK;
}
int dtb_encoder_state_dim(const dtb_encoder* e) { return e ? e->state_dim : DTB_ERR_BAD_ARG; }
int dtb_encoder_param_entries(const dtb_encoder* e) { return e ? e->n_entries : DTB_ERR_BAD_ARG; }
long long dtb_encoder_param_layout(const dtb_encoder* e, long long* offsets, long long* numels) {
  if (!e) return DTB_ERR_BAD_ARG;
  for (int i = 0; i < e->n_entries; ++i) {
    if (offsets) offsets[i] = e->off[i];
    if (numels) numels[i] = e->num[i];
  }
  return e->total;
}
int dtb_encoder_fc_grad_range(const dtb_encoder* e, long long* begin, long long* end) {
  DTB_REQUIRE(e && begin && end, DTB_ERR_BAD_ARG, "dtb_encoder_fc_grad_range: null argument");
  *begin = e->fc_begin;
  *end = e->total;
  return DTB_OK;
}
int dtb_encoder_wait_fc_grads(const dtb_encoder* e, void* stream) {
  DTB_REQUIRE(e, DTB_ERR_BAD_ARG, "dtb_encoder_wait_fc_grads: null handle");
  DTB_REQUIRE(e->fc_ev[0] && e->fc_ev[1], DTB_ERR_BAD_ARG, "dtb_encoder_wait_fc_grads: no dtb_encoder_backward call has been made on this handle");
  for (int ti = 0; ti < 2; ++ti) DTB_CUDA(cudaStreamWaitEvent((cudaStream_t)stream, e->fc_ev[ti], 0));
  return DTB_OK;
}
long long dtb_encoder_staged_elems(const dtb_encoder* e) { return e ? e->staged_total : DTB_ERR_BAD_ARG; }
size_t dtb_encoder_workspace_bytes(const dtb_encoder* e, int N) {
  if (!e || N < 0) return 0;
  EncWs w;
  carve_ws(e, N < 1 ? 1 : N, nullptr, w);
  return w.total + 256;
}

int dtb_encoder_stage_params(const dtb_encoder* e, void* stream, const float* params, void* staged) {
  DTB_REQUIRE(e && params && staged, DTB_ERR_BAD_ARG, "dtb_encoder_stage_params: null argument");
  cudaStream_t st = (cudaStream_t)stream;
  bf16* S = (bf16*)staged;
  for (int ti = 0; ti < 2; ++ti) {
    const Tower& t = e->t[ti];
    for (int i = 0; i < t.nconv; ++i) {
      const ConvSpec& cs = t.conv[i];
      if (i == 0) DTB_TRY(cast_bf16(st, params + t.w_off[0], (size_t)cs.cout * cs.cin * cs.k * cs.k, S + t.w_off[0]));
      else DTB_TRY(permute_stage(st, params + t.w_off[i], cs.cout, cs.cin, cs.k * cs.k, S + t.w_off[i]));
      if (i > 0 && t.implicit[i]) DTB_TRY(convn_stage_dgrad(st, params + t.w_off[i], cs.cin, cs.cout, cs.k, cs.stride, S + t.dg_off[i]));
    }
    const ConvSpec& cl = t.conv[t.nconv - 1];
    DTB_TRY(permute_stage(st, params + t.w_off[3], t.feat, cl.cout, cl.hout * cl.wout, S + t.w_off[3]));
  }
  return DTB_OK;
}

static int enc_ws(const dtb_encoder* e, int N, void* workspace, size_t bytes, EncWs& w) {
  DTB_REQUIRE(workspace != nullptr, DTB_ERR_WORKSPACE, "encoder: null workspace");
  uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(workspace) + 255) & ~uintptr_t(255));
  carve_ws(e, N < 1 ? 1 : N, base, w);
  DTB_REQUIRE((size_t)(base - reinterpret_cast<uint8_t*>(workspace)) + w.total <= bytes, DTB_ERR_WORKSPACE,
              "encoder: workspace too small (%zu bytes given, %zu needed)", bytes, w.total + 256);
  return DTB_OK;
}

static int encoder_forward_impl(const dtb_encoder* e, void* stream, const float* params, const void* staged, int N, const void* frames,
                                int frames_u8, const float* audio, void* workspace, size_t workspace_bytes, void* states_bf16) {
  DTB_REQUIRE(e && params && staged && frames && audio && states_bf16, DTB_ERR_BAD_ARG, "dtb_encoder_forward: null argument");
  DTB_TRY(require_sm100());
  if (N == 0) return DTB_OK;
  cudaStream_t st = (cudaStream_t)stream;
  EncWs w;
  DTB_TRY(enc_ws(e, N, workspace, workspace_bytes, w));
  DTB_TRY(e->side.init());
  const bf16* Ps = (const bf16*)staged;
  bf16* out = (bf16*)states_bf16;
  if (e->c.fusion == 0) {  // concat: each tower's FC epilogue writes its half of [N, state_dim]  (agents.py:137-139, dim=1 semantics)
    // the two towers are independent: audio runs on the side branch, video on the caller's stream
    DTB_TRY(e->side.fork(st));
    DTB_TRY(tower_forward(e, 1, e->side.pick(st), params, Ps, N, audio, 0, w.t[1], out + e->t[0].feat, e->state_dim));
    DTB_TRY(tower_forward(e, 0, st, params, Ps, N, frames, frames_u8, w.t[0], out, e->state_dim));
    DTB_TRY(e->side.join(st));
  } else {  // sum (agents.py:140-141)
    DTB_TRY(e->side.fork(st));
    DTB_TRY(tower_forward(e, 1, e->side.pick(st), params, Ps, N, audio, 0, w.t[1], w.feat[1], e->t[1].feat));
    DTB_TRY(tower_forward(e, 0, st, params, Ps, N, frames, frames_u8, w.t[0], w.feat[0], e->t[0].feat));
    DTB_TRY(e->side.join(st));
    DTB_TRY(add_bf16(st, w.feat[0], w.feat[1], (size_t)N * e->t[0].feat, out));
  }
  return DTB_OK;
}

static int encoder_backward_impl(const dtb_encoder* e, void* stream, const float* params, const void* staged, int N, const void* frames,
                                 int frames_u8, const float* audio, const void* states_bf16, const void* d_states_bf16, int d_states_masked,
                                 void* workspace, size_t workspace_bytes, float* grads) {
  DTB_REQUIRE(e && params && staged && d_states_bf16 && grads, DTB_ERR_BAD_ARG, "dtb_encoder_backward: null argument");
  cudaStream_t st = (cudaStream_t)stream;
  for (int ti = 0; ti < 2; ++ti)
    if (!e->fc_ev[ti]) DTB_CUDA(cudaEventCreateWithFlags(&e->fc_ev[ti], cudaEventDisableTiming));
  if (N == 0) {
    DTB_CUDA(cudaMemsetAsync(grads, 0, (size_t)e->total * sizeof(float), st));
    for (int ti = 0; ti < 2; ++ti) DTB_CUDA(cudaEventRecord(e->fc_ev[ti], st));
    return DTB_OK;
  }
  EncWs w;
  DTB_TRY(enc_ws(e, N, workspace, workspace_bytes, w));
  DTB_TRY(e->side.init());
  const bf16* Ps = (const bf16*)staged;
  const bf16* dS = (const bf16*)d_states_bf16;
  if (e->c.fusion == 0) {
    // concat: the wide state IS [relu(video) | relu(audio)], so a gradient masked by (state > 0) is already dZ of both FCs
    DTB_REQUIRE(d_states_masked || states_bf16, DTB_ERR_BAD_ARG, "encoder backward: states needed to apply the ReLU mask");
    bf16* d = const_cast<bf16*>(dS);
    if (!d_states_masked)
      DTB_TRY(relu_mask_bf16(st, dS, e->state_dim, (const bf16*)states_bf16, e->state_dim, N, e->state_dim, d, e->state_dim));
    // the towers are independent: the audio tower's dgrad chain runs on its own stream next to the video tower's; both feed their
    // weight gradients to the shared side stream (which serialises the shared split-K / column-sum scratch)
    DTB_TRY(e->side2.init());
    DTB_TRY(e->side3.init());
    DTB_TRY(e->side2.fork(st));
    DTB_TRY(tower_backward(e, 1, e->side2.pick(st), Ps, N, audio, 0, d + e->t[0].feat, e->state_dim, w.t[1], w, grads, e->side3));
    DTB_TRY(tower_backward(e, 0, st, Ps, N, frames, frames_u8, d, e->state_dim, w.t[0], w, grads, e->side));
    DTB_TRY(e->side3.join(e->side2.pick(st)));
    DTB_TRY(e->side2.join(st));
  } else {
    // sum: the gradient reaches both towers unchanged; each applies its own ReLU mask
    const int F = e->t[0].feat;
    for (int ti = 0; ti < 2; ++ti) {
      DTB_TRY(relu_mask_bf16(st, dS, F, w.feat[ti], F, N, F, w.dfeat, F));
      DTB_TRY(tower_backward(e, ti, st, Ps, N, ti == 0 ? frames : (const void*)audio, ti == 0 ? frames_u8 : 0, w.dfeat, F, w.t[ti], w, grads,
                             e->side));
      DTB_TRY(e->side.join(st));  // w.dfeat is reused by the next tower
    }
  }
  return e->side.join(st);
}

int dtb_encoder_forward(const dtb_encoder* e, void* stream, const float* params, const void* staged, int N, const float* frames,
                        const float* audio, void* workspace, size_t workspace_bytes, void* states_bf16) {
  return encoder_forward_impl(e, stream, params, staged, N, frames, 0, audio, workspace, workspace_bytes, states_bf16);
}
int dtb_encoder_forward_u8(const dtb_encoder* e, void* stream, const float* params, const void* staged, int N, const unsigned char* frames_u8,
                           const float* audio, void* workspace, size_t workspace_bytes, void* states_bf16) {
  return encoder_forward_impl(e, stream, params, staged, N, frames_u8, 1, audio, workspace, workspace_bytes, states_bf16);
}
int dtb_encoder_backward(const dtb_encoder* e, void* stream, const float* params, const void* staged, int N, const float* frames,
                         const float* audio, const void* states_bf16, const void* d_states_bf16, int d_states_masked, void* workspace,
                         size_t workspace_bytes, float* grads) {
  return encoder_backward_impl(e, stream, params, staged, N, frames, 0, audio, states_bf16, d_states_bf16, d_states_masked, workspace,
                               workspace_bytes, grads);
}
int dtb_encoder_backward_u8(const dtb_encoder* e, void* stream, const float* params, const void* staged, int N,
                            const unsigned char* frames_u8, const float* audio, const void* states_bf16, const void* d_states_bf16,
                            int d_states_masked, void* workspace, size_t workspace_bytes, float* grads) {
  return encoder_backward_impl(e, stream, params, staged, N, frames_u8, 1, audio, states_bf16, d_states_bf16, d_states_masked, workspace,
                               workspace_bytes, grads);
}

}  // extern "C"
