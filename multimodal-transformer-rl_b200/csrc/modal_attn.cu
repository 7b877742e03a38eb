// Modality attention over the two tower features (REF src/agents.py:58-75 CaslAttention, :78-90 NewAttention).
//   casl: z = tanh(video_fc(v) + audio_fc(a) + state_fc(h))            [32]
//   new : z = tanh(fc([v | a | h]))                                    [128]
//   w = softmax(attention(z)) [2];  v' = w0 * v, a' = w1 * a;  fusion: concat [v' | a'] or sum v' + a' (agents.py:137-143)
// Both are one linear map of the row x = [v | a | h] (K = 2F + 128): for casl the three weight matrices sit side by side along K and
// the three biases add.  The linear map and its data / weight gradients run on the tcgen05 GEMM (features are the bf16 tower output,
// read in place); the per-sample tail (tanh, 2-way softmax, scaling, fusion) and its backward are one warp per sample.
// h is the recurrent hidden state of the reference's LSTM agent (128 wide); the Decision Transformer has none and passes NULL (zeros:
// only state_fc.bias / the fc bias contribute), a caller that has one passes fp32 [N, 128].
#include "../../include/dtb200.h"
#include "kernels.h"

#include <string.h>

namespace dtb {
int require_sm100();

namespace {

constexpr int HID = 128;  // LSTM hidden width of the reference agent (agents.py:62, 81, 121)

struct MaLayout {
  int type, F, fusion, J, Kc;
  // flat fp32 parameter offsets (8-element padded), PyTorch layouts
  long long w_off[3], b_off[3];  // casl: audio_fc, video_fc, state_fc ([32,F], [32,F], [32,128]); new: fc [128, 2F+128] in slot 0
  long long wat_off, bat_off;    // attention.weight [2, J], attention.bias [2]
  long long total;
  int n_entries;
  long long off[8], num[8];
};

inline long long pad8(long long n) { return (n + 7) & ~7LL; }
inline size_t al256(size_t n) { return (n + 255) & ~size_t(255); }

int make_layout(const dtb_modal_attn_config* c, MaLayout& L) {
  DTB_REQUIRE(c != nullptr, DTB_ERR_BAD_ARG, "modal_attn: null config");
  DTB_REQUIRE(c->type == 0 || c->type == 1, DTB_ERR_BAD_ARG, "modal_attn: type must be 0 (casl) or 1 (new)");
  DTB_REQUIRE(c->fusion == 0 || c->fusion == 1, DTB_ERR_BAD_ARG, "modal_attn: fusion must be 0 (concat) or 1 (sum)");
  DTB_REQUIRE(c->feat >= 64 && c->feat <= 1024 && c->feat % 64 == 0, DTB_ERR_UNSUPPORTED, "modal_attn: feature size %d unsupported", c->feat);
  memset(&L, 0, sizeof(L));
  L.type = c->type; L.F = c->feat; L.fusion = c->fusion;
  L.J = c->type == 0 ? 32 : 128;
  L.Kc = 2 * L.F + HID;
  long long o = 0;
  int n = 0;
  auto add = [&](long long cnt) { L.off[n] = o; L.num[n] = cnt; ++n; const long long at = o; o += pad8(cnt); return at; };
  if (c->type == 0) {  // module order of CaslAttention.__init__: audio_fc, video_fc, state_fc, attention
    L.w_off[0] = add((long long)32 * L.F); L.b_off[0] = add(32);
    L.w_off[1] = add((long long)32 * L.F); L.b_off[1] = add(32);
    L.w_off[2] = add((long long)32 * HID); L.b_off[2] = add(32);
  } else {
    L.w_off[0] = add((long long)128 * L.Kc); L.b_off[0] = add(128);
  }
  L.wat_off = add(2LL * L.J);
  L.bat_off = add(2);
  L.total = o;
  L.n_entries = n;
  return DTB_OK;
}

constexpr int RED_SLICES = 64;  // sample slices of the attention.weight gradient reduction (stage 1), summed in index order (stage 2)

struct MaWs {
  bf16* Wc;       // [J, Kc] staged weights, K order [video | audio | hidden]
  float* bc;      // [J] combined bias
  bf16* hid;      // [N, 128] hidden as a GEMM operand
  float* pre1;    // [N, J] feature part of the pre-activation (+ bias)
  float* z;       // [N, J] tanh output (forward), kept for backward
  float* w;       // [N, 2] attention weights
  bf16* dpre;     // [N, J]
  float* dl;      // [N, 2] logit gradients
  float* R;       // [N, 2F] direct feature gradient (w_i * dS, ReLU-masked): residual of the data-gradient GEMM
  float* red;     // [RED_SLICES, 2, J + 1]
  float* partial; size_t partial_floats;
  float* cs_scratch;
  size_t total;
};

template <typename T>
T* carve(uint8_t*& p, size_t count) {
  T* r = reinterpret_cast<T*>(p);
  p += al256(count * sizeof(T));
  return r;
}

void carve_ws(const MaLayout& L, int N, uint8_t* base, MaWs& w) {
  uint8_t* p = base;
  const size_t n = (size_t)(N < 1 ? 1 : N);
  w.Wc = carve<bf16>(p, (size_t)L.J * L.Kc);
  w.bc = carve<float>(p, L.J);
  w.hid = carve<bf16>(p, n * HID);
  w.pre1 = carve<float>(p, n * L.J);
  w.z = carve<float>(p, n * L.J);
  w.w = carve<float>(p, n * 2);
  w.dpre = carve<bf16>(p, n * L.J);
  w.dl = carve<float>(p, n * 2);
  w.R = carve<float>(p, n * 2 * L.F);
  w.red = carve<float>(p, (size_t)RED_SLICES * 2 * (L.J + 1));
  size_t pf = 0;
  const int nouts[3] = {L.type == 0 ? L.F : 2 * L.F, HID, 0};
  for (int i = 0; i < 2; ++i) {
    const size_t s = (size_t)gemm_wgrad_split(L.J, nouts[i], (int)n, (size_t)-1);
    if (s > 1 && s * L.J * nouts[i] > pf) pf = s * (size_t)L.J * nouts[i];
  }
  w.partial_floats = pf;
  w.partial = carve<float>(p, pf);
  w.cs_scratch = carve<float>(p, colsum_scratch_floats(L.J));
  w.total = (size_t)(p - base);
}

int get_ws(const MaLayout& L, int N, void* workspace, size_t bytes, MaWs& w) {
  DTB_REQUIRE(workspace != nullptr, DTB_ERR_WORKSPACE, "modal_attn: null workspace");
  uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(workspace) + 255) & ~uintptr_t(255));
  carve_ws(L, N, base, w);
  DTB_REQUIRE((size_t)(base - reinterpret_cast<uint8_t*>(workspace)) + w.total <= bytes, DTB_ERR_WORKSPACE,
              "modal_attn: workspace too small (%zu bytes given, %zu needed)", bytes, w.total + 256);
  return DTB_OK;
}

// ---- staging: fp32 parameters -> bf16 [J, Kc] in K order [video | audio | hidden] + combined bias; hidden fp32 -> bf16 ----------
// casl: Wc[j, 0:F] = video_fc.weight[j], Wc[j, F:2F] = audio_fc.weight[j], Wc[j, 2F:] = state_fc.weight[j]; new: a plain cast.
__global__ void __launch_bounds__(256) ma_stage_kernel(const float* __restrict__ P, MaLayout L, bf16* __restrict__ Wc,
                                                        float* __restrict__ bc, const float* __restrict__ hidden, long long n_hid,
                                                        bf16* __restrict__ hid) {
  pdl_trigger();
  pdl_wait();
  const long long nW = (long long)L.J * L.Kc;
  const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x, nth = (long long)gridDim.x * blockDim.x;
  for (long long i = tid; i < nW; i += nth) {
    const int j = (int)(i / L.Kc), k = (int)(i % L.Kc);
    float v;
    if (L.type == 1) v = P[L.w_off[0] + i];
    else if (k < L.F) v = P[L.w_off[1] + (long long)j * L.F + k];
    else if (k < 2 * L.F) v = P[L.w_off[0] + (long long)j * L.F + (k - L.F)];
    else v = P[L.w_off[2] + (long long)j * HID + (k - 2 * L.F)];
    Wc[i] = __float2bfloat16(v);
  }
  for (long long i = tid; i < L.J; i += nth)
    bc[i] = L.type == 1 ? P[L.b_off[0] + i] : (P[L.b_off[1] + i] + P[L.b_off[0] + i]) + P[L.b_off[2] + i];  // (video + audio) + state
  if (hidden)
    for (long long i = tid; i < n_hid; i += nth) hid[i] = __float2bfloat16(hidden[i]);
}

// ---- forward tail: one warp per sample -------------------------------------------------------------------------------------------
// `pre` and `z_out` may be the same buffer (each element is read, then written, by the same thread): no __restrict__ on those two.
template <int J>
__global__ void __launch_bounds__(256) ma_fwd_kernel(const float* pre, const float* __restrict__ Wat,
                                                      const float* __restrict__ bat, const bf16* __restrict__ feat, int N, int F,
                                                      int fusion, float* z_out, float* __restrict__ w_ws,
                                                      float* __restrict__ w_out, bf16* __restrict__ states) {
  pdl_trigger();
  pdl_wait();
  constexpr int PER = J / 32;
  const int lane = threadIdx.x & 31;
  const int warps = (gridDim.x * blockDim.x) >> 5;
  float wa0[PER], wa1[PER];
#pragma unroll
  for (int i = 0; i < PER; ++i) {
    wa0[i] = Wat[lane + 32 * i];
    wa1[i] = Wat[J + lane + 32 * i];
  }
  const float b0 = bat[0], b1 = bat[1];
  for (int n = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; n < N; n += warps) {
    float l0 = 0.f, l1 = 0.f;
#pragma unroll
    for (int i = 0; i < PER; ++i) {
      const float z = tanhf(pre[(size_t)n * J + lane + 32 * i]);
      z_out[(size_t)n * J + lane + 32 * i] = z;
      l0 = fmaf(wa0[i], z, l0);
      l1 = fmaf(wa1[i], z, l1);
    }
    l0 = warp_sum(l0) + b0;
    l1 = warp_sum(l1) + b1;
    const float m = fmaxf(l0, l1);
    const float e0 = expf(l0 - m), e1 = expf(l1 - m);
    const float inv = 1.f / (e0 + e1);
    const float w0 = e0 * inv, w1 = e1 * inv;
    if (lane == 0) {
      w_ws[2 * n] = w0; w_ws[2 * n + 1] = w1;
      if (w_out) { w_out[2 * n] = w0; w_out[2 * n + 1] = w1; }
    }
    const bf16* fv = feat + (size_t)n * 2 * F;
    const bf16* fa = fv + F;
    for (int c = lane * 8; c < F; c += 256) {  // 16-byte chunks
      const uint4 v = ldg_u4(fv + c), a = ldg_u4(fa + c);
      const uint32_t vv[4] = {v.x, v.y, v.z, v.w}, aa[4] = {a.x, a.y, a.z, a.w};
      uint32_t ov[4], oa[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float2 x = unpack_bf16x2(vv[q]), y = unpack_bf16x2(aa[q]);
        if (fusion == 0) {
          ov[q] = pack_bf16x2(w0 * x.x, w0 * x.y);
          oa[q] = pack_bf16x2(w1 * y.x, w1 * y.y);
        } else {
          ov[q] = pack_bf16x2(fmaf(w0, x.x, w1 * y.x), fmaf(w0, x.y, w1 * y.y));
        }
      }
      if (fusion == 0) {
        *reinterpret_cast<uint4*>(states + (size_t)n * 2 * F + c) = make_uint4(ov[0], ov[1], ov[2], ov[3]);
        *reinterpret_cast<uint4*>(states + (size_t)n * 2 * F + F + c) = make_uint4(oa[0], oa[1], oa[2], oa[3]);
      } else {
        *reinterpret_cast<uint4*>(states + (size_t)n * F + c) = make_uint4(ov[0], ov[1], ov[2], ov[3]);
      }
    }
  }
}

// ---- backward tail: one warp per sample ------------------------------------------------------------------------------------------
// g_i = <dS_i, feat_i>;  dl_i = w_i (g_i - (w0 g0 + w1 g1));  dz = Wat^T dl;  dpre = dz (1 - z^2);  R = relu'(feat) * w_i * dS_i
template <int J>
__global__ void __launch_bounds__(256) ma_bwd_kernel(const bf16* __restrict__ dS, const bf16* __restrict__ feat,
                                                      const float* __restrict__ z_in, const float* __restrict__ w_ws,
                                                      const float* __restrict__ Wat, int N, int F, int fusion, bf16* __restrict__ dpre,
                                                      float* __restrict__ dl_out, float* __restrict__ R) {
  pdl_trigger();
  pdl_wait();
  constexpr int PER = J / 32;
  const int lane = threadIdx.x & 31;
  const int warps = (gridDim.x * blockDim.x) >> 5;
  float wa0[PER], wa1[PER];
#pragma unroll
  for (int i = 0; i < PER; ++i) {
    wa0[i] = Wat[lane + 32 * i];
    wa1[i] = Wat[J + lane + 32 * i];
  }
  for (int n = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; n < N; n += warps) {
    const float w0 = w_ws[2 * n], w1 = w_ws[2 * n + 1];
    const bf16* fv = feat + (size_t)n * 2 * F;
    const bf16* fa = fv + F;
    const bf16* dv = fusion == 0 ? dS + (size_t)n * 2 * F : dS + (size_t)n * F;
    const bf16* da = fusion == 0 ? dv + F : dv;
    float* Rv = R + (size_t)n * 2 * F;
    float* Ra = Rv + F;
    float g0 = 0.f, g1 = 0.f;
    for (int c = lane * 8; c < F; c += 256) {
      const uint4 v = ldg_u4(fv + c), a = ldg_u4(fa + c), gv = ldg_u4(dv + c), ga = ldg_u4(da + c);
      const uint32_t vv[4] = {v.x, v.y, v.z, v.w}, aa[4] = {a.x, a.y, a.z, a.w};
      const uint32_t gvv[4] = {gv.x, gv.y, gv.z, gv.w}, gaa[4] = {ga.x, ga.y, ga.z, ga.w};
      float rv[8], ra[8];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float2 x = unpack_bf16x2(vv[q]), y = unpack_bf16x2(aa[q]);
        const float2 dx = unpack_bf16x2(gvv[q]), dy = unpack_bf16x2(gaa[q]);
        g0 = fmaf(dx.x, x.x, g0); g0 = fmaf(dx.y, x.y, g0);
        g1 = fmaf(dy.x, y.x, g1); g1 = fmaf(dy.y, y.y, g1);
        rv[2 * q] = x.x > 0.f ? w0 * dx.x : 0.f; rv[2 * q + 1] = x.y > 0.f ? w0 * dx.y : 0.f;
        ra[2 * q] = y.x > 0.f ? w1 * dy.x : 0.f; ra[2 * q + 1] = y.y > 0.f ? w1 * dy.y : 0.f;
      }
      *reinterpret_cast<float4*>(Rv + c) = make_float4(rv[0], rv[1], rv[2], rv[3]);
      *reinterpret_cast<float4*>(Rv + c + 4) = make_float4(rv[4], rv[5], rv[6], rv[7]);
      *reinterpret_cast<float4*>(Ra + c) = make_float4(ra[0], ra[1], ra[2], ra[3]);
      *reinterpret_cast<float4*>(Ra + c + 4) = make_float4(ra[4], ra[5], ra[6], ra[7]);
    }
    g0 = warp_sum(g0);
    g1 = warp_sum(g1);
    const float mean = w0 * g0 + w1 * g1;
    const float dl0 = w0 * (g0 - mean), dl1 = w1 * (g1 - mean);
    if (lane == 0) { dl_out[2 * n] = dl0; dl_out[2 * n + 1] = dl1; }
#pragma unroll
    for (int i = 0; i < PER; ++i) {
      const float z = z_in[(size_t)n * J + lane + 32 * i];
      const float dz = wa0[i] * dl0 + wa1[i] * dl1;
      dpre[(size_t)n * J + lane + 32 * i] = __float2bfloat16(dz * (1.f - z * z));
    }
  }
}

// ---- attention.weight / attention.bias gradient: dWat[i, j] = sum_n dl[n, i] z[n, j], dbat[i] = sum_n dl[n, i] ------------------
// stage 1: grid (J / 32, RED_SLICES); lanes = columns, the 8 warps of a block walk the slice's samples, combined in warp order.
__global__ void __launch_bounds__(256) ma_wat_partial_kernel(const float* __restrict__ dl, const float* __restrict__ z, int N, int J,
                                                              float* __restrict__ red) {
  pdl_trigger();
  pdl_wait();
  __shared__ float sm[8][3][32];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int j = blockIdx.x * 32 + lane;
  const int per = (N + gridDim.y - 1) / gridDim.y;
  const int n0 = blockIdx.y * per, n1 = min(N, n0 + per);
  float a0 = 0.f, a1 = 0.f, c0 = 0.f, c1 = 0.f;
  for (int n = n0 + wid; n < n1; n += 8) {
    const float d0 = dl[2 * n], d1 = dl[2 * n + 1];
    const float zz = z[(size_t)n * J + j];
    a0 = fmaf(d0, zz, a0);
    a1 = fmaf(d1, zz, a1);
    c0 += d0; c1 += d1;
  }
  sm[wid][0][lane] = a0; sm[wid][1][lane] = a1; sm[wid][2][lane] = lane == 0 ? c0 : (lane == 1 ? c1 : 0.f);
  __syncthreads();
  if (wid == 0) {
    float s0 = 0.f, s1 = 0.f, sb = 0.f;
    for (int k = 0; k < 8; ++k) { s0 += sm[k][0][lane]; s1 += sm[k][1][lane]; sb += sm[k][2][lane]; }
    float* r = red + (size_t)blockIdx.y * 2 * (J + 1);
    r[j] = s0;
    r[(J + 1) + j] = s1;
    if (blockIdx.x == 0 && lane < 2) r[lane * (J + 1) + J] = sb;  // bias partial of logit `lane`
  }
}
// stage 2: one thread per output, slices added in index order
__global__ void __launch_bounds__(128) ma_wat_final_kernel(const float* __restrict__ red, int slices, int J, float* __restrict__ dWat,
                                                            float* __restrict__ dbat) {
  pdl_trigger();
  pdl_wait();
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= 2 * (J + 1)) return;
  const int i = t / (J + 1), j = t % (J + 1);
  float s = 0.f;
  for (int k = 0; k < slices; ++k) s += red[(size_t)k * 2 * (J + 1) + t];
  if (j < J) dWat[i * J + j] = s;
  else dbat[i] = s;
}

__global__ void __launch_bounds__(256) ma_copy_kernel(const float* __restrict__ src, int n, float* __restrict__ d0, float* __restrict__ d1) {
  pdl_trigger();
  pdl_wait();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    const float v = src[i];
    d0[i] = v;
    if (d1) d1[i] = v;
  }
}
__global__ void __launch_bounds__(256) ma_zero_kernel(float* __restrict__ p, int n) {
  pdl_trigger();
  pdl_wait();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = 0.f;
}

GemmEpilogue epz() {
  GemmEpilogue ep;
  memset(&ep, 0, sizeof(ep));
  ep.alpha = 1.f;
  return ep;
}

int warp_grid(int N) {
  int blocks = (N + 7) / 8;
  const int cap = num_sms() * 8;
  return blocks < 1 ? 1 : (blocks > cap ? cap : blocks);
}

}  // namespace
}  // namespace dtb

using namespace dtb;

extern "C" {

int dtb_modal_attn_param_entries(const dtb_modal_attn_config* cfg) {
  MaLayout L;
  DTB_TRY(make_layout(cfg, L));
  return L.n_entries;
}
long long dtb_modal_attn_param_layout(const dtb_modal_attn_config* cfg, long long* offsets, long long* numels) {
  MaLayout L;
  DTB_TRY(make_layout(cfg, L));
  for (int i = 0; i < L.n_entries; ++i) {
    if (offsets) offsets[i] = L.off[i];
    if (numels) numels[i] = L.num[i];
  }
  return L.total;
}
int dtb_modal_attn_state_dim(const dtb_modal_attn_config* cfg) {
  MaLayout L;
  DTB_TRY(make_layout(cfg, L));
  return L.fusion == 0 ? 2 * L.F : L.F;
}
size_t dtb_modal_attn_workspace_bytes(const dtb_modal_attn_config* cfg, int N) {
  MaLayout L;
  if (make_layout(cfg, L) != DTB_OK || N < 0) return 0;
  MaWs w;
  carve_ws(L, N, nullptr, w);
  return w.total + 256;
}

int dtb_modal_attn_forward(const dtb_modal_attn_config* cfg, void* stream, const float* params, int N, const void* feat_bf16,
                           const float* hidden, void* workspace, size_t workspace_bytes, void* states_bf16, float* attn_weights) {
  MaLayout L;
  DTB_TRY(make_layout(cfg, L));
  DTB_REQUIRE(params && feat_bf16 && states_bf16, DTB_ERR_BAD_ARG, "dtb_modal_attn_forward: null argument");
  DTB_TRY(require_sm100());
  if (N == 0) return DTB_OK;
  cudaStream_t st = (cudaStream_t)stream;
  MaWs w;
  DTB_TRY(get_ws(L, N, workspace, workspace_bytes, w));
  const bf16* feat = (const bf16*)feat_bf16;
  {
    const long long work = (long long)L.J * L.Kc + (hidden ? (long long)N * HID : 0);
    long long blocks = (work + 255) / 256;
    if (blocks > num_sms() * 8) blocks = num_sms() * 8;
    DTB_LAUNCH(ma_stage_kernel, (unsigned)blocks, 256, 0, st, params, L, w.Wc, w.bc, hidden, (long long)N * HID, w.hid);
    DTB_CHECK_LAUNCH();
  }
  // pre = [v | a] Wc[:, :2F]^T + bc (+ h Wc[:, 2F:]^T)
  float* pre = w.pre1;
  {
    GemmEpilogue ep = epz();
    ep.bias = w.bc;
    ep.out_f32 = w.pre1;
    ep.ldo_f = L.J;
    const int rc = gemm_bf16(st, feat, 2 * L.F, 0, w.Wc, L.Kc, 0, N, L.J, 2 * L.F, ep, 1, 0);
    if (rc < 0) return rc;
  }
  if (hidden) {
    GemmEpilogue ep = epz();
    ep.residual = w.pre1;
    ep.ldr = L.J;
    ep.out_f32 = w.z;  // overwritten in place by the tail kernel (each element is read and written by the same thread)
    ep.ldo_f = L.J;
    const int rc = gemm_bf16(st, w.hid, HID, 0, w.Wc + 2 * L.F, L.Kc, 0, N, L.J, HID, ep, 1, 0);
    if (rc < 0) return rc;
    pre = w.z;
  }
  const float* Wat = params + L.wat_off;
  const float* bat = params + L.bat_off;
  if (L.J == 32)
    DTB_LAUNCH(ma_fwd_kernel<32>, warp_grid(N), 256, 0, st, (const float*)pre, Wat, bat, feat, N, L.F, L.fusion, w.z, w.w, attn_weights,
               (bf16*)states_bf16);
  else
    DTB_LAUNCH(ma_fwd_kernel<128>, warp_grid(N), 256, 0, st, (const float*)pre, Wat, bat, feat, N, L.F, L.fusion, w.z, w.w, attn_weights,
               (bf16*)states_bf16);
  DTB_CHECK_LAUNCH();
  return DTB_OK;
}

int dtb_modal_attn_backward(const dtb_modal_attn_config* cfg, void* stream, const float* params, int N, const void* feat_bf16,
                            int has_hidden, const void* d_states_bf16, void* workspace, size_t workspace_bytes, void* d_feat_bf16,
                            float* grads) {
  MaLayout L;
  DTB_TRY(make_layout(cfg, L));
  DTB_REQUIRE(params && feat_bf16 && d_states_bf16 && d_feat_bf16 && grads, DTB_ERR_BAD_ARG, "dtb_modal_attn_backward: null argument");
  cudaStream_t st = (cudaStream_t)stream;
  if (N == 0) {
    DTB_CUDA(cudaMemsetAsync(grads, 0, (size_t)L.total * sizeof(float), st));
    return DTB_OK;
  }
  MaWs w;
  DTB_TRY(get_ws(L, N, workspace, workspace_bytes, w));  // same carve as the forward: Wc, hid, z, w are still there
  const bf16* feat = (const bf16*)feat_bf16;
  const float* Wat = params + L.wat_off;
  const int F = L.F, J = L.J;
  if (J == 32)
    DTB_LAUNCH(ma_bwd_kernel<32>, warp_grid(N), 256, 0, st, (const bf16*)d_states_bf16, feat, (const float*)w.z, (const float*)w.w, Wat, N, F,
               L.fusion, w.dpre, w.dl, w.R);
  else
    DTB_LAUNCH(ma_bwd_kernel<128>, warp_grid(N), 256, 0, st, (const bf16*)d_states_bf16, feat, (const float*)w.z, (const float*)w.w, Wat, N, F,
               L.fusion, w.dpre, w.dl, w.R);
  DTB_CHECK_LAUNCH();
  // d_feat = relu'(feat) * (dpre Wc[:, :2F]) + R     (W [J][2F] read in place as the MN-major B operand)
  {
    GemmEpilogue ep = epz();
    ep.dact = 1;
    ep.aux = feat;
    ep.ldaux = 2 * F;
    ep.residual = w.R;
    ep.ldr = 2 * F;
    ep.out_bf16 = (bf16*)d_feat_bf16;
    ep.ldo_b = 2 * F;
    const int rc = gemm_bf16(st, w.dpre, J, 0, w.Wc, L.Kc, 1, N, 2 * F, J, ep, 1, 0);
    if (rc < 0) return rc;
  }
  // attention.{weight, bias}
  {
    const int slices = N < RED_SLICES * 8 ? (N + 7) / 8 : RED_SLICES;
    DTB_LAUNCH(ma_wat_partial_kernel, dim3(J / 32, slices), 256, 0, st, (const float*)w.dl, (const float*)w.z, N, J, w.red);
    DTB_CHECK_LAUNCH();
    DTB_LAUNCH(ma_wat_final_kernel, (2 * (J + 1) + 127) / 128, 128, 0, st, (const float*)w.red, slices, J, grads + L.wat_off,
               grads + L.bat_off);
    DTB_CHECK_LAUNCH();
  }
  // weight gradients dW = dpre^T x (both operands read in place), bias gradient = column sums of dpre
  if (L.type == 0) {
    DTB_TRY(gemm_wgrad(st, w.dpre, J, feat, 2 * F, J, F, N, grads + L.w_off[1], F, w.partial, w.partial_floats));      // video_fc
    DTB_TRY(gemm_wgrad(st, w.dpre, J, feat + F, 2 * F, J, F, N, grads + L.w_off[0], F, w.partial, w.partial_floats));  // audio_fc
    if (has_hidden) DTB_TRY(gemm_wgrad(st, w.dpre, J, w.hid, HID, J, HID, N, grads + L.w_off[2], HID, w.partial, w.partial_floats));
    else {
      DTB_LAUNCH(ma_zero_kernel, (J * HID + 255) / 256, 256, 0, st, grads + L.w_off[2], J * HID);
      DTB_CHECK_LAUNCH();
    }
    DTB_TRY(colsum_bf16(st, w.dpre, N, J, J, grads + L.b_off[0], 0, w.cs_scratch));
    DTB_LAUNCH(ma_copy_kernel, 1, 256, 0, st, (const float*)(grads + L.b_off[0]), J, grads + L.b_off[1], grads + L.b_off[2]);
    DTB_CHECK_LAUNCH();
  } else {
    float* dW = grads + L.w_off[0];
    DTB_TRY(gemm_wgrad(st, w.dpre, J, feat, 2 * F, J, 2 * F, N, dW, L.Kc, w.partial, w.partial_floats));
    if (has_hidden) DTB_TRY(gemm_wgrad(st, w.dpre, J, w.hid, HID, J, HID, N, dW + 2 * F, L.Kc, w.partial, w.partial_floats));
    else {
      // zero the hidden columns [J, 128] of fc.weight's gradient (row stride Kc)
      DTB_CUDA(cudaMemset2DAsync(dW + 2 * F, (size_t)L.Kc * sizeof(float), 0, HID * sizeof(float), J, st));
    }
    DTB_TRY(colsum_bf16(st, w.dpre, N, J, J, grads + L.b_off[0], 0, w.cs_scratch));
  }
  return DTB_OK;
}

}  // extern "C"
