// Fused attention backward (autograd of HF:47-65 under the causal + padding mask rule of attention.cu).
// Two deterministic passes, both recomputing P = exp2(S * scale_log2 - lse) tile by tile (nothing T x T in HBM):
//   pass 1 (per 64-query block): delta = rowsum(dO * O);  dQ = sum_j dS_ij K_j          -> dqkv[:, :, 0:H]
//   pass 2 (per 64-key block)  : dV = P^T dO;  dK = dS^T Q                               -> dqkv[:, :, H:3H]
// with dS = P * (dP - delta) / sqrt(d), dP = dO V^T.  No atomics, fixed summation order.
#include "kernels.h"
#include "attention_common.cuh"

namespace dtb {

template <int D>
__global__ void __launch_bounds__(128) attn_bwd_dq_kernel(const bf16* __restrict__ qkv, const float* __restrict__ mask,
                                                          const bf16* __restrict__ out, const bf16* __restrict__ dout,
                                                          const float* __restrict__ lse, int T, int nh, int Kmask, float scale,
                                                          float scale_log2, bf16* __restrict__ dqkv, float* __restrict__ delta,
                                                          const DropSite drop) {
  pdl_trigger();
  pdl_wait();
  const uint32_t drop_key = drop.state ? drop_group_key(drop_site_key(drop), blockIdx.z * nh + blockIdx.y) : 0u;
  constexpr int P = D + 8;
  extern __shared__ __align__(16) uint8_t smem_att[];
  bf16* sQ = reinterpret_cast<bf16*>(smem_att);
  bf16* sdO = sQ + 64 * P;
  bf16* sK = sdO + 64 * P;
  bf16* sV = sK + 64 * P;
  float* sValid = reinterpret_cast<float*>(sV + 64 * P);
  float* sLse = sValid + 64;
  float* sDelta = sLse + 64;

  const int qb = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
  const int H = nh * D, ld = 3 * H;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t4 = lane & 3;
  const int q0 = qb * 64;
  const bf16* base = qkv + (size_t)b * T * ld + h * D;
  const bf16* obase = out + (size_t)b * T * H + h * D;
  const bf16* dobase = dout + (size_t)b * T * H + h * D;
  const size_t row_off = ((size_t)b * nh + h) * T;

  load_tile<D>(sQ, base, ld, q0, T);
  load_tile<D>(sdO, dobase, H, q0, T);
  load_tile<D>(sK, obase, H, q0, T);  // O staged in the K buffer for the delta pass
  __syncthreads();
  {
    const int r = threadIdx.x >> 1, half = threadIdx.x & 1;
    float acc = 0.f;
#pragma unroll
    for (int c = half * (D / 2); c < (half + 1) * (D / 2); c += 2) {
      const float2 a = unpack_bf16x2(*reinterpret_cast<const uint32_t*>(sdO + r * P + c));
      const float2 o = unpack_bf16x2(*reinterpret_cast<const uint32_t*>(sK + r * P + c));
      acc += a.x * o.x + a.y * o.y;
    }
    acc += __shfl_xor_sync(0xffffffffu, acc, 1);
    if (half == 0) {
      sDelta[r] = acc;
      const bool ok = q0 + r < T;
      if (ok) delta[row_off + q0 + r] = acc;
      sLse[r] = ok ? lse[row_off + q0 + r] : 0.f;
    }
  }

  float dq[D / 8][4];
#pragma unroll
  for (int i = 0; i < D / 8; ++i) dq[i][0] = dq[i][1] = dq[i][2] = dq[i][3] = 0.f;
  const int r0 = warp * 16 + g, r1 = r0 + 8;
  const int qi0 = q0 + r0, qi1 = q0 + r1;

  for (int kb = 0; kb <= qb; ++kb) {
    const int k0 = kb * 64;
    __syncthreads();
    load_tile<D>(sK, base + H, ld, k0, T);
    load_tile<D>(sV, base + 2 * H, ld, k0, T);
    if (threadIdx.x < 64) {
      const int kj = k0 + threadIdx.x;
      sValid[threadIdx.x] = (kj < T && (mask == nullptr || mask[(size_t)b * Kmask + kj / 3] > 0.f)) ? 1.f : 0.f;
    }
    __syncthreads();
    const float lse0 = sLse[r0], lse1 = sLse[r1], dl0 = sDelta[r0], dl1 = sDelta[r1];

    float s[8][4], dp[8][4];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      s[j][0] = s[j][1] = s[j][2] = s[j][3] = 0.f;
      dp[j][0] = dp[j][1] = dp[j][2] = dp[j][3] = 0.f;
    }
#pragma unroll
    for (int kk = 0; kk < D / 16; ++kk) {
      uint32_t qf[4], dof[4];
      ldsm_x4(qf, sQ + (warp * 16 + (lane & 15)) * P + kk * 16 + (lane >> 4) * 8);
      ldsm_x4(dof, sdO + (warp * 16 + (lane & 15)) * P + kk * 16 + (lane >> 4) * 8);
#pragma unroll
      for (int jp = 0; jp < 4; ++jp) {
        uint32_t kf[4], vf[4];
        const int off = (jp * 16 + (lane >> 4) * 8 + (lane & 7)) * P + kk * 16 + ((lane >> 3) & 1) * 8;
        ldsm_x4(kf, sK + off);
        ldsm_x4(vf, sV + off);
        mma_16816(s[2 * jp], qf, kf[0], kf[1]);
        mma_16816(s[2 * jp + 1], qf, kf[2], kf[3]);
        mma_16816(dp[2 * jp], dof, vf[0], vf[1]);
        mma_16816(dp[2 * jp + 1], dof, vf[2], vf[3]);
      }
    }
    uint32_t dsf[4][4];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float ds[4];
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const int kl = j * 8 + 2 * t4 + e;
        const int kj = k0 + kl;
        const bool kv = sValid[kl] > 0.f;
        const float p0 = (kv && kj <= qi0) ? exp2f(s[j][e] * scale_log2 - lse0) : 0.f;
        const float p1 = (kv && kj <= qi1) ? exp2f(s[j][2 + e] * scale_log2 - lse1) : 0.f;
        float g0 = dp[j][e], g1 = dp[j][2 + e];
        if (drop.state) {  // dP = D o (dO V^T), D = keep / (1 - p): the bits of the forward pass
          g0 = drop_apply(g0, drop_key, (uint32_t)qi0 * (uint32_t)T + (uint32_t)kj, drop.thr, drop.scale);
          g1 = drop_apply(g1, drop_key, (uint32_t)qi1 * (uint32_t)T + (uint32_t)kj, drop.thr, drop.scale);
        }
        ds[e] = p0 * (g0 - dl0) * scale;
        ds[2 + e] = p1 * (g1 - dl1) * scale;
      }
      dsf[j >> 1][(j & 1) * 2 + 0] = pack_bf16x2(ds[0], ds[1]);
      dsf[j >> 1][(j & 1) * 2 + 1] = pack_bf16x2(ds[2], ds[3]);
    }
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
#pragma unroll
      for (int ip = 0; ip < D / 16; ++ip) {
        uint32_t bfr[4];
        ldsm_x4_t(bfr, sK + (kk * 16 + ((lane >> 3) & 1) * 8 + (lane & 7)) * P + ip * 16 + (lane >> 4) * 8);
        mma_16816(dq[2 * ip], dsf[kk], bfr[0], bfr[1]);
        mma_16816(dq[2 * ip + 1], dsf[kk], bfr[2], bfr[3]);
      }
    }
  }
  bf16* dqb = dqkv + (size_t)b * T * ld + h * D;
#pragma unroll
  for (int i = 0; i < D / 8; ++i) {
    const int c = i * 8 + 2 * t4;
    if (qi0 < T) *reinterpret_cast<uint32_t*>(dqb + (size_t)qi0 * ld + c) = pack_bf16x2(dq[i][0], dq[i][1]);
    if (qi1 < T) *reinterpret_cast<uint32_t*>(dqb + (size_t)qi1 * ld + c) = pack_bf16x2(dq[i][2], dq[i][3]);
  }
}

template <int D>
__global__ void __launch_bounds__(128) attn_bwd_dkv_kernel(const bf16* __restrict__ qkv, const float* __restrict__ mask,
                                                           const bf16* __restrict__ dout, const float* __restrict__ lse,
                                                           const float* __restrict__ delta, int T, int nh, int Kmask, float scale,
                                                           float scale_log2, bf16* __restrict__ dqkv, const DropSite drop) {
  pdl_trigger();
  pdl_wait();
  const uint32_t drop_key = drop.state ? drop_group_key(drop_site_key(drop), blockIdx.z * nh + blockIdx.y) : 0u;
  constexpr int P = D + 8;
  extern __shared__ __align__(16) uint8_t smem_att[];
  bf16* sK = reinterpret_cast<bf16*>(smem_att);
  bf16* sV = sK + 64 * P;
  bf16* sQ = sV + 64 * P;
  bf16* sdO = sQ + 64 * P;
  float* sLse = reinterpret_cast<float*>(sdO + 64 * P);
  float* sDelta = sLse + 64;

  const int kb = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
  const int H = nh * D, ld = 3 * H;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t4 = lane & 3;
  const int k0 = kb * 64;
  const bf16* base = qkv + (size_t)b * T * ld + h * D;
  const bf16* dobase = dout + (size_t)b * T * H + h * D;
  const size_t row_off = ((size_t)b * nh + h) * T;
  const int nqb = (T + 63) / 64;

  load_tile<D>(sK, base + H, ld, k0, T);
  load_tile<D>(sV, base + 2 * H, ld, k0, T);

  const int kr0 = warp * 16 + g, kr1 = kr0 + 8;
  const int kj0 = k0 + kr0, kj1 = k0 + kr1;
  const bool kv0 = kj0 < T && (mask == nullptr || mask[(size_t)b * Kmask + kj0 / 3] > 0.f);
  const bool kv1 = kj1 < T && (mask == nullptr || mask[(size_t)b * Kmask + kj1 / 3] > 0.f);

  float dk[D / 8][4], dv[D / 8][4];
#pragma unroll
  for (int i = 0; i < D / 8; ++i) {
    dk[i][0] = dk[i][1] = dk[i][2] = dk[i][3] = 0.f;
    dv[i][0] = dv[i][1] = dv[i][2] = dv[i][3] = 0.f;
  }

  for (int qb = kb; qb < nqb; ++qb) {
    const int q0 = qb * 64;
    __syncthreads();
    load_tile<D>(sQ, base, ld, q0, T);
    load_tile<D>(sdO, dobase, H, q0, T);
    if (threadIdx.x < 64) {
      const bool ok = q0 + threadIdx.x < T;
      sLse[threadIdx.x] = ok ? lse[row_off + q0 + threadIdx.x] : 0.f;
      sDelta[threadIdx.x] = ok ? delta[row_off + q0 + threadIdx.x] : 0.f;
    }
    __syncthreads();

    float st[8][4], dpt[8][4];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      st[j][0] = st[j][1] = st[j][2] = st[j][3] = 0.f;
      dpt[j][0] = dpt[j][1] = dpt[j][2] = dpt[j][3] = 0.f;
    }
#pragma unroll
    for (int kk = 0; kk < D / 16; ++kk) {
      uint32_t kf[4], vf[4];
      ldsm_x4(kf, sK + (warp * 16 + (lane & 15)) * P + kk * 16 + (lane >> 4) * 8);
      ldsm_x4(vf, sV + (warp * 16 + (lane & 15)) * P + kk * 16 + (lane >> 4) * 8);
#pragma unroll
      for (int jp = 0; jp < 4; ++jp) {
        uint32_t qf[4], dof[4];
        const int off = (jp * 16 + (lane >> 4) * 8 + (lane & 7)) * P + kk * 16 + ((lane >> 3) & 1) * 8;
        ldsm_x4(qf, sQ + off);
        ldsm_x4(dof, sdO + off);
        mma_16816(st[2 * jp], kf, qf[0], qf[1]);
        mma_16816(st[2 * jp + 1], kf, qf[2], qf[3]);
        mma_16816(dpt[2 * jp], vf, dof[0], dof[1]);
        mma_16816(dpt[2 * jp + 1], vf, dof[2], dof[3]);
      }
    }
    uint32_t pf[4][4], dsf[4][4];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float p[4], ds[4];
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const int ql = j * 8 + 2 * t4 + e;
        const int qi = q0 + ql;
        const float l = sLse[ql], dl = sDelta[ql];
        const bool qok = qi < T;
        p[e] = (qok && kv0 && kj0 <= qi) ? exp2f(st[j][e] * scale_log2 - l) : 0.f;
        p[2 + e] = (qok && kv1 && kj1 <= qi) ? exp2f(st[j][2 + e] * scale_log2 - l) : 0.f;
        float g0 = dpt[j][e], g1 = dpt[j][2 + e];
        if (drop.state) {
          const uint32_t base_i = drop_key + (uint32_t)qi * (uint32_t)T;
          const bool k0_keep = drop_keep(base_i, (uint32_t)kj0, drop.thr), k1_keep = drop_keep(base_i, (uint32_t)kj1, drop.thr);
          g0 = k0_keep ? g0 * drop.scale : 0.f;
          g1 = k1_keep ? g1 * drop.scale : 0.f;
          ds[e] = p[e] * (g0 - dl) * scale;
          ds[2 + e] = p[2 + e] * (g1 - dl) * scale;
          p[e] = k0_keep ? p[e] * drop.scale : 0.f;          // dV = (P o D)^T dO
          p[2 + e] = k1_keep ? p[2 + e] * drop.scale : 0.f;
        } else {
          ds[e] = p[e] * (g0 - dl) * scale;
          ds[2 + e] = p[2 + e] * (g1 - dl) * scale;
        }
      }
      pf[j >> 1][(j & 1) * 2 + 0] = pack_bf16x2(p[0], p[1]);
      pf[j >> 1][(j & 1) * 2 + 1] = pack_bf16x2(p[2], p[3]);
      dsf[j >> 1][(j & 1) * 2 + 0] = pack_bf16x2(ds[0], ds[1]);
      dsf[j >> 1][(j & 1) * 2 + 1] = pack_bf16x2(ds[2], ds[3]);
    }
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
#pragma unroll
      for (int ip = 0; ip < D / 16; ++ip) {
        uint32_t b1[4], b2[4];
        const int off = (kk * 16 + ((lane >> 3) & 1) * 8 + (lane & 7)) * P + ip * 16 + (lane >> 4) * 8;
        ldsm_x4_t(b1, sdO + off);
        ldsm_x4_t(b2, sQ + off);
        mma_16816(dv[2 * ip], pf[kk], b1[0], b1[1]);
        mma_16816(dv[2 * ip + 1], pf[kk], b1[2], b1[3]);
        mma_16816(dk[2 * ip], dsf[kk], b2[0], b2[1]);
        mma_16816(dk[2 * ip + 1], dsf[kk], b2[2], b2[3]);
      }
    }
  }
  bf16* dkb = dqkv + (size_t)b * T * ld + H + h * D;
  bf16* dvb = dkb + H;
#pragma unroll
  for (int i = 0; i < D / 8; ++i) {
    const int c = i * 8 + 2 * t4;
    if (kj0 < T) {
      *reinterpret_cast<uint32_t*>(dkb + (size_t)kj0 * ld + c) = pack_bf16x2(dk[i][0], dk[i][1]);
      *reinterpret_cast<uint32_t*>(dvb + (size_t)kj0 * ld + c) = pack_bf16x2(dv[i][0], dv[i][1]);
    }
    if (kj1 < T) {
      *reinterpret_cast<uint32_t*>(dkb + (size_t)kj1 * ld + c) = pack_bf16x2(dk[i][2], dk[i][3]);
      *reinterpret_cast<uint32_t*>(dvb + (size_t)kj1 * ld + c) = pack_bf16x2(dv[i][2], dv[i][3]);
    }
  }
}

template <int D>
static int launch_bwd(cudaStream_t stream, const bf16* qkv, const float* mask, const bf16* out, const bf16* dout, const float* lse, int B,
                      int T, int nh, bf16* dqkv, float* delta, const DropSite& drop) {
  const int smem = 4 * 64 * (D + 8) * 2 + 3 * 64 * 4;
  static bool attr = false;
  if (!attr) {
    DTB_CUDA(cudaFuncSetAttribute(attn_bwd_dq_kernel<D>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    DTB_CUDA(cudaFuncSetAttribute(attn_bwd_dkv_kernel<D>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    attr = true;
  }
  const float scale = 1.f / sqrtf((float)D);
  const float scale_log2 = 1.4426950408889634f * scale;
  dim3 grid((T + 63) / 64, nh, B);
  ProfScope prof(stream, PF_ATTN_BWD, 5.0 * B * nh * (double)T * T * D);
  DTB_LAUNCH(attn_bwd_dq_kernel<D>, grid, 128, smem, stream, qkv, mask, out, dout, lse, T, nh, T / 3, scale, scale_log2, dqkv, delta, drop);
  DTB_CHECK_LAUNCH();
  DTB_LAUNCH(attn_bwd_dkv_kernel<D>, grid, 128, smem, stream, qkv, mask, dout, lse, delta, T, nh, T / 3, scale, scale_log2, dqkv, drop);
  DTB_CHECK_LAUNCH();
  return DTB_OK;
}

int attn_bwd(cudaStream_t stream, const bf16* qkv, const float* mask, const bf16* out, const bf16* dout, const float* lse, int B, int T,
             int nh, int d, bf16* dqkv, float* delta, const DropSite& drop) {
  DTB_REQUIRE(T % 3 == 0 || mask == nullptr, DTB_ERR_BAD_ARG, "attention bwd: T (%d) must be 3*K when a [B,K] mask is given", T);
  if (B == 0 || T == 0) return DTB_OK;
  if (attn_tc_supported(T, d)) return attn_bwd_tc(stream, qkv, mask, out, dout, lse, B, T, nh, d, dqkv, delta, drop);
  switch (d) {
    case 32: return launch_bwd<32>(stream, qkv, mask, out, dout, lse, B, T, nh, dqkv, delta, drop);
    case 64: return launch_bwd<64>(stream, qkv, mask, out, dout, lse, B, T, nh, dqkv, delta, drop);
    case 128: return launch_bwd<128>(stream, qkv, mask, out, dout, lse, B, T, nh, dqkv, delta, drop);
    default: set_error("attention bwd: head_dim %d unsupported (32, 64, 128)", d); return DTB_ERR_UNSUPPORTED;
  }
}

}  // namespace dtb
