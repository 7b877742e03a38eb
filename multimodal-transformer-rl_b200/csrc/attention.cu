// Fused causal self-attention with padding mask for the interleaved (R,s,a) token stream — forward.
// Replaces eager_attention_forward (HF:47-65) + the materialised [B,1,T,T] mask (HF:408-414, 631-635):
//   allowed(i, j) = (j <= i) and attention_mask[b, j / 3] > 0        (integer-exact index rule, A3)
// Flash-style: online softmax in registers with warp-shuffle row reductions, S/P never touch HBM,
// only the log-sum-exp (base 2) per row is saved for backward.  mma.sync m16n8k16 bf16 tiles.
// Layout: qkv bf16 [B, T, 3H] exactly as the c_attn GEMM writes it (q | k | v, head h at h*d); out bf16 [B, T, H].
#include "kernels.h"
#include "attention_common.cuh"

namespace dtb {

template <int D>
__global__ void __launch_bounds__(128) attn_fwd_kernel(const bf16* __restrict__ qkv, const float* __restrict__ mask, int T, int nh,
                                                       int Kmask, float scale_log2, bf16* __restrict__ out, float* __restrict__ lse,
                                                       const DropSite drop) {
  pdl_trigger();
  pdl_wait();
  const uint32_t drop_key = drop.state ? drop_group_key(drop_site_key(drop), blockIdx.z * nh + blockIdx.y) : 0u;
  constexpr int P = D + 8;  // smem row pitch (elements): 16-byte pad keeps ldmatrix conflict-free
  extern __shared__ __align__(16) uint8_t smem_att[];
  bf16* sQ = reinterpret_cast<bf16*>(smem_att);
  bf16* sK = sQ + 64 * P;
  bf16* sV = sK + 64 * P;
  float* sValid = reinterpret_cast<float*>(sV + 64 * P);  // [64] key validity of the current tile

  const int qb = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
  const int H = nh * D, ld = 3 * H;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t4 = lane & 3;
  const int q0 = qb * 64;
  const bf16* base = qkv + (size_t)b * T * ld + h * D;

  // Q and the first key / value tile are fetched in one sweep (one L2 round trip instead of two before the first MMA)
  auto load_kv = [&](int k0) {
    load_tile<D>(sK, base + H, ld, k0, T);
    load_tile<D>(sV, base + 2 * H, ld, k0, T);
    if (threadIdx.x < 64) {
      const int kj = k0 + threadIdx.x;
      sValid[threadIdx.x] = (kj < T && (mask == nullptr || mask[(size_t)b * Kmask + kj / 3] > 0.f)) ? 1.f : 0.f;
    }
  };
  load_tile<D>(sQ, base, ld, q0, T);
  load_kv(0);
  __syncthreads();
  uint32_t qf[D / 16][4];
#pragma unroll
  for (int kk = 0; kk < D / 16; ++kk) ldsm_x4(qf[kk], sQ + (warp * 16 + (lane & 15)) * P + kk * 16 + (lane >> 4) * 8);

  float o[D / 8][4];
#pragma unroll
  for (int i = 0; i < D / 8; ++i) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;
  float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;
  const int qi0 = q0 + warp * 16 + g, qi1 = qi0 + 8;

  for (int kb = 0; kb <= qb; ++kb) {
    const int k0 = kb * 64;
    if (kb > 0) {
      __syncthreads();  // previous tile fully consumed
      load_kv(k0);
      __syncthreads();
    }

    float s[8][4];
#pragma unroll
    for (int j = 0; j < 8; ++j) s[j][0] = s[j][1] = s[j][2] = s[j][3] = 0.f;
#pragma unroll
    for (int kk = 0; kk < D / 16; ++kk) {
#pragma unroll
      for (int jp = 0; jp < 4; ++jp) {  // two 8-key n-tiles per ldmatrix.x4
        uint32_t bfr[4];
        ldsm_x4(bfr, sK + (jp * 16 + (lane >> 4) * 8 + (lane & 7)) * P + kk * 16 + ((lane >> 3) & 1) * 8);
        mma_16816(s[2 * jp], qf[kk], bfr[0], bfr[1]);
        mma_16816(s[2 * jp + 1], qf[kk], bfr[2], bfr[3]);
      }
    }
    // mask + online softmax (base-2 domain)
    float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const int kl = j * 8 + 2 * t4 + e;
        const int kj = k0 + kl;
        const bool kv = sValid[kl] > 0.f;
        s[j][e] = (kv && kj <= qi0) ? s[j][e] * scale_log2 : -INFINITY;
        s[j][2 + e] = (kv && kj <= qi1) ? s[j][2 + e] * scale_log2 : -INFINITY;
        mx0 = fmaxf(mx0, s[j][e]);
        mx1 = fmaxf(mx1, s[j][2 + e]);
      }
    }
    mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
    mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
    mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
    mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
    const float mn0 = fmaxf(m0, mx0), mn1 = fmaxf(m1, mx1);
    const float ms0 = mn0 == -INFINITY ? 0.f : mn0, ms1 = mn1 == -INFINITY ? 0.f : mn1;  // fully-masked rows stay finite
    const float c0 = exp2f(m0 - ms0), c1 = exp2f(m1 - ms1);
    m0 = mn0;
    m1 = mn1;
    float rs0 = 0.f, rs1 = 0.f;
    uint32_t pf[4][4];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float p00 = exp2f(s[j][0] - ms0), p01 = exp2f(s[j][1] - ms0);
      float p10 = exp2f(s[j][2] - ms1), p11 = exp2f(s[j][3] - ms1);
      rs0 += p00 + p01;
      rs1 += p10 + p11;
      if (drop.state) {  // attn_dropout on the softmax weights: the normaliser keeps the undropped sum (HF:60-62)
        const uint32_t kj = (uint32_t)(k0 + j * 8 + 2 * t4);
        const uint32_t i0 = drop_key + (uint32_t)qi0 * (uint32_t)T + kj, i1 = drop_key + (uint32_t)qi1 * (uint32_t)T + kj;
        p00 = drop_apply(p00, i0, 0u, drop.thr, drop.scale); p01 = drop_apply(p01, i0, 1u, drop.thr, drop.scale);
        p10 = drop_apply(p10, i1, 0u, drop.thr, drop.scale); p11 = drop_apply(p11, i1, 1u, drop.thr, drop.scale);
      }
      pf[j >> 1][(j & 1) * 2 + 0] = pack_bf16x2(p00, p01);
      pf[j >> 1][(j & 1) * 2 + 1] = pack_bf16x2(p10, p11);
    }
    l0 = l0 * c0 + rs0;
    l1 = l1 * c1 + rs1;
#pragma unroll
    for (int i = 0; i < D / 8; ++i) {
      o[i][0] *= c0; o[i][1] *= c0; o[i][2] *= c1; o[i][3] *= c1;
    }
    // O += P V
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
#pragma unroll
      for (int ip = 0; ip < D / 16; ++ip) {
        uint32_t bfr[4];
        ldsm_x4_t(bfr, sV + (kk * 16 + ((lane >> 3) & 1) * 8 + (lane & 7)) * P + ip * 16 + (lane >> 4) * 8);
        mma_16816(o[2 * ip], pf[kk], bfr[0], bfr[1]);
        mma_16816(o[2 * ip + 1], pf[kk], bfr[2], bfr[3]);
      }
    }
  }
  // finalize: each quad holds partial row sums
  l0 += __shfl_xor_sync(0xffffffffu, l0, 1);
  l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
  l1 += __shfl_xor_sync(0xffffffffu, l1, 1);
  l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
  const float inv0 = l0 > 0.f ? 1.f / l0 : 0.f, inv1 = l1 > 0.f ? 1.f / l1 : 0.f;
  bf16* ob = out + (size_t)b * T * H + h * D;
#pragma unroll
  for (int i = 0; i < D / 8; ++i) {
    const int c = i * 8 + 2 * t4;
    if (qi0 < T) *reinterpret_cast<uint32_t*>(ob + (size_t)qi0 * H + c) = pack_bf16x2(o[i][0] * inv0, o[i][1] * inv0);
    if (qi1 < T) *reinterpret_cast<uint32_t*>(ob + (size_t)qi1 * H + c) = pack_bf16x2(o[i][2] * inv1, o[i][3] * inv1);
  }
  if (t4 == 0) {
    float* lp = lse + ((size_t)b * nh + h) * T;
    if (qi0 < T) lp[qi0] = l0 > 0.f ? m0 + log2f(l0) : 0.f;
    if (qi1 < T) lp[qi1] = l1 > 0.f ? m1 + log2f(l1) : 0.f;
  }
}

template <int D>
static int launch_fwd(cudaStream_t stream, const bf16* qkv, const float* mask, int B, int T, int nh, bf16* out, float* lse,
                      const DropSite& drop) {
  const int smem = 3 * 64 * (D + 8) * 2 + 64 * 4;
  static bool attr = false;
  if (!attr) {
    DTB_CUDA(cudaFuncSetAttribute(attn_fwd_kernel<D>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    attr = true;
  }
  const float scale_log2 = 1.4426950408889634f / sqrtf((float)D);
  dim3 grid((T + 63) / 64, nh, B);
  ProfScope prof(stream, PF_ATTN_FWD, 2.0 * B * nh * (double)T * T * D);
  DTB_LAUNCH(attn_fwd_kernel<D>, grid, 128, smem, stream, qkv, mask, T, nh, T / 3, scale_log2, out, lse, drop);
  DTB_CHECK_LAUNCH();
  return DTB_OK;
}

int attn_fwd(cudaStream_t stream, const bf16* qkv, const float* mask, int B, int T, int nh, int d, bf16* out, float* lse,
             const DropSite& drop) {
  DTB_REQUIRE(T % 3 == 0 || mask == nullptr, DTB_ERR_BAD_ARG, "attention: T (%d) must be 3*K when a [B,K] mask is given", T);
  DTB_REQUIRE(drop.state == nullptr || (double)T * T < 4294967296.0, DTB_ERR_UNSUPPORTED, "attention dropout: T too large");
  if (B == 0 || T == 0) return DTB_OK;
  if (attn_tc_supported(T, d)) return attn_fwd_tc(stream, qkv, mask, B, T, nh, d, out, lse, drop);
  switch (d) {
    case 32: return launch_fwd<32>(stream, qkv, mask, B, T, nh, out, lse, drop);
    case 64: return launch_fwd<64>(stream, qkv, mask, B, T, nh, out, lse, drop);
    case 128: return launch_fwd<128>(stream, qkv, mask, B, T, nh, out, lse, drop);
    default: set_error("attention: head_dim %d unsupported (32, 64, 128)", d); return DTB_ERR_UNSUPPORTED;
  }
}

}  // namespace dtb
