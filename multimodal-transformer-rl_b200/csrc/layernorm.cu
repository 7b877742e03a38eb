// LayerNorm forward / backward for the fp32 residual stream (HBM-bound; one warp per token row).
// Replaces nn.LayerNorm at HF:272 (ln_1), HF:303 (ln_2) and its autograd backward.
//   fwd : x fp32 [M,H] -> y bf16 [M,H] (GEMM operand), mean/rstd fp32 [M] saved for backward
//   bwd : dx = dres + LN'(dy) written as fp32 (gradient residual stream) and bf16 (next GEMM operand);
//         dgamma/dbeta via per-block partials reduced in index order (deterministic).
#include "kernels.h"

namespace dtb {

constexpr int LN_MAX_VEC = 8;  // H <= 8 * 128 = 1024

__global__ void __launch_bounds__(256) ln_fwd_kernel(const float* __restrict__ x, const float* __restrict__ gamma,
                                                     const float* __restrict__ beta, int M, int H, float eps, bf16* __restrict__ y,
                                                     float* __restrict__ mean_out, float* __restrict__ rstd_out) {
  pdl_trigger();
  pdl_wait();
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= M) return;
  const float* xr = x + (size_t)row * H;
  float4 v[LN_MAX_VEC];
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < LN_MAX_VEC; ++i) {
    const int c = i * 128 + lane * 4;
    if (c < H) {
      v[i] = ldg_f4(xr + c);
      s += v[i].x + v[i].y + v[i].z + v[i].w;
    }
  }
  const float mean = warp_sum(s) / H;
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < LN_MAX_VEC; ++i) {
    const int c = i * 128 + lane * 4;
    if (c < H) {
      const float a = v[i].x - mean, b = v[i].y - mean, cc = v[i].z - mean, d = v[i].w - mean;
      q += a * a + b * b + cc * cc + d * d;
    }
  }
  const float rstd = rsqrtf(warp_sum(q) / H + eps);
  if (lane == 0) {
    if (mean_out) mean_out[row] = mean;
    if (rstd_out) rstd_out[row] = rstd;
  }
  bf16* yr = y + (size_t)row * H;
#pragma unroll
  for (int i = 0; i < LN_MAX_VEC; ++i) {
    const int c = i * 128 + lane * 4;
    if (c < H) {
      const float4 g = ldg_f4(gamma + c), b = ldg_f4(beta + c);
      const float o0 = (v[i].x - mean) * rstd * g.x + b.x, o1 = (v[i].y - mean) * rstd * g.y + b.y;
      const float o2 = (v[i].z - mean) * rstd * g.z + b.z, o3 = (v[i].w - mean) * rstd * g.w + b.w;
      *reinterpret_cast<uint2*>(yr + c) = make_uint2(pack_bf16x2(o0, o1), pack_bf16x2(o2, o3));
    }
  }
}

int ln_fwd(cudaStream_t stream, const float* x, const float* gamma, const float* beta, int M, int H, float eps, bf16* y, float* mean,
           float* rstd) {
  DTB_REQUIRE(H % 4 == 0 && H <= LN_MAX_VEC * 128, DTB_ERR_UNSUPPORTED, "layernorm: hidden size %d unsupported (multiple of 4, <= 1024)", H);
  if (M == 0) return DTB_OK;
  ProfScope prof(stream, PF_LN_FWD, 6.0 * M * H);
  DTB_LAUNCH(ln_fwd_kernel, (M + 7) / 8, 256, 0, stream, x, gamma, beta, M, H, eps, y, mean, rstd);
  DTB_CHECK_LAUNCH();
  return DTB_OK;
}

// ---------------------------------------------------------------------------------------------------
// One warp per token row, NV = ceil(H / 128) float4 per lane (compile-time so the per-lane arrays are exact).  All of a row's loads
// (x, dy, dres) are issued before the first reduction so ~10 KB per warp is in flight; two 256-thread CTAs per SM.
// Column reductions (dgamma, dbeta and — CS — the column sums of the bf16 output, i.e. the bias gradient of the c_proj whose output
// gradient this is) leave the kernel as per-block partials, summed in block order by ln_bwd_finalize.
template <int NV, bool CS>
__global__ void __launch_bounds__(256, (NV > 4 && CS) ? 1 : 2) ln_bwd_kernel(const bf16* __restrict__ dy, const float* __restrict__ x,
                                                        const float* __restrict__ mean, const float* __restrict__ rstd,
                                                        const float* __restrict__ gamma, const float* __restrict__ dres, int M, int H,
                                                        float* __restrict__ dx_out, bf16* __restrict__ dx_bf16,
                                                        float* __restrict__ scratch /* [grid][NF][H] */, const DropSite drop) {
  pdl_trigger();
  pdl_wait();
  constexpr int NF = CS ? 3 : 2;
  const uint32_t drop_key = drop.state ? drop_site_key(drop) : 0u;
  extern __shared__ float sred[];  // [NF][H]
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  float4 dg[NV], db[NV], dc[CS ? NV : 1];
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    dg[i] = make_float4(0, 0, 0, 0);
    db[i] = make_float4(0, 0, 0, 0);
    if (CS) dc[i] = make_float4(0, 0, 0, 0);
  }
  const int row_step = gridDim.x * nwarp;
  for (int row = blockIdx.x * nwarp + warp; row < M; row += row_step) {
    const float* xr = x + (size_t)row * H;
    const bf16* dyr = dy + (size_t)row * H;
    float4 xv[NV];
    uint2 dyp[NV];
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = i * 128 + lane * 4;
      if (c < H) {
        xv[i] = ldg_f4(xr + c);
        dyp[i] = __ldg(reinterpret_cast<const uint2*>(dyr + c));
      }
    }
    // register-free look-ahead: this row's dres and the warp's next row are pulled into L2 while the reductions run
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = i * 128 + lane * 4;
      if (c < H) {
        if (dres) prefetch_l2(dres + (size_t)row * H + c);
        if (row + row_step < M) {
          prefetch_l2(xr + (size_t)row_step * H + c);
          if ((lane & 1) == 0) prefetch_l2(dyr + (size_t)row_step * H + c);
        }
      }
    }
    const float mu = mean[row], rs = rstd[row];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = i * 128 + lane * 4;
      if (c < H) {
        const float2 d01 = unpack_bf16x2(dyp[i].x), d23 = unpack_bf16x2(dyp[i].y);
        xv[i] = make_float4((xv[i].x - mu) * rs, (xv[i].y - mu) * rs, (xv[i].z - mu) * rs, (xv[i].w - mu) * rs);  // xhat
        const float4 gm = ldg_f4(gamma + c);
        const float g0 = d01.x * gm.x, g1 = d01.y * gm.y, g2 = d23.x * gm.z, g3 = d23.y * gm.w;
        s1 += g0 + g1 + g2 + g3;
        s2 += g0 * xv[i].x + g1 * xv[i].y + g2 * xv[i].z + g3 * xv[i].w;
        dg[i].x += d01.x * xv[i].x; dg[i].y += d01.y * xv[i].y; dg[i].z += d23.x * xv[i].z; dg[i].w += d23.y * xv[i].w;
        db[i].x += d01.x; db[i].y += d01.y; db[i].z += d23.x; db[i].w += d23.y;
      }
    }
    const float c1 = warp_sum(s1) / H, c2 = warp_sum(s2) / H;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = i * 128 + lane * 4;
      if (c < H) {
        const float2 d01 = unpack_bf16x2(dyp[i].x), d23 = unpack_bf16x2(dyp[i].y);
        const float4 gm = ldg_f4(gamma + c);
        float4 o = make_float4(rs * (d01.x * gm.x - c1 - xv[i].x * c2), rs * (d01.y * gm.y - c1 - xv[i].y * c2),
                               rs * (d23.x * gm.z - c1 - xv[i].z * c2), rs * (d23.y * gm.w - c1 - xv[i].w * c2));
        if (dres) {
          const float4 r = ldg_f4(dres + (size_t)row * H + c);
          o.x += r.x; o.y += r.y; o.z += r.z; o.w += r.w;
        }
        if (dx_out) *reinterpret_cast<float4*>(dx_out + (size_t)row * H + c) = o;
        if (dx_bf16 || CS) {
          if (drop.state) {  // the bf16 copy feeds the c_proj backward of a dropout site: d(drop(y)) = mask * scale * d
            const uint32_t k = drop_key + (uint32_t)row * (uint32_t)H + (uint32_t)c;
            o.x = drop_apply(o.x, k, 0u, drop.thr, drop.scale); o.y = drop_apply(o.y, k, 1u, drop.thr, drop.scale);
            o.z = drop_apply(o.z, k, 2u, drop.thr, drop.scale); o.w = drop_apply(o.w, k, 3u, drop.thr, drop.scale);
          }
          if (dx_bf16) *reinterpret_cast<uint2*>(dx_bf16 + (size_t)row * H + c) = make_uint2(pack_bf16x2(o.x, o.y), pack_bf16x2(o.z, o.w));
          if (CS) { dc[i].x += o.x; dc[i].y += o.y; dc[i].z += o.z; dc[i].w += o.w; }
        }
      }
    }
  }
  // deterministic cross-warp reduction of the column partials
  for (int i = threadIdx.x; i < NF * H; i += blockDim.x) sred[i] = 0.f;
  __syncthreads();
  for (int w = 0; w < nwarp; ++w) {
    if (warp == w) {
#pragma unroll
      for (int i = 0; i < NV; ++i) {
        const int c = i * 128 + lane * 4;
        if (c < H) {
          sred[c] += dg[i].x; sred[c + 1] += dg[i].y; sred[c + 2] += dg[i].z; sred[c + 3] += dg[i].w;
          sred[H + c] += db[i].x; sred[H + c + 1] += db[i].y; sred[H + c + 2] += db[i].z; sred[H + c + 3] += db[i].w;
          if (CS) { sred[2 * H + c] += dc[i].x; sred[2 * H + c + 1] += dc[i].y; sred[2 * H + c + 2] += dc[i].z; sred[2 * H + c + 3] += dc[i].w; }
        }
      }
    }
    __syncthreads();
  }
  for (int i = threadIdx.x; i < NF * H; i += blockDim.x) scratch[(size_t)blockIdx.x * NF * H + i] = sred[i];
}

// Column partials -> dgamma | dbeta | (dcolsum): 32 columns x 32 row-lanes per block (short dependent-load chains: nblk / 32 deep), lanes
// combined in index order (deterministic).  A separate small launch on purpose: one CTA summing all partials at the end of ln_bwd_kernel
// ("last block finishes") was measured 5x slower than this parallel pass (75 vs 13 us per LayerNorm backward at c2).
__global__ void __launch_bounds__(1024) ln_bwd_finalize(const float* __restrict__ scratch, int nblk, int H, int NF, float* __restrict__ dgamma,
                                                        float* __restrict__ dbeta, float* __restrict__ dcolsum) {
  pdl_trigger();
  pdl_wait();
  __shared__ float red[32][33];
  const int c = blockIdx.x * 32 + threadIdx.x;
  float s = 0.f;
  if (c < NF * H)
    for (int b = threadIdx.y; b < nblk; b += 32) s += scratch[(size_t)b * NF * H + c];
  red[threadIdx.y][threadIdx.x] = s;
  __syncthreads();
  if (threadIdx.y == 0 && c < NF * H) {
    for (int i = 1; i < 32; ++i) s += red[i][threadIdx.x];
    if (c < H) dgamma[c] = s;
    else if (c < 2 * H) dbeta[c - H] = s;
    else dcolsum[c - 2 * H] = s;
  }
}

static int ln_bwd_blocks(int M, int occ) {
  int b = (M + 7) / 8;
  const int cap = num_sms() * occ;  // resident CTAs per SM; scratch is sized for 8 x SMs partial rows
  return b > cap ? cap : (b < 1 ? 1 : b);
}
size_t ln_bwd_scratch_floats(int H) { return (size_t)DTB_NUM_SMS_FALLBACK * 8 * 3 * H; }

int ln_bwd(cudaStream_t stream, const bf16* dy, const float* x, const float* mean, const float* rstd, const float* gamma,
           const float* dres, int M, int H, float* dx_out, bf16* dx_out_bf16, float* dgamma, float* dbeta, float* scratch,
           const DropSite& drop, float* dcolsum) {
  DTB_REQUIRE(H % 4 == 0 && H <= LN_MAX_VEC * 128, DTB_ERR_UNSUPPORTED, "layernorm bwd: hidden size %d unsupported", H);
  const int nv = (H + 127) / 128;
  const int nf = dcolsum ? 3 : 2;
  const int blocks = ln_bwd_blocks(M, (nv > 4 && dcolsum) ? 1 : 2);
  const size_t sm = (size_t)nf * H * sizeof(float);
  ProfScope prof(stream, PF_LN_BWD, 16.0 * M * H);
#define DTB_LN_BWD(NV)                                                                                                                  \
  do {                                                                                                                                  \
    if (dcolsum) DTB_LAUNCH((ln_bwd_kernel<NV, true>), blocks, 256, sm, stream, dy, x, mean, rstd, gamma, dres, M, H, dx_out, dx_out_bf16, \
                            scratch, drop);                                                                                             \
    else DTB_LAUNCH((ln_bwd_kernel<NV, false>), blocks, 256, sm, stream, dy, x, mean, rstd, gamma, dres, M, H, dx_out, dx_out_bf16,       \
                    scratch, drop);                                                                                                     \
  } while (0)
  if (nv <= 1) DTB_LN_BWD(1);
  else if (nv <= 2) DTB_LN_BWD(2);
  else if (nv <= 4) DTB_LN_BWD(4);
  else if (nv <= 6) DTB_LN_BWD(6);
  else DTB_LN_BWD(8);
#undef DTB_LN_BWD
  DTB_CHECK_LAUNCH();
  DTB_LAUNCH(ln_bwd_finalize, (nf * H + 31) / 32, dim3(32, 32), 0, stream, scratch, blocks, H, nf, dgamma, dbeta, dcolsum);
  DTB_CHECK_LAUNCH();
  return DTB_OK;
}

}  // namespace dtb
