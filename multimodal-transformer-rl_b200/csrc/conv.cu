// Convolution data movement for the multimodal state towers (REF src/agents.py:30-55, A11): the convolutions
// themselves run on the tcgen05 GEMM (rows = output pixels, K = kh*kw*Cin, N = Cout, bias + ReLU in the epilogue,
// which makes the GEMM output the NHWC activation of the next layer).  This file holds the patch gather / scatter
// and the weight re-ordering between PyTorch's layouts (NCHW activations, [Cout,Cin,kh,kw] weights, (c,h,w) flatten)
// and the channels-last layouts the tensor-core path uses.
//   conv1 : fp32 NCHW input -> bf16 patches [N*ho*wo, Cin*k*k], K order (c,kh,kw) = PyTorch weight order; `scale`
//           folds the reference's `/ 255.0` input scaling (agents.py:133-134).
//   conv2+: bf16 NHWC input -> bf16 patches [N*ho*wo, k*k*C], K order (kh,kw,c) (weights staged in the same order).
//   col2im: gather form (no atomics, deterministic) fused with the ReLU mask of the layer that produced the input.
#include "kernels.h"

namespace dtb {

// ---- conv1 patches from fp32 NCHW (k = 8, any stride with W % 4 == 0 and stride % 4 == 0) -------------------------
__global__ void __launch_bounds__(256) im2col_nchw_k8_kernel(const float* __restrict__ x, int N, int C, int H, int W, int stride, int ho,
                                                             int wo, float scale, bf16* __restrict__ p) {
  pdl_trigger();
  pdl_wait();
  const size_t total = (size_t)N * ho * wo * C * 8;  // one thread per (row, c, kh): 8 contiguous kw
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int kh = (int)(i & 7);
    size_t r = i >> 3;
    const int c = (int)(r % C);
    r /= C;  // row = (n, oh, ow)
    const int ow = (int)(r % wo);
    const size_t t = r / wo;
    const int oh = (int)(t % ho), n = (int)(t / ho);
    const float* src = x + (((size_t)n * C + c) * H + (oh * stride + kh)) * W + ow * stride;
    const float4 a = ldg_f4(src), b = ldg_f4(src + 4);
    uint4 o = make_uint4(pack_bf16x2(a.x * scale, a.y * scale), pack_bf16x2(a.z * scale, a.w * scale), pack_bf16x2(b.x * scale, b.y * scale),
                         pack_bf16x2(b.z * scale, b.w * scale));
    *reinterpret_cast<uint4*>(p + r * (size_t)(C * 64) + c * 64 + kh * 8) = o;
  }
}

int im2col_nchw_k8(cudaStream_t stream, const float* x, int N, int C, int H, int W, int stride, float scale, bf16* p) {
  DTB_REQUIRE(W % 4 == 0 && stride % 4 == 0 && H >= 8 && W >= 8, DTB_ERR_UNSUPPORTED,
              "conv1: needs W %% 4 == 0 and stride %% 4 == 0 (got W=%d stride=%d)", W, stride);
  DTB_REQUIRE((reinterpret_cast<uintptr_t>(x) & 15) == 0, DTB_ERR_BAD_ARG, "conv1: input must be 16-byte aligned");
  const int ho = (H - 8) / stride + 1, wo = (W - 8) / stride + 1;
  const size_t total = (size_t)N * ho * wo * C * 8;
  if (total == 0) return DTB_OK;
  size_t blocks = (total + 255) / 256;
  const size_t cap = (size_t)num_sms() * 16;
  if (blocks > cap) blocks = cap;
  DTB_LAUNCH(im2col_nchw_k8_kernel, (unsigned)blocks, 256, 0, stream, x, N, C, H, W, stride, ho, wo, scale, p);
  DTB_CHECK_LAUNCH();
  return DTB_OK;
}

// ---- patches from bf16 NHWC: for a fixed (row, kh) the kw*C elements are contiguous in both source and destination --
// One warp per patch row (output pixel), two rows in flight per warp: the (n, oh, ow) split is warp-uniform 32-bit arithmetic and the
// lanes walk the row's 16-byte chunks, so every load and store instruction is a contiguous run (the previous one-thread-per-chunk
// form spent ~500 integer instructions of 64-bit div / mod per 16 bytes and reached 24 % of the HBM roofline).
__global__ void __launch_bounds__(256) im2col_nhwc_kernel(const bf16* __restrict__ x, int H, int W, int C, int k, int stride, int ho,
                                                          int wo, unsigned rows, bf16* __restrict__ p) {
  pdl_trigger();
  pdl_wait();
  const unsigned run = (unsigned)(k * C) / 8u;   // 16-byte chunks per (row, kh)
  const unsigned chunks = run * (unsigned)k;     // per patch row
  const unsigned lane = threadIdx.x & 31u, nwarp = blockDim.x >> 5;
  const unsigned warp = blockIdx.x * nwarp + (threadIdx.x >> 5), wstep = gridDim.x * nwarp;
  const size_t in_row = (size_t)W * C;           // elements per input image row
  for (unsigned r0 = warp * 2u; r0 < rows; r0 += wstep * 2u) {
    const uint4* src[2];
    uint4* dst[2];
#pragma unroll
    for (int q = 0; q < 2; ++q) {
      const unsigned r = r0 + q < rows ? r0 + q : r0;
      const unsigned ow = r % (unsigned)wo, t = r / (unsigned)wo, oh = t % (unsigned)ho, n = t / (unsigned)ho;
      src[q] = reinterpret_cast<const uint4*>(x + ((size_t)n * H + oh * stride) * in_row + (size_t)ow * stride * C);
      dst[q] = reinterpret_cast<uint4*>(p + (size_t)r * chunks * 8);
    }
    const bool two = r0 + 1 < rows;
    for (unsigned j = lane; j < chunks; j += 32u) {
      const unsigned kh = j / run, off = j - kh * run;
      const size_t so = (size_t)kh * (in_row / 8) + off;
      const uint4 v0 = __ldg(src[0] + so);
      uint4 v1 = make_uint4(0, 0, 0, 0);
      if (two) v1 = __ldg(src[1] + so);
      dst[0][j] = v0;
      if (two) dst[1][j] = v1;
    }
  }
}

int im2col_nhwc(cudaStream_t stream, const bf16* x, int N, int H, int W, int C, int k, int stride, bf16* p) {
  DTB_REQUIRE(C % 8 == 0, DTB_ERR_UNSUPPORTED, "im2col: channels (%d) must be a multiple of 8", C);
  const int ho = (H - k) / stride + 1, wo = (W - k) / stride + 1;
  const size_t rows = (size_t)N * ho * wo;
  DTB_REQUIRE(rows < (1u << 31), DTB_ERR_UNSUPPORTED, "im2col: %zu patch rows exceed the 32-bit row index", rows);
  const size_t total = rows * k * (k * C / 8);
  if (total == 0) return DTB_OK;
  size_t blocks = (rows + 15) / 16;  // 8 warps x 2 rows per block and iteration
  const size_t cap = (size_t)num_sms() * 8;
  if (blocks > cap) blocks = cap;
  ProfScope prof(stream, PF_IM2COL, 2.0 * N * H * W * C + 2.0 * (double)total * 8);
  DTB_LAUNCH(im2col_nhwc_kernel, (unsigned)blocks, 256, 0, stream, x, H, W, C, k, stride, ho, wo, (unsigned)rows, p);
  DTB_CHECK_LAUNCH();
  return DTB_OK;
}

// ---- col2im (gather) + ReLU mask: dz[n,h,w,c] = (y[n,h,w,c] > 0) * sum_{kh,kw valid} dp[(n,(h-kh)/s,(w-kw)/s), (kh,kw,c)] ---------
// TK = ceil(k / stride) candidate taps per axis (kh = h % s + a s, kw = w % s + b s): the reference's layers are (k, s) = (4, 2) and
// (3, 1), i.e. TK = 2 and 3.  All TK^2 loads of a thread are issued first (predicated, independent) and summed afterwards: the previous
// data-dependent tap loops kept one 16-byte load in flight per thread and reached 18 % of the HBM roofline.  TK = 0: generic loops.
template <int TK>
__global__ void __launch_bounds__(256) col2im_nhwc_relu_kernel(const bf16* __restrict__ dp, const bf16* __restrict__ y, int N, int H, int W,
                                                               int C, int k, int stride, int ho, int wo, bf16* __restrict__ dz) {
  pdl_trigger();
  pdl_wait();
  const unsigned cg = (unsigned)C / 8u;
  const unsigned total = (unsigned)N * H * W * cg;  // < 2^31 (checked by the host): the pixel split is 32-bit arithmetic
  const size_t ldp = (size_t)k * k * C;
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const unsigned c8 = i % cg, r = i / cg;
    const int w = (int)(r % (unsigned)W);
    const unsigned t = r / (unsigned)W;
    const int h = (int)(t % (unsigned)H), n = (int)(t / (unsigned)H);
    const uint4 yu = ldg_u4(y + (size_t)r * C + c8 * 8);  // ReLU mask of the producing layer: in flight while the gather runs
    float acc[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[e] = 0.f;
    if (TK > 0) {
      uint4 u[TK > 0 ? TK * TK : 1];
      const int kh0 = h % stride, kw0 = w % stride;
#pragma unroll
      for (int a = 0; a < TK; ++a) {
        const int kh = kh0 + a * stride, oh = (h - kh) / stride;
        const bool okh = kh < k && h - kh >= 0 && oh < ho;
#pragma unroll
        for (int b = 0; b < TK; ++b) {
          const int kw = kw0 + b * stride, ow = (w - kw) / stride;
          const bool ok = okh && kw < k && w - kw >= 0 && ow < wo;
          u[a * TK + b] = ok ? ldg_u4(dp + (((size_t)n * ho + oh) * wo + ow) * ldp + ((size_t)kh * k + kw) * C + c8 * 8) : make_uint4(0, 0, 0, 0);
        }
      }
#pragma unroll
      for (int q = 0; q < TK * TK; ++q) {  // same (kh-major, kw-minor) summation order as the generic loops: bit-identical
        const float2 a = unpack_bf16x2(u[q].x), b = unpack_bf16x2(u[q].y), c = unpack_bf16x2(u[q].z), d = unpack_bf16x2(u[q].w);
        acc[0] += a.x; acc[1] += a.y; acc[2] += b.x; acc[3] += b.y; acc[4] += c.x; acc[5] += c.y; acc[6] += d.x; acc[7] += d.y;
      }
    } else {
      for (int kh = h % stride; kh < k; kh += stride) {
        const int oh = (h - kh) / stride;
        if (h - kh < 0 || oh >= ho) continue;
        for (int kw = w % stride; kw < k; kw += stride) {
          const int ow = (w - kw) / stride;
          if (w - kw < 0 || ow >= wo) continue;
          const uint4 u = ldg_u4(dp + (((size_t)n * ho + oh) * wo + ow) * ldp + ((size_t)kh * k + kw) * C + c8 * 8);
          const float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), c = unpack_bf16x2(u.z), d = unpack_bf16x2(u.w);
          acc[0] += a.x; acc[1] += a.y; acc[2] += b.x; acc[3] += b.y; acc[4] += c.x; acc[5] += c.y; acc[6] += d.x; acc[7] += d.y;
        }
      }
    }
    const float2 ya = unpack_bf16x2(yu.x), yb = unpack_bf16x2(yu.y), yc = unpack_bf16x2(yu.z), yd = unpack_bf16x2(yu.w);
    const float m[8] = {ya.x, ya.y, yb.x, yb.y, yc.x, yc.y, yd.x, yd.y};
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[e] = m[e] > 0.f ? acc[e] : 0.f;
    *reinterpret_cast<uint4*>(dz + (size_t)r * C + c8 * 8) =
        make_uint4(pack_bf16x2(acc[0], acc[1]), pack_bf16x2(acc[2], acc[3]), pack_bf16x2(acc[4], acc[5]), pack_bf16x2(acc[6], acc[7]));
  }
}

int col2im_nhwc_relu(cudaStream_t stream, const bf16* dp, const bf16* y, int N, int H, int W, int C, int k, int stride, bf16* dz) {
  DTB_REQUIRE(C % 8 == 0, DTB_ERR_UNSUPPORTED, "col2im: channels (%d) must be a multiple of 8", C);
  const int ho = (H - k) / stride + 1, wo = (W - k) / stride + 1;
  const size_t total = (size_t)N * H * W * (C / 8);
  DTB_REQUIRE(total < (1u << 31), DTB_ERR_UNSUPPORTED, "col2im: %zu pixel groups exceed the 32-bit index", total);
  if (total == 0) return DTB_OK;
  size_t blocks = (total + 255) / 256;
  const size_t cap = (size_t)num_sms() * 8;
  if (blocks > cap) blocks = cap;
  ProfScope prof(stream, PF_COL2IM, 4.0 * N * H * W * C + 2.0 * (double)N * ho * wo * k * k * C);
  const int tk = (k + stride - 1) / stride;
  if (tk == 2) DTB_LAUNCH(col2im_nhwc_relu_kernel<2>, (unsigned)blocks, 256, 0, stream, dp, y, N, H, W, C, k, stride, ho, wo, dz);
  else if (tk == 3) DTB_LAUNCH(col2im_nhwc_relu_kernel<3>, (unsigned)blocks, 256, 0, stream, dp, y, N, H, W, C, k, stride, ho, wo, dz);
  else DTB_LAUNCH(col2im_nhwc_relu_kernel<0>, (unsigned)blocks, 256, 0, stream, dp, y, N, H, W, C, k, stride, ho, wo, dz);
  DTB_CHECK_LAUNCH();
  return DTB_OK;
}

// ---- weight re-ordering: src fp32 [O][A][B] <-> [O][B][A] --------------------------------------------------------------
// stage  : dst bf16 [O][B][A] = src fp32 [O][A][B]      ([Cout][Cin][kh*kw] -> [Cout][kh*kw][Cin];  fc [F][C][h*w] -> [F][h*w][C])
// unstage: dst fp32 [O][A][B] = src fp32 [O][B][A]      (gradients back to the PyTorch order)
__global__ void permute_stage_kernel(const float* __restrict__ src, int O, int A, int B, bf16* __restrict__ dst) {
  pdl_trigger();
  pdl_wait();
  const size_t total = (size_t)O * A * B;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int a = (int)(i % A);
    const size_t r = i / A;
    const int b = (int)(r % B);
    const size_t o = r / B;
    dst[i] = __float2bfloat16(src[(o * A + a) * B + b]);
  }
}
__global__ void permute_unstage_kernel(const float* __restrict__ src, int O, int A, int B, float* __restrict__ dst) {
  pdl_trigger();
  pdl_wait();
  const size_t total = (size_t)O * A * B;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int b = (int)(i % B);
    const size_t r = i / B;
    const int a = (int)(r % A);
    const size_t o = r / A;
    dst[i] = src[(o * B + b) * A + a];
  }
}
static unsigned perm_blocks(size_t total) {
  size_t blocks = (total + 255) / 256;
  const size_t cap = (size_t)num_sms() * 8;
  return (unsigned)(blocks > cap ? cap : (blocks < 1 ? 1 : blocks));
}
int permute_stage(cudaStream_t stream, const float* src, int O, int A, int B, bf16* dst) {
  DTB_LAUNCH(permute_stage_kernel, perm_blocks((size_t)O * A * B), 256, 0, stream, src, O, A, B, dst);
  DTB_CHECK_LAUNCH();
  return DTB_OK;
}
int permute_unstage(cudaStream_t stream, const float* src, int O, int A, int B, float* dst) {
  DTB_LAUNCH(permute_unstage_kernel, perm_blocks((size_t)O * A * B), 256, 0, stream, src, O, A, B, dst);
  DTB_CHECK_LAUNCH();
  return DTB_OK;
}

// out = a + b (bf16), and the ReLU-masked variant used by the 'sum' fusion backward: dz = (y > 0) * dy
__global__ void add_bf16_kernel(const bf16* __restrict__ a, const bf16* __restrict__ b, size_t n8, bf16* __restrict__ out) {
  pdl_trigger();
  pdl_wait();
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n8; i += (size_t)gridDim.x * blockDim.x) {
    const uint4 ua = ldg_u4(a + i * 8), ub = ldg_u4(b + i * 8);
    const uint32_t wa[4] = {ua.x, ua.y, ua.z, ua.w}, wb[4] = {ub.x, ub.y, ub.z, ub.w};
    uint32_t o[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float2 x = unpack_bf16x2(wa[e]), y = unpack_bf16x2(wb[e]);
      o[e] = pack_bf16x2(x.x + y.x, x.y + y.y);
    }
    *reinterpret_cast<uint4*>(out + i * 8) = make_uint4(o[0], o[1], o[2], o[3]);
  }
}
int add_bf16(cudaStream_t stream, const bf16* a, const bf16* b, size_t n, bf16* out) {
  DTB_REQUIRE(n % 8 == 0, DTB_ERR_BAD_ARG, "add_bf16: n must be a multiple of 8");
  if (n == 0) return DTB_OK;
  DTB_LAUNCH(add_bf16_kernel, perm_blocks(n / 8), 256, 0, stream, a, b, n / 8, out);
  DTB_CHECK_LAUNCH();
  return DTB_OK;
}
__global__ void relu_mask_bf16_kernel(const bf16* __restrict__ dy, int ldy, const bf16* __restrict__ y, int ld, int rows, int cols8,
                                      bf16* __restrict__ dz, int ldz) {
  pdl_trigger();
  pdl_wait();
  const size_t total = (size_t)rows * cols8;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % cols8) * 8;
    const size_t r = i / cols8;
    const uint4 ud = ldg_u4(dy + r * ldy + c), uy = ldg_u4(y + r * ld + c);
    const uint32_t wd[4] = {ud.x, ud.y, ud.z, ud.w}, wy[4] = {uy.x, uy.y, uy.z, uy.w};
    uint32_t o[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float2 g = unpack_bf16x2(wd[e]), v = unpack_bf16x2(wy[e]);
      o[e] = pack_bf16x2(v.x > 0.f ? g.x : 0.f, v.y > 0.f ? g.y : 0.f);
    }
    *reinterpret_cast<uint4*>(dz + r * ldz + c) = make_uint4(o[0], o[1], o[2], o[3]);
  }
}
int relu_mask_bf16(cudaStream_t stream, const bf16* dy, int ldy, const bf16* y, int ld, int rows, int cols, bf16* dz, int ldz) {
  DTB_REQUIRE(cols % 8 == 0 && ldy % 8 == 0 && ld % 8 == 0 && ldz % 8 == 0, DTB_ERR_BAD_ARG, "relu_mask: dims must be multiples of 8");
  if (rows * cols == 0) return DTB_OK;
  DTB_LAUNCH(relu_mask_bf16_kernel, perm_blocks((size_t)rows * cols / 8), 256, 0, stream, dy, ldy, y, ld, rows, cols / 8, dz, ldz);
  DTB_CHECK_LAUNCH();
  return DTB_OK;
}

}  // namespace dtb
