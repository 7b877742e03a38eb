// Implicit-GEMM convolutions of the towers' inner layers on tcgen05 — forward, data gradient and weight gradient of
// Conv2d(32, 64, 4, stride=2), Conv2d(64, 64, 3, stride=1) ('big', REF src/agents.py:35-37) and Conv2d(16, 32, 4, stride=2) ('small', :48).
// Activations are channels-last bf16 ([N, H, W, C]: the previous layer's GEMM-shaped output), so for a fixed output pixel and kernel row
// the kw * C patch elements are one contiguous run of memory: producer warps copy 16-byte chunks straight into the 128B-swizzled
// shared-memory image the tensor core reads.  No patch matrix (im2col) and no patch-gradient matrix (col2im) ever exists in HBM.
//   forward : Y[rows, Cout] = relu(P W^T + b)        rows = output pixels, K = (kh, kw, c) in 64-wide k-blocks; weights resident in smem
//   dgrad   : dX[pix, Cin]  = relu'(X) * sum over the taps that reach the input pixel of dZ[out pixel(tap), :] W[:, tap, :]
//             input pixels are grouped by their parity class (ih mod stride, iw mod stride): all pixels of a class see the same set
//             of kernel taps, so each class is a dense GEMM with K = taps(class) * Cout — no multiplications by structural zeros
//   wgrad   : dW[Cout, K] and db[Cout] = sum over output pixels of dZ^T [P | 1]: P tiles gathered again (MN-major A operand = the very
//             same smem image), dZ tiles by TMA; the bias gradient is one extra accumulator row fed by a constant "ones" chunk;
//             fp32 accumulators stay in TMEM across the CTA's whole row range; per-CTA partials are reduced in CTA order and written
//             straight in PyTorch's [Cout, Cin, kh, kw] order by one small kernel.
#include "kernels.h"
#include <stdlib.h>
#include <string.h>

namespace dtb {

constexpr int CN_PRODUCERS = 256;               // warps 0-7 gather
constexpr int CN_THREADS = CN_PRODUCERS + 160;  // warps 8-11 epilogue (TMEM lane quadrant = warp & 3), warp 12 MMA + TMEM alloc
constexpr int CN_EPI_WARP0 = CN_PRODUCERS / 32, CN_MMA_WARP = CN_EPI_WARP0 + 4;
constexpr int CN_KB128 = 128 * 128;  // one k-block image of a 128-row tile
constexpr int CN_KB64 = 64 * 128;    // ... of a 64-row tile (wgrad)
constexpr int CN_MAX_CLS = 4;
constexpr int CN_STAGES = 6, CN_LAG = 4;  // fwd / dgrad k-block ring; stages announced LAG issues later

struct ConvNParams {
  const bf16* x;   // fwd / wgrad: input activation NHWC [N, hin, win, cin];  dgrad: dZ of this layer [N, ho, wo, cout]
  int N, hin, win, cin, k, stride, ho, wo, cout;
  int nkb;         // 64-wide k-blocks (fwd / wgrad: k*k*cin / 64)
  long long rows;  // GEMM rows
  int tiles, stages;
  const float* bias;
  bf16* y;          // fwd: output [rows, cout];  dgrad: dX [N*hin*win, cin]
  const bf16* act;  // dgrad: forward activation of the input layer (its ReLU mask)
  float* partial;   // wgrad: [grid][nmb * 128][64]
  int nmb;          // wgrad: 128-row accumulator blocks (pairs of 64-wide chunks: the k-blocks, then the ones chunk)
  // dgrad parity classes
  int ncls;
  int cls_tile0[CN_MAX_CLS + 1];  // first tile of each class (prefix sums)
  int cls_ph[CN_MAX_CLS], cls_pw[CN_MAX_CLS], cls_na[CN_MAX_CLS], cls_nb[CN_MAX_CLS];
  int cls_ntw[CN_MAX_CLS];   // taps along kw in this class (taps of the class are ordered (dkh, dkw))
  int cls_kb0[CN_MAX_CLS], cls_nkb[CN_MAX_CLS];  // k-block range of the class inside the staged dgrad weights
};

// 16-byte asynchronous copy global -> shared (LDGSTS); src_bytes = 0 writes zeros (structural zeros / rows beyond the tensor)
__device__ __forceinline__ void cn_cp_async16(uint32_t dst, const void* src, uint32_t src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cn_cp_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cn_cp_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ uint32_t cn_chunk_addr(uint32_t tile_base, int r, int u) { return tile_base + r * 128 + ((u ^ (r & 7)) << 4); }

__device__ __forceinline__ void cn_store_chunk(uint32_t tile_base, int r, int u, const uint4& v) {
  const uint32_t addr = tile_base + r * 128 + ((u ^ (r & 7)) << 4);
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// ================================================================================================================================
// forward + data gradient share one kernel skeleton: 128-row tiles, one k-block (128 rows x 64 k) per pipeline stage; producer lanes are
// laid out so that every load instruction reads whole 128-byte lines (first version: one row per thread = 32 lines per instruction,
// 256 mbarrier arrivals per k-block: 3.2 us per k-block at c2).
// ================================================================================================================================
template <bool DGRAD>
__global__ void __launch_bounds__(CN_THREADS, 1) convn_kernel(const __grid_constant__ CUtensorMap tmW, const ConvNParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int nout = DGRAD ? p.cin : p.cout;           // GEMM N
  const int wkb = nout * 128;                         // bytes of one weight k-block
  const int wkb_total = DGRAD ? (p.k * p.k * p.cout) / 64 : p.nkb;
  uint8_t* smW = smem;                                // [k-blocks][nout rows][128 B]
  uint8_t* smA = smem + ((wkb_total * wkb + 1023) & ~1023);  // [stages][128 rows][128 B]
  uint64_t* bars = reinterpret_cast<uint64_t*>(smA + (size_t)p.stages * CN_KB128);
  uint64_t* full_bar = bars;         // [8]
  uint64_t* empty_bar = bars + 8;    // [8]
  uint64_t* tfull_bar = bars + 16;   // [2]
  uint64_t* tempty_bar = bars + 18;  // [2]
  uint64_t* w_bar = bars + 20;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 21);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < p.stages; ++s) { mbar_init(&full_bar[s], CN_PRODUCERS / 32); mbar_init(&empty_bar[s], 1); }  // one arrival per producer warp
    for (int s = 0; s < 2; ++s) { mbar_init(&tfull_bar[s], 1); mbar_init(&tempty_bar[s], 4); }
    mbar_init(w_bar, 1);
    fence_mbar_init();
  }
  if (warp == CN_MMA_WARP) { tmem_alloc(tmem_slot, 128); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_trigger();
  pdl_wait();

  // tile -> (class, first row) ; forward has a single "class"
  auto tile_class = [&](int t) {
    int c = 0;
    if (DGRAD) while (c + 1 < p.ncls && t >= p.cls_tile0[c + 1]) ++c;
    return c;
  };

  if (warp < CN_EPI_WARP0) {
    // ===== patch producers =====
    // Warp w owns rows 16 w .. 16 w + 15 of the tile; one load instruction covers 4 rows x 8 chunks = four whole 128-byte lines
    // (lane = 8 * row + chunk), so a k-block costs every thread 4 loads (rows rq, rq + 4, rq + 8, rq + 12) and 4 conflict-free swizzled
    // stores.  The copies are asynchronous (cp.async, zero-filled where the patch has no data): a stage is announced to the MMA warp
    // CN_LAG stages after it was issued, so 4 x 4 x 16 B per thread (64 KB per SM) are in flight; one mbarrier arrival per warp.
    const int u = lane & 7, rq = warp * 16 + (lane >> 3);
    uint32_t it = 0;  // k-block stages issued so far (runs straight through tile boundaries)
    for (int t = blockIdx.x; t < p.tiles; t += gridDim.x) {
      const int cls = tile_class(t);
      const int nkb = DGRAD ? p.cls_nkb[cls] : p.nkb;
      // per row: element offset of the row's anchor in p.x (fwd: first patch element; dgrad: dZ pixel (n, a, b) channel 0), or -1
      long long base[4];
      int ra[4], rb[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int r = rq + 4 * j;
        if (!DGRAD) {
          const long long m = (long long)t * 128 + r;
          const bool ok = m < p.rows;
          const uint32_t mu = (uint32_t)(ok ? m : 0);
          const int ow = (int)(mu % (uint32_t)p.wo);
          const uint32_t q = mu / (uint32_t)p.wo;
          const int oh = (int)(q % (uint32_t)p.ho), n = (int)(q / (uint32_t)p.ho);
          base[j] = ok ? (((long long)n * p.hin + (long long)oh * p.stride) * p.win + (long long)ow * p.stride) * p.cin : -1;
          ra[j] = rb[j] = 0;
        } else {
          const long long m = (long long)(t - p.cls_tile0[cls]) * 128 + r;
          const int na = p.cls_na[cls], nb = p.cls_nb[cls];
          const bool ok = m < (long long)p.N * na * nb;
          const uint32_t mu = (uint32_t)(ok ? m : 0);
          rb[j] = (int)(mu % (uint32_t)nb);
          const uint32_t q = mu / (uint32_t)nb;
          ra[j] = (int)(q % (uint32_t)na);
          const int n = (int)(q / (uint32_t)na);
          base[j] = ok ? (((long long)n * p.ho + ra[j]) * p.wo + rb[j]) * p.cout : -1;
        }
      }
      // k-block kb of this tile -> stage: four asynchronous 16-byte copies per thread, straight into the swizzled image
      auto issue4 = [&](uint32_t tile_base, int kb) {
        if (!DGRAD) {
          const int tap0 = kb * (64 / p.cin);  // the k-block's taps are consecutive kw of one kernel row: contiguous in NHWC
          const int kh = tap0 / p.k, kw0 = tap0 - kh * p.k;
          const long long off = ((long long)kh * p.win + kw0) * p.cin + u * 8;
#pragma unroll
          for (int j = 0; j < 4; ++j)
            cn_cp_async16(cn_chunk_addr(tile_base, rq + 4 * j, u), base[j] >= 0 ? p.x + base[j] + off : p.x, base[j] >= 0 ? 16u : 0u);
        } else {
          const int e0 = kb * 64 + u * 8;  // K index (tap, cout) of this lane's chunk
          const int tt = e0 / p.cout, co = e0 - tt * p.cout;
          const int ntw = p.cls_ntw[cls];
          const int dkh = tt / ntw, dkw = tt - dkh * ntw;  // kh = ph + dkh * stride  ->  oh = (ih - kh) / stride = a - dkh
          const long long off = -((long long)dkh * p.wo + dkw) * p.cout + co;
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const int oh = ra[j] - dkh, ow = rb[j] - dkw;
            const bool ok = base[j] >= 0 && oh >= 0 && ow >= 0 && oh < p.ho && ow < p.wo;  // else: the tap falls outside dZ (structural zero)
            cn_cp_async16(cn_chunk_addr(tile_base, rq + 4 * j, u), ok ? p.x + base[j] + off : p.x, ok ? 16u : 0u);
          }
        }
      };
      for (int kb = 0; kb < nkb; ++kb, ++it) {
        const uint32_t stage = it % CN_STAGES, phase = (it / CN_STAGES) & 1;
        mbar_wait(&empty_bar[stage], phase ^ 1);
        issue4(smem_u32(smA + (size_t)stage * CN_KB128), kb);
        cn_cp_commit();
        if (it >= CN_LAG) {  // the copies of stage it - LAG have landed (at most LAG younger groups are still in flight)
          cn_cp_wait<CN_LAG>();
          fence_proxy_async();  // generic-proxy writes -> visible to the tensor core (async proxy)
          __syncwarp();
          if (lane == 0) mbar_arrive(&full_bar[(it - CN_LAG) % CN_STAGES]);
        }
      }
    }
    // drain: the last LAG stages
    cn_cp_wait<0>();
    fence_proxy_async();
    __syncwarp();
    if (lane == 0)
      for (uint32_t j = it > CN_LAG ? it - CN_LAG : 0; j < it; ++j) mbar_arrive(&full_bar[j % CN_STAGES]);
  } else if (warp == CN_MMA_WARP) {
    if (lane == 0) {
      tma_prefetch_desc(&tmW);
      mbar_arrive_expect_tx(w_bar, (uint32_t)(wkb_total * wkb));
      for (int c = 0; c < wkb_total; ++c) tma_load_2d(smW + (size_t)c * wkb, &tmW, w_bar, c * 64, 0);
      mbar_wait(w_bar, 0);
      const uint32_t idesc = make_idesc_bf16(128, nout, 0, 0);
      uint32_t stage = 0, phase = 0, acc = 0, acc_phase = 0;
      for (int t = blockIdx.x; t < p.tiles; t += gridDim.x) {
        const int cls = tile_class(t);
        const int nkb = DGRAD ? p.cls_nkb[cls] : p.nkb, kb0 = DGRAD ? p.cls_kb0[cls] : 0;
        mbar_wait(&tempty_bar[acc], acc_phase ^ 1);
        tc_fence_after();
        for (int kb = 0; kb < nkb; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t a0 = smem_u32(smA + (size_t)stage * CN_KB128), w0 = smem_u32(smW + (size_t)(kb0 + kb) * wkb);
#pragma unroll
          for (int kk = 0; kk < 4; ++kk)
            umma_bf16(tmem_base + acc * 64, make_smem_desc_sw128(a0 + kk * 32, 16, 1024), make_smem_desc_sw128(w0 + kk * 32, 16, 1024), idesc,
                      (kb > 0 || kk > 0) ? 1u : 0u);
          umma_commit(&empty_bar[stage]);
          if (++stage == (uint32_t)p.stages) { stage = 0; phase ^= 1; }
        }
        umma_commit(&tfull_bar[acc]);
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
  } else {
    // ===== epilogue: one accumulator row per thread =====
    const int q = warp & 3, r = q * 32 + lane;
    uint32_t acc = 0, acc_phase = 0;
    for (int t = blockIdx.x; t < p.tiles; t += gridDim.x) {
      const int cls = tile_class(t);
      long long out_row = -1;
      if (!DGRAD) {
        const long long m = (long long)t * 128 + r;
        if (m < p.rows) out_row = m;
      } else {
        const long long m = (long long)(t - p.cls_tile0[cls]) * 128 + r;
        const int na = p.cls_na[cls], nb = p.cls_nb[cls];
        if (m < (long long)p.N * na * nb) {
          const uint32_t mu = (uint32_t)m;
          const int b = (int)(mu % (uint32_t)nb);
          const uint32_t qq = mu / (uint32_t)nb;
          const int a = (int)(qq % (uint32_t)na), n = (int)(qq / (uint32_t)na);
          out_row = ((long long)n * p.hin + (a * p.stride + p.cls_ph[cls])) * p.win + (b * p.stride + p.cls_pw[cls]);
        }
      }
      mbar_wait(&tfull_bar[acc], acc_phase);
      tc_fence_after();
      for (int c = 0; c < nout; c += 32) {
        uint32_t v[32];
        tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + acc * 64 + c, v);
        tmem_ld_wait();
        if (c + 32 >= nout) {  // accumulator fully read: hand it back before the global traffic
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&tempty_bar[acc]);
        }
        if (out_row >= 0) {
          bf16* dst = p.y + out_row * nout + c;
#pragma unroll
          for (int j = 0; j < 32; j += 8) {
            if (c + j >= nout) break;  // nout = 16: half a chunk
            float f[8];
            if (!DGRAD) {
#pragma unroll
              for (int e = 0; e < 8; ++e) f[e] = fmaxf(__uint_as_float(v[j + e]) + __ldg(p.bias + c + j + e), 0.f);
            } else {
              const uint4 m4 = __ldg(reinterpret_cast<const uint4*>(p.act + out_row * nout + c + j));
              const uint32_t mw[4] = {m4.x, m4.y, m4.z, m4.w};
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                const float2 xa = unpack_bf16x2(mw[e]);
                f[2 * e] = xa.x > 0.f ? __uint_as_float(v[j + 2 * e]) : 0.f;
                f[2 * e + 1] = xa.y > 0.f ? __uint_as_float(v[j + 2 * e + 1]) : 0.f;
              }
            }
            *reinterpret_cast<uint4*>(dst + j) = make_uint4(pack_bf16x2(f[0], f[1]), pack_bf16x2(f[2], f[3]), pack_bf16x2(f[4], f[5]), pack_bf16x2(f[6], f[7]));
          }
        }
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == CN_MMA_WARP) { tc_fence_after(); tmem_dealloc(tmem_base, 128); }
}

// ================================================================================================================================
// weight (+ bias) gradient.  64 output pixels per pipeline stage (two stages); a stage holds all k-blocks of the patch rows, the dZ
// rows (TMA) — and, constant for the whole kernel, a "ones" chunk (element (row, m = 0) = 1) and a zero chunk.
//   A (MN-major, M = 128) = two 64-wide chunks out of [k-block 0, ..., k-block nkb-1, ones, zero]   (LBO = distance between the two)
//   B (MN-major, N = 64)  = dZ rows (columns >= cout arrive as zeros)
//   D[mb] (TMEM columns mb * 64 ..) : lanes = K index (kh, kw, c) of the block, columns = cout; lane `nkb * 64` of it all = db
// ================================================================================================================================
__global__ void __launch_bounds__(CN_THREADS, 1) convn_wgrad_kernel(const __grid_constant__ CUtensorMap tmZ, const ConvNParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int stage_bytes = (p.nkb + 1) * CN_KB64;  // patch k-blocks + the dZ tile
  uint8_t* smS = smem;                                       // [2][nkb patch images | dZ image]
  uint8_t* smOnes = smS + 2 * (size_t)stage_bytes;
  uint8_t* smZero = smOnes + CN_KB64;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smZero + CN_KB64);
  uint64_t* full_bar = bars;       // [2] producers (256) + TMA expect_tx arrival
  uint64_t* empty_bar = bars + 2;  // [2]
  uint64_t* done_bar = bars + 4;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 5);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < 2 * CN_KB64 / 16; i += blockDim.x) reinterpret_cast<uint4*>(smOnes)[i] = make_uint4(0, 0, 0, 0);
  __syncthreads();
  if (threadIdx.x < 64) {  // element (row r, m = 0): first bf16 of 16-byte unit 0, which the 128B swizzle places at unit (0 ^ (r & 7))
    const int r = threadIdx.x;
    *reinterpret_cast<uint16_t*>(smOnes + r * 128 + ((r & 7) << 4)) = 0x3F80;  // bf16 1.0
  }
  if (threadIdx.x == 0) {
    for (int s = 0; s < 2; ++s) { mbar_init(&full_bar[s], CN_PRODUCERS / 32 + 1); mbar_init(&empty_bar[s], 1); }  // producer warps + the TMA's expect_tx arrival
    mbar_init(done_bar, 1);
    fence_mbar_init();
  }
  if (warp == CN_MMA_WARP) { tmem_alloc(tmem_slot, 512); tmem_relinquish(); }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_trigger();
  pdl_wait();
  const int my_tiles = p.tiles > (int)blockIdx.x ? (p.tiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;

  if (warp < CN_EPI_WARP0) {
    // warp w owns patch rows 8 w .. 8 w + 7 of the 64-row tile; a load instruction covers 4 rows x 8 chunks (whole 128-byte lines)
    const int u = lane & 7, rq = warp * 8 + (lane >> 3);
    uint32_t stage = 0, phase = 0, issued = 0;
    for (int t = blockIdx.x; t < p.tiles; t += gridDim.x) {
      long long base[2];
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const long long m = (long long)t * 64 + rq + 4 * j;
        const bool ok = m < p.rows;
        const uint32_t mu = (uint32_t)(ok ? m : 0);
        const int ow = (int)(mu % (uint32_t)p.wo);
        const uint32_t q = mu / (uint32_t)p.wo;
        const int oh = (int)(q % (uint32_t)p.ho), n = (int)(q / (uint32_t)p.ho);
        base[j] = ok ? (((long long)n * p.hin + (long long)oh * p.stride) * p.win + (long long)ow * p.stride) * p.cin + u * 8 : -1;
      }
      mbar_wait(&empty_bar[stage], phase ^ 1);
      uint8_t* st = smS + (size_t)stage * stage_bytes;
      if (threadIdx.x == 0) {  // dZ rows of this tile: box [64 columns][64 rows]; columns >= cout and rows >= rows arrive as zeros
        mbar_arrive_expect_tx(&full_bar[stage], CN_KB64);
        tma_load_2d(st + (size_t)p.nkb * CN_KB64, &tmZ, &full_bar[stage], 0, t * 64);
      }
      for (int kb = 0; kb < p.nkb; ++kb) {  // all k-blocks of the tile: 2 asynchronous 16-byte copies per thread and k-block
        const int tap0 = kb * (64 / p.cin);
        const int kh = tap0 / p.k, kw0 = tap0 - kh * p.k;
        const long long off = ((long long)kh * p.win + kw0) * p.cin;
        const uint32_t img = smem_u32(st + (size_t)kb * CN_KB64);
#pragma unroll
        for (int j = 0; j < 2; ++j)
          cn_cp_async16(cn_chunk_addr(img, rq + 4 * j, u), base[j] >= 0 ? p.x + base[j] + off : p.x, base[j] >= 0 ? 16u : 0u);
      }
      cn_cp_commit();
      if (issued > 0) {  // announce the PREVIOUS tile's stage: its copies have landed once at most one younger group is pending
        cn_cp_wait<1>();
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) mbar_arrive(&full_bar[stage ^ 1]);
      }
      ++issued;
      if (++stage == 2) { stage = 0; phase ^= 1; }
    }
    if (issued > 0) {
      cn_cp_wait<0>();
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(&full_bar[stage ^ 1]);
    }
  } else if (warp == CN_MMA_WARP) {
    if (lane == 0) {
      tma_prefetch_desc(&tmZ);
      const uint32_t idesc = make_idesc_bf16(128, 64, 1, 1);
      const uint32_t ones_addr = smem_u32(smOnes), zero_addr = smem_u32(smZero);
      uint32_t stage = 0, phase = 0;
      for (int i = 0; i < my_tiles; ++i) {
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        const uint32_t s0 = smem_u32(smS + (size_t)stage * stage_bytes), z0 = s0 + (uint32_t)p.nkb * CN_KB64;
        for (int mb = 0; mb < p.nmb; ++mb) {
          const int c0 = 2 * mb, c1 = 2 * mb + 1;  // chunk list: k-blocks, then ones, then zero
          const uint32_t a0 = c0 < p.nkb ? s0 + c0 * CN_KB64 : (c0 == p.nkb ? ones_addr : zero_addr);
          const uint32_t a1 = c1 < p.nkb ? s0 + c1 * CN_KB64 : (c1 == p.nkb ? ones_addr : zero_addr);
#pragma unroll
          for (int ks = 0; ks < 4; ++ks)  // 16 patch rows per MMA
            umma_bf16(tmem_base + mb * 64, make_smem_desc_sw128(a0 + ks * 2048, a1 - a0, 1024), make_smem_desc_sw128(z0 + ks * 2048, 16, 1024),
                      idesc, (i > 0 || ks > 0) ? 1u : 0u);
        }
        umma_commit(&empty_bar[stage]);
        if (++stage == 2) { stage = 0; phase ^= 1; }
      }
      umma_commit(done_bar);
    }
  } else {
    const int q = warp & 3;
    mbar_wait(done_bar, 0);
    tc_fence_after();
    for (int mb = 0; mb < p.nmb; ++mb) {
      float* dst = p.partial + (((size_t)blockIdx.x * p.nmb + mb) * 128 + q * 32 + lane) * 64;
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        uint32_t v[32];
        tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + mb * 64 + h * 32, v);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; j += 4)
          *reinterpret_cast<float4*>(dst + h * 32 + j) =
              my_tiles > 0 ? make_float4(__uint_as_float(v[j]), __uint_as_float(v[j + 1]), __uint_as_float(v[j + 2]), __uint_as_float(v[j + 3]))
                           : make_float4(0.f, 0.f, 0.f, 0.f);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == CN_MMA_WARP) { tc_fence_after(); tmem_dealloc(tmem_base, 512); }
}

// partial [ncta][nmb * 128][64] -> dW in PyTorch order [cout][cin][kh][kw] (+ db).  32 outputs x 8 slice-lanes per block: lane l sums CTAs
// l, l + 8, ... (independent loads), the 8 lanes are combined in index order (deterministic).
__global__ void __launch_bounds__(256) convn_wgrad_reduce_kernel(const float* __restrict__ partial, int ncta, int nmb, int nkb, int cin, int k,
                                                                 int cout, float* __restrict__ dw, float* __restrict__ db) {
  pdl_trigger();
  pdl_wait();
  __shared__ float red[8][33];
  const int rows = nkb * 64 + 1;  // K indices + the ones row
  const size_t per = (size_t)nmb * 128 * 64;
  const int i = blockIdx.x * 32 + threadIdx.x;
  const int m = i >> 6, co = i & 63;
  const bool ok = m < rows && co < cout;
  float s = 0.f;
  if (ok)
    for (int c = threadIdx.y; c < ncta; c += 8) s += __ldg(partial + (size_t)c * per + (size_t)m * 64 + co);
  red[threadIdx.y][threadIdx.x] = s;
  __syncthreads();
  if (threadIdx.y == 0 && ok) {
    for (int l = 1; l < 8; ++l) s += red[l][threadIdx.x];
    if (m == nkb * 64) {
      db[co] = s;
    } else {
      const int tap = m / cin, ci = m - tap * cin;  // staged K order (kh, kw, c)
      dw[((size_t)co * cin + ci) * (k * k) + tap] = s;
    }
  }
}

// Staged data-gradient weights: for each parity class, the matrix B[cin][(tap in class, cout)] (K-major, K contiguous), classes side by side
// along K in the order of ConvNParams' class table.  src = fp32 PyTorch weight [cout][cin][k][k].
__global__ void __launch_bounds__(256) convn_stage_dgrad_kernel(const float* __restrict__ w, int cin, int cout, int k, int stride,
                                                                bf16* __restrict__ dst) {
  pdl_trigger();
  pdl_wait();
  const int ktot = k * k * cout;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < cin * ktot; i += gridDim.x * blockDim.x) {
    const int ci = i / ktot;
    int col = i - ci * ktot;
    // walk the classes (ph, pw) in order until `col` falls into one
    int kh = 0, kw = 0, co = 0;
    for (int ph = 0; ph < stride; ++ph)
      for (int pw = 0; pw < stride; ++pw) {
        const int nth = (k - ph + stride - 1) / stride, ntw = (k - pw + stride - 1) / stride;
        const int kc = nth * ntw * cout;
        if (col >= 0 && col < kc) {
          const int tt = col / cout;
          co = col - tt * cout;
          kh = ph + (tt / ntw) * stride;
          kw = pw + (tt % ntw) * stride;
          col = -1;
        } else if (col >= 0) {
          col -= kc;
        }
      }
    dst[i] = __float2bfloat16(w[(((size_t)co * cin + ci) * k + kh) * k + kw]);
  }
}

// ---- host ------------------------------------------------------------------------------------------------------------------------
bool convn_supported(int cin, int cout, int k, int stride) {
  if (!opt_conv_implicit()) return false;  // dtb_set_option("conv_implicit", 0) / DTB_CONV_EXPLICIT=1: the im2col + GEMM + col2im path
  if (!(cin == 16 || cin == 32 || cin == 64) || !(cout == 32 || cout == 64)) return false;
  if ((k * k * cin) % 64 != 0 || k % (64 / cin) != 0 || stride < 1 || stride > 2 || k > 8) return false;
  // data gradient: every parity class must fill whole k-blocks
  for (int ph = 0; ph < stride; ++ph)
    for (int pw = 0; pw < stride; ++pw) {
      const int nth = (k - ph + stride - 1) / stride, ntw = (k - pw + stride - 1) / stride;
      if ((nth * ntw * cout) % 64 != 0) return false;
    }
  return true;
}

static void cn_fill(ConvNParams& p, const bf16* x, int N, int hin, int win, int cin, int k, int stride, int cout) {
  memset(&p, 0, sizeof(p));
  p.x = x; p.N = N; p.hin = hin; p.win = win; p.cin = cin; p.k = k; p.stride = stride; p.cout = cout;
  p.ho = (hin - k) / stride + 1; p.wo = (win - k) / stride + 1;
  p.nkb = k * k * cin / 64;
}

static int cn_attr() {
  static bool attr = false;
  if (!attr) {
    DTB_CUDA(cudaFuncSetAttribute(convn_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 225 * 1024));
    DTB_CUDA(cudaFuncSetAttribute(convn_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 225 * 1024));
    DTB_CUDA(cudaFuncSetAttribute(convn_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 225 * 1024));
    attr = true;
  }
  return DTB_OK;
}

// y[rows, cout] = relu(conv(x) + bias);  w_staged bf16 [cout][(kh, kw, c)]
int convn_fwd(cudaStream_t stream, const bf16* x, int N, int hin, int win, int cin, int k, int stride, const bf16* w_staged, const float* bias,
              int cout, bf16* y) {
  ConvNParams p;
  cn_fill(p, x, N, hin, win, cin, k, stride, cout);
  p.rows = (long long)N * p.ho * p.wo;
  if (p.rows == 0) return DTB_OK;
  DTB_REQUIRE(p.rows < (1LL << 31), DTB_ERR_UNSUPPORTED, "conv: %lld output pixels exceed the 32-bit row index", p.rows);
  p.tiles = (int)((p.rows + 127) / 128);
  p.bias = bias; p.y = y;
  const int wbytes = (p.nkb * cout * 128 + 1023) & ~1023;
  p.stages = CN_STAGES;
  DTB_REQUIRE(wbytes + CN_STAGES * CN_KB128 <= 200 * 1024, DTB_ERR_UNSUPPORTED, "conv: weights too large for shared memory");
  CUtensorMap tmW;
  DTB_TRY(make_tmap(&tmW, w_staged, (uint64_t)k * k * cin, (uint64_t)cout, (uint64_t)k * k * cin, 64, (uint32_t)cout));
  DTB_TRY(cn_attr());
  const int smem = 1024 + wbytes + p.stages * CN_KB128 + 512;
  const int grid = p.tiles < num_sms() ? p.tiles : num_sms();
  ProfScope prof(stream, PF_CONVN_FWD, 2.0 * ((double)N * hin * win * cin + (double)p.rows * cout));
  DTB_LAUNCH(convn_kernel<false>, grid, CN_THREADS, smem, stream, tmW, p);
  DTB_CHECK_LAUNCH();
  return DTB_OK;
}

size_t convn_dgrad_weight_elems(int cin, int cout, int k) { return (size_t)cin * k * k * cout; }

int convn_stage_dgrad(cudaStream_t stream, const float* w, int cin, int cout, int k, int stride, bf16* dst) {
  const int total = cin * k * k * cout;
  DTB_LAUNCH(convn_stage_dgrad_kernel, (total + 255) / 256, 256, 0, stream, w, cin, cout, k, stride, dst);
  DTB_CHECK_LAUNCH();
  return DTB_OK;
}

// dx[N*hin*win, cin] = relu'(act) * conv_transpose(dz, W);  wd_staged from convn_stage_dgrad
int convn_dgrad(cudaStream_t stream, const bf16* dz, int N, int hin, int win, int cin, int k, int stride, const bf16* wd_staged, int cout,
                const bf16* act, bf16* dx) {
  ConvNParams p;
  cn_fill(p, dz, N, hin, win, cin, k, stride, cout);
  if (N == 0) return DTB_OK;
  p.act = act; p.y = dx;
  int tile0 = 0, kb0 = 0;
  p.ncls = 0;
  for (int ph = 0; ph < stride; ++ph)
    for (int pw = 0; pw < stride; ++pw) {
      const int c = p.ncls++;
      const int nth = (k - ph + stride - 1) / stride, ntw = (k - pw + stride - 1) / stride;
      p.cls_ph[c] = ph; p.cls_pw[c] = pw;
      p.cls_na[c] = (hin - ph + stride - 1) / stride;
      p.cls_nb[c] = (win - pw + stride - 1) / stride;
      p.cls_ntw[c] = ntw;
      p.cls_kb0[c] = kb0;
      p.cls_nkb[c] = nth * ntw * cout / 64;
      kb0 += p.cls_nkb[c];
      p.cls_tile0[c] = tile0;
      const long long rows = (long long)N * p.cls_na[c] * p.cls_nb[c];
      DTB_REQUIRE(rows < (1LL << 31), DTB_ERR_UNSUPPORTED, "conv dgrad: too many pixels");
      tile0 += (int)((rows + 127) / 128);
    }
  p.cls_tile0[p.ncls] = tile0;
  p.tiles = tile0;
  const int wkb_total = k * k * cout / 64;
  const int wbytes = (wkb_total * cin * 128 + 1023) & ~1023;
  p.stages = CN_STAGES;
  DTB_REQUIRE(wbytes + CN_STAGES * CN_KB128 <= 200 * 1024, DTB_ERR_UNSUPPORTED, "conv dgrad: weights too large for shared memory");
  CUtensorMap tmW;
  DTB_TRY(make_tmap(&tmW, wd_staged, (uint64_t)k * k * cout, (uint64_t)cin, (uint64_t)k * k * cout, 64, (uint32_t)cin));
  DTB_TRY(cn_attr());
  const int smem = 1024 + wbytes + p.stages * CN_KB128 + 512;
  const int grid = p.tiles < num_sms() ? p.tiles : num_sms();
  ProfScope prof(stream, PF_CONVN_DGRAD, 2.0 * ((double)N * p.ho * p.wo * cout + 2.0 * (double)N * hin * win * cin));
  DTB_LAUNCH(convn_kernel<true>, grid, CN_THREADS, smem, stream, tmW, p);
  DTB_CHECK_LAUNCH();
  return DTB_OK;
}

static int cn_wgrad_grid(long long rows) {
  const long long tiles = (rows + 63) / 64;
  long long g = (tiles + 3) / 4;  // at least ~4 tiles per CTA: fewer partial slices to write and reduce
  if (g > num_sms()) g = num_sms();
  if (g > DTB_NUM_SMS_FALLBACK) g = DTB_NUM_SMS_FALLBACK;
  return (int)(g < 1 ? 1 : g);
}
static int cn_nmb(int nkb) { return (nkb + 1 + 1) / 2; }  // chunks: nkb k-blocks + ones (+ zero padding)
size_t convn_wgrad_partial_floats(int cin, int k) { return (size_t)DTB_NUM_SMS_FALLBACK * cn_nmb(k * k * cin / 64) * 128 * 64; }

// dw [cout][cin][k][k] (PyTorch order) and db [cout] from dz [rows, cout] and the layer input x (NHWC)
int convn_wgrad(cudaStream_t stream, const bf16* x, int N, int hin, int win, int cin, int k, int stride, const bf16* dz, int cout,
                float* partial, float* dw, float* db) {
  ConvNParams p;
  cn_fill(p, x, N, hin, win, cin, k, stride, cout);
  p.rows = (long long)N * p.ho * p.wo;
  if (p.rows == 0) {
    DTB_CUDA(cudaMemsetAsync(dw, 0, (size_t)cout * cin * k * k * sizeof(float), stream));
    DTB_CUDA(cudaMemsetAsync(db, 0, (size_t)cout * sizeof(float), stream));
    return DTB_OK;
  }
  DTB_REQUIRE(p.rows < (1LL << 31), DTB_ERR_UNSUPPORTED, "conv wgrad: too many pixels");
  p.tiles = (int)((p.rows + 63) / 64);
  p.partial = partial;
  p.nmb = cn_nmb(p.nkb);
  DTB_REQUIRE(p.nmb * 64 <= 512, DTB_ERR_UNSUPPORTED, "conv wgrad: K = %d exceeds the TMEM accumulator", k * k * cin);
  const int smem = 1024 + 2 * (p.nkb + 1) * CN_KB64 + 2 * CN_KB64 + 512;
  DTB_REQUIRE(smem <= 225 * 1024, DTB_ERR_UNSUPPORTED, "conv wgrad: K = %d exceeds shared memory", k * k * cin);
  CUtensorMap tmZ;
  DTB_TRY(make_tmap(&tmZ, dz, (uint64_t)cout, (uint64_t)p.rows, (uint64_t)cout, 64, 64));
  DTB_TRY(cn_attr());
  const int grid = cn_wgrad_grid(p.rows);
  {
    ProfScope prof(stream, PF_CONVN_WGRAD, 2.0 * ((double)N * hin * win * cin + (double)p.rows * cout));
    DTB_LAUNCH(convn_wgrad_kernel, grid, CN_THREADS, smem, stream, tmZ, p);
    DTB_CHECK_LAUNCH();
  }
  const int total = (p.nkb * 64 + 1) * 64;
  DTB_LAUNCH(convn_wgrad_reduce_kernel, (total + 31) / 32, dim3(32, 8), 0, stream, partial, grid, p.nmb, p.nkb, cin, k, cout, dw, db);
  DTB_CHECK_LAUNCH();
  return DTB_OK;
}

}  // namespace dtb
