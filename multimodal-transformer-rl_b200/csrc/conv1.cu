// Implicit-GEMM first convolution of the frame / spectrogram towers on tcgen05 (REF src/agents.py:33, :46: Conv2d(C, 32|16, 8, stride=4)).
// The 8x8xC patches are never written to HBM: producer warps read the fp32 NCHW input (the reference's obs layout), apply the
// `/255` input scale, convert to bf16 and store straight into the 128B-swizzled shared-memory tile the tensor core reads;
// K order (c, kh, kw) = PyTorch's weight order, one 64-wide k-block (= one 128-byte swizzle row) per input channel.
//   forward : Y[rows, cout] = relu(P W^T + b)          P tile = A operand (K-major), weights resident in smem via TMA
//   wgrad   : dW[cout, 64C] = dZ^T P                   P tile = B operand (MN-major: the very same smem image), dZ^T via TMA,
//             fp32 accumulators for all C k-blocks stay in TMEM across the CTA's whole row range; per-CTA partials reduced in order.
// HBM-bound by construction (N = 32): algorithmic bytes = input read once + output written once.
#include "kernels.h"
#include <string.h>

namespace dtb {

constexpr int C1_PRODUCERS = 384;  // warps 0-11 gather patches: enough 128-bit loads in flight per SM to cover HBM latency
constexpr int C1_THREADS = C1_PRODUCERS + 160;  // warps 12-15 epilogue (TMEM quadrant = warp & 3), warp 16 MMA + TMEM alloc
constexpr int C1_EPI_WARP0 = C1_PRODUCERS / 32, C1_MMA_WARP = C1_EPI_WARP0 + 4;
constexpr int C1_MAXC = 4;
constexpr int C1_KB_BYTES = 128 * 128;  // one k-block of a 128-row tile

struct Conv1Params {
  const float* x;
  const uint8_t* xu8;  // non-null: uint8 pixels, decoded as p*2/255-1 (REF environments/Minecraft/Minecraft.py:333-335), x unused
  int N, C, H, W, stride, ho, wo, cout;
  long long rows;
  int tiles, stages;
  float scale;
  const float* bias;
  bf16* y;          // forward output [rows, cout]
  float* partial;   // wgrad: [grid][cout][64C]
};

// Gather a 128-row patch tile (all C k-blocks) into the swizzled image: row r of k-block c lives at c*16KB + r*128, 16-byte chunk kh at
// (kh ^ (r & 7)).  Work item = (row, channel, half of the 8 kernel rows): 8 independent 128-bit loads, then 4 shared stores.
__device__ __forceinline__ void c1_gather_tile(const Conv1Params& p, uint8_t* stage, long long m0) {
  const uint32_t base = smem_u32(stage);
  constexpr int NB = 2;  // work items per thread whose loads are all issued before the first conversion (16 x 128-bit in flight)
  const int nitems = 256 * p.C;
  for (int it0 = threadIdx.x; it0 < nitems; it0 += C1_PRODUCERS * NB) {
    float4 a[NB][4], b[NB][4];
#pragma unroll
    for (int j = 0; j < NB; ++j) {
      const int item = it0 + j * C1_PRODUCERS;
      const int r = item & 127, half = (item >> 7) & 1, c = item >> 8;
      const long long m = m0 + r;
      if (item < nitems && m < p.rows) {
        const uint32_t mu = (uint32_t)m;  // rows < 2^31 (checked on the host): 32-bit index arithmetic
        const int ow = (int)(mu % (uint32_t)p.wo);
        const uint32_t t = mu / (uint32_t)p.wo;
        const int oh = (int)(t % (uint32_t)p.ho), n = (int)(t / (uint32_t)p.ho);
        const float* src = p.x + (((size_t)n * p.C + c) * p.H + (size_t)oh * p.stride + half * 4) * p.W + (size_t)ow * p.stride;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          a[j][k] = ldg_f4(src + (size_t)k * p.W);
          b[j][k] = ldg_f4(src + (size_t)k * p.W + 4);
        }
      } else {
#pragma unroll
        for (int k = 0; k < 4; ++k) a[j][k] = b[j][k] = make_float4(0.f, 0.f, 0.f, 0.f);
      }
    }
    const float s = p.scale;
#pragma unroll
    for (int j = 0; j < NB; ++j) {
      const int item = it0 + j * C1_PRODUCERS;
      if (item >= nitems) break;
      const int r = item & 127, half = (item >> 7) & 1, c = item >> 8;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int kh = half * 4 + k;
        const uint32_t addr = base + c * C1_KB_BYTES + r * 128 + ((kh ^ (r & 7)) << 4);
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(pack_bf16x2(a[j][k].x * s, a[j][k].y * s)),
                     "r"(pack_bf16x2(a[j][k].z * s, a[j][k].w * s)), "r"(pack_bf16x2(b[j][k].x * s, b[j][k].y * s)),
                     "r"(pack_bf16x2(b[j][k].z * s, b[j][k].w * s))
                     : "memory");
      }
    }
  }
}

// uint8-pixel variant: the reference's pixel decode (img * 2 / 255 - 1, fp32) and the `/255` input scale in registers.  The IEEE
// division n / 255 is reproduced exactly by a reciprocal multiply plus one FMA residual correction (verified for all 256 pixel values:
// tests/test_encoder_gpu.py::test_uint8_pixel_frames_bit_exact), so the tile image is bit-identical to the fp32-input path.
__device__ __forceinline__ float c1_decode_px(uint32_t byte, float scale) {
  const float n = (float)byte * 2.f;  // exact
  const float r = 1.f / 255.f;
  float q = __fmul_rn(n, r);
  const float rem = __fmaf_rn(-q, 255.f, n);
  q = __fmaf_rn(rem, r, q);  // == n / 255 correctly rounded
  return __fmul_rn(__fsub_rn(q, 1.f), scale);
}
__device__ __forceinline__ uint32_t c1_decode_pair(uint32_t word, int lo_byte, float scale) {
  return pack_bf16x2(c1_decode_px((word >> (8 * lo_byte)) & 255u, scale), c1_decode_px((word >> (8 * lo_byte + 8)) & 255u, scale));
}
__device__ __forceinline__ void c1_gather_tile_u8(const Conv1Params& p, uint8_t* stage, long long m0) {
  const uint32_t base = smem_u32(stage);
  constexpr int NB = 3;  // all of a thread's work items of the tile (C <= 4: at most 3): every load is in flight before the first lookup
  const int nitems = 256 * p.C;
  for (int it0 = threadIdx.x; it0 < nitems; it0 += C1_PRODUCERS * NB) {
    uint32_t a[NB][4], b[NB][4];
    bool ok[NB];
#pragma unroll
    for (int j = 0; j < NB; ++j) {
      const int item = it0 + j * C1_PRODUCERS;
      const int r = item & 127, half = (item >> 7) & 1, c = item >> 8;
      const long long m = m0 + r;
      ok[j] = item < nitems && m < p.rows;
      if (ok[j]) {
        const uint32_t mu = (uint32_t)m;  // rows < 2^31 (checked on the host): 32-bit index arithmetic
        const int ow = (int)(mu % (uint32_t)p.wo);
        const uint32_t t = mu / (uint32_t)p.wo;
        const int oh = (int)(t % (uint32_t)p.ho), n = (int)(t / (uint32_t)p.ho);
        const uint8_t* src = p.xu8 + (((size_t)n * p.C + c) * p.H + (size_t)oh * p.stride + half * 4) * p.W + (size_t)ow * p.stride;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          a[j][k] = __ldg(reinterpret_cast<const uint32_t*>(src + (size_t)k * p.W));
          b[j][k] = __ldg(reinterpret_cast<const uint32_t*>(src + (size_t)k * p.W + 4));
        }
      } else {
#pragma unroll
        for (int k = 0; k < 4; ++k) a[j][k] = b[j][k] = 0u;
      }
    }
#pragma unroll
    for (int j = 0; j < NB; ++j) {
      const int item = it0 + j * C1_PRODUCERS;
      if (item >= nitems) break;
      const int r = item & 127, half = (item >> 7) & 1, c = item >> 8;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int kh = half * 4 + k;
        const uint32_t addr = base + c * C1_KB_BYTES + r * 128 + ((kh ^ (r & 7)) << 4);
        uint32_t w0 = 0u, w1 = 0u, w2 = 0u, w3 = 0u;
        if (ok[j]) {
          const uint32_t av = a[j][k], bv = b[j][k];
          w0 = c1_decode_pair(av, 0, p.scale); w1 = c1_decode_pair(av, 2, p.scale);
          w2 = c1_decode_pair(bv, 0, p.scale); w3 = c1_decode_pair(bv, 2, p.scale);
        }
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(w0), "r"(w1), "r"(w2), "r"(w3) : "memory");
      }
    }
  }
}

__global__ void __launch_bounds__(C1_THREADS, 1) conv1_fwd_kernel(const __grid_constant__ CUtensorMap tmW, const Conv1Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int wkb = p.cout * 128;  // bytes of one weight k-block
  uint8_t* smW = smem;                                       // [C][cout rows][128 B]
  uint8_t* smA = smem + ((p.C * wkb + 1023) & ~1023);        // [stages][C][128 rows][128 B]
  const int stage_bytes = p.C * C1_KB_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smA + (size_t)p.stages * stage_bytes);
  uint64_t* full_bar = bars;             // [stages] producers -> MMA
  uint64_t* empty_bar = bars + 4;        // [stages] MMA -> producers
  uint64_t* tfull_bar = bars + 8;        // [2]
  uint64_t* tempty_bar = bars + 10;      // [2]
  uint64_t* w_bar = bars + 12;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 13);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < p.stages; ++s) { mbar_init(&full_bar[s], C1_PRODUCERS); mbar_init(&empty_bar[s], 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(&tfull_bar[s], 1); mbar_init(&tempty_bar[s], 4); }
    mbar_init(w_bar, 1);
    fence_mbar_init();
  }
  if (warp == C1_MMA_WARP) { tmem_alloc(tmem_slot, 64); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_trigger();  // prologue done: the next kernel of the stream may start its own
  pdl_wait();     // the previous kernel's results are complete and visible from here on

  if (warp < C1_EPI_WARP0) {
    // ===== patch producers =====
    uint32_t stage = 0, phase = 0;
    for (int t = blockIdx.x; t < p.tiles; t += gridDim.x) {
      mbar_wait(&empty_bar[stage], phase ^ 1);
      if (p.xu8) c1_gather_tile_u8(p, smA + (size_t)stage * stage_bytes, (long long)t * 128);
      else c1_gather_tile(p, smA + (size_t)stage * stage_bytes, (long long)t * 128);
      fence_proxy_async();  // generic-proxy stores -> visible to the tensor core (async proxy)
      mbar_arrive(&full_bar[stage]);
      if (++stage == (uint32_t)p.stages) { stage = 0; phase ^= 1; }
    }
  } else if (warp == C1_MMA_WARP) {
    if (lane == 0) {
      tma_prefetch_desc(&tmW);
      mbar_arrive_expect_tx(w_bar, (uint32_t)(p.C * wkb));
      for (int c = 0; c < p.C; ++c) tma_load_2d(smW + c * wkb, &tmW, w_bar, c * 64, 0);
      mbar_wait(w_bar, 0);
      const uint32_t idesc = make_idesc_bf16(128, p.cout, 0, 0);
      uint32_t stage = 0, phase = 0, acc = 0, acc_phase = 0;
      for (int t = blockIdx.x; t < p.tiles; t += gridDim.x) {
        mbar_wait(&tempty_bar[acc], acc_phase ^ 1);
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        const uint32_t a0 = smem_u32(smA + (size_t)stage * stage_bytes), w0 = smem_u32(smW);
        for (int c = 0; c < p.C; ++c)
#pragma unroll
          for (int kk = 0; kk < 4; ++kk)
            umma_bf16(tmem_base + acc * 32, make_smem_desc_sw128(a0 + c * C1_KB_BYTES + kk * 32, 16, 1024),
                      make_smem_desc_sw128(w0 + c * wkb + kk * 32, 16, 1024), idesc, (c > 0 || kk > 0) ? 1u : 0u);
        umma_commit(&empty_bar[stage]);
        umma_commit(&tfull_bar[acc]);
        if (++stage == (uint32_t)p.stages) { stage = 0; phase ^= 1; }
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
  } else {
    // ===== epilogue: TMEM -> bias + ReLU -> bf16 NHWC rows =====
    const int q = warp & 3;
    uint32_t acc = 0, acc_phase = 0;
    for (int t = blockIdx.x; t < p.tiles; t += gridDim.x) {
      const long long row = (long long)t * 128 + q * 32 + lane;
      mbar_wait(&tfull_bar[acc], acc_phase);
      tc_fence_after();
      uint32_t v[32];
      tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + acc * 32, v);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty_bar[acc]);
      if (row < p.rows) {
        bf16* dst = p.y + row * p.cout;
        for (int j = 0; j < p.cout; j += 8) {
          float f[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) f[e] = fmaxf(__uint_as_float(v[j + e]) + __ldg(p.bias + j + e), 0.f);
          *reinterpret_cast<uint4*>(dst + j) =
              make_uint4(pack_bf16x2(f[0], f[1]), pack_bf16x2(f[2], f[3]), pack_bf16x2(f[4], f[5]), pack_bf16x2(f[6], f[7]));
        }
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == C1_MMA_WARP) { tc_fence_after(); tmem_dealloc(tmem_base, 64); }
}

// ---- weight gradient: dW[cout, 64C] = sum over patch rows of dZ[row, cout]^T P[row, :] --------------------------------------------
// A = dZ^T (MN-major: smem lines = patch rows, 64 m-elements per 128-byte line, only the first `cout` are real; the second 64-wide
// M chunk of the 128-row MMA is a shared all-zero region), B = P (MN-major: the producers' tile image), D[c] = TMEM cols c*64..
__global__ void __launch_bounds__(C1_THREADS, 1) conv1_wgrad_kernel(const __grid_constant__ CUtensorMap tmZ, const Conv1Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int stage_bytes = p.C * C1_KB_BYTES;
  uint8_t* smP = smem;                                            // [stages][C][128 lines][128 B]
  uint8_t* smZ = smP + (size_t)p.stages * stage_bytes;            // [stages][128 lines][128 B]
  uint8_t* smZero = smZ + (size_t)p.stages * C1_KB_BYTES;         // [128 lines][128 B] zeros
  uint8_t* smOnes = smZero + C1_KB_BYTES;                         // [128 lines][128 B]: element (line, n = 0) = 1 -> D[C][:, 0] = db
  uint64_t* bars = reinterpret_cast<uint64_t*>(smOnes + C1_KB_BYTES);
  uint64_t* full_bar = bars;         // [stages] 128 producer arrivals + 1 TMA (expect_tx) arrival
  uint64_t* empty_bar = bars + 4;    // [stages]
  uint64_t* done_bar = bars + 8;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 9);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < 2 * C1_KB_BYTES / 16; i += blockDim.x) reinterpret_cast<uint4*>(smZero)[i] = make_uint4(0, 0, 0, 0);
  __syncthreads();
  if (threadIdx.x < 128) {  // n = 0 is the first bf16 of 16-byte unit 0, which the 128B swizzle places at unit (0 ^ (line & 7))
    const int r = threadIdx.x;
    *reinterpret_cast<uint16_t*>(smOnes + r * 128 + ((r & 7) << 4)) = 0x3F80;  // bf16 1.0
  }
  if (threadIdx.x == 0) {
    for (int s = 0; s < p.stages; ++s) { mbar_init(&full_bar[s], C1_PRODUCERS + 1); mbar_init(&empty_bar[s], 1); }
    mbar_init(done_bar, 1);
    fence_mbar_init();
  }
  const uint32_t tmem_cols = (p.C + 1) * 64 <= 256 ? 256u : 512u;
  if (warp == C1_MMA_WARP) { tmem_alloc(tmem_slot, tmem_cols); tmem_relinquish(); }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_trigger();  // prologue done: the next kernel of the stream may start its own
  pdl_wait();     // the previous kernel's results are complete and visible from here on
  const int my_tiles = p.tiles > (int)blockIdx.x ? (p.tiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;

  if (warp < C1_EPI_WARP0) {
    uint32_t stage = 0, phase = 0;
    for (int t = blockIdx.x; t < p.tiles; t += gridDim.x) {
      mbar_wait(&empty_bar[stage], phase ^ 1);
      if (threadIdx.x == 0) {  // dZ^T tile: box [64 m][128 k] of the (cout, rows) tensor; columns >= cout and rows >= rows arrive as zeros
        mbar_arrive_expect_tx(&full_bar[stage], C1_KB_BYTES);
        tma_load_2d(smZ + (size_t)stage * C1_KB_BYTES, &tmZ, &full_bar[stage], 0, t * 128);
      }
      if (p.xu8) c1_gather_tile_u8(p, smP + (size_t)stage * stage_bytes, (long long)t * 128);
      else c1_gather_tile(p, smP + (size_t)stage * stage_bytes, (long long)t * 128);
      fence_proxy_async();
      mbar_arrive(&full_bar[stage]);
      if (++stage == (uint32_t)p.stages) { stage = 0; phase ^= 1; }
    }
  } else if (warp == C1_MMA_WARP) {
    if (lane == 0) {
      tma_prefetch_desc(&tmZ);
      const uint32_t idesc = make_idesc_bf16(128, 64, 1, 1);
      uint32_t stage = 0, phase = 0;
      const uint32_t zero_addr = smem_u32(smZero);
      for (int i = 0; i < my_tiles; ++i) {
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        const uint32_t z0 = smem_u32(smZ + (size_t)stage * C1_KB_BYTES), p0 = smem_u32(smP + (size_t)stage * stage_bytes);
        for (int c = 0; c < p.C; ++c)
#pragma unroll
          for (int ks = 0; ks < 8; ++ks)  // 16 patch rows per MMA
            umma_bf16(tmem_base + c * 64, make_smem_desc_sw128(z0 + ks * 2048, zero_addr - z0, 1024),
                      make_smem_desc_sw128(p0 + c * C1_KB_BYTES + ks * 2048, 16, 1024), idesc, (i > 0 || ks > 0) ? 1u : 0u);
        // bias gradient: the same dZ^T tile against the constant ones chunk (column 0 of accumulator block C = sum over the patch rows)
#pragma unroll
        for (int ks = 0; ks < 8; ++ks)
          umma_bf16(tmem_base + p.C * 64, make_smem_desc_sw128(z0 + ks * 2048, zero_addr - z0, 1024),
                    make_smem_desc_sw128(smem_u32(smOnes) + ks * 2048, 16, 1024), idesc, (i > 0 || ks > 0) ? 1u : 0u);
        umma_commit(&empty_bar[stage]);
        if (++stage == (uint32_t)p.stages) { stage = 0; phase ^= 1; }
      }
      umma_commit(done_bar);
    }
  } else if (warp == C1_EPI_WARP0) {
    // quadrant 0 holds output rows 0..31 = cout channels
    mbar_wait(done_bar, 0);
    tc_fence_after();
    const int K = 64 * (p.C + 1);  // + the ones block (only its column 0 is meaningful)
    float* dst = p.partial + ((size_t)blockIdx.x * p.cout + lane) * K;
    for (int c = 0; c <= p.C; ++c)
      for (int h = 0; h < 2; ++h) {
        uint32_t v[32];
        tmem_ld_32x32(tmem_base + c * 64 + h * 32, v);
        tmem_ld_wait();
        if (lane < p.cout) {
#pragma unroll
          for (int j = 0; j < 32; j += 4)
            *reinterpret_cast<float4*>(dst + c * 64 + h * 32 + j) =
                my_tiles > 0 ? make_float4(__uint_as_float(v[j]), __uint_as_float(v[j + 1]), __uint_as_float(v[j + 2]), __uint_as_float(v[j + 3]))
                             : make_float4(0.f, 0.f, 0.f, 0.f);
        }
      }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == C1_MMA_WARP) { tc_fence_after(); tmem_dealloc(tmem_base, tmem_cols); }
}

// per-CTA partials [ncta][cout][64 (C + 1)] -> dw [cout][64 C] (PyTorch order) and db [cout] (column 64 C).  32 outputs x 8 slice-lanes per
// block: lane l sums CTAs l, l + 8, ... (independent loads), the 8 lanes are combined in index order (deterministic).
__global__ void __launch_bounds__(256) conv1_wgrad_reduce_kernel(const float* __restrict__ partial, int ncta, int cout, int C,
                                                                 float* __restrict__ dw, float* __restrict__ db) {
  pdl_trigger();
  pdl_wait();
  __shared__ float red[8][33];
  const int Kp = 64 * (C + 1), K = 64 * C, total = cout * (K + 1);
  const int i = blockIdx.x * 32 + threadIdx.x;
  const int m = i / (K + 1), kk = i - m * (K + 1);
  float s = 0.f;
  if (i < total)
    for (int c = threadIdx.y; c < ncta; c += 8) s += __ldg(partial + ((size_t)c * cout + m) * Kp + kk);
  red[threadIdx.y][threadIdx.x] = s;
  __syncthreads();
  if (threadIdx.y == 0 && i < total) {
    for (int l = 1; l < 8; ++l) s += red[l][threadIdx.x];
    if (kk < K) dw[(size_t)m * K + kk] = s;
    else if (db) db[m] = s;
  }
}

// ---- host ------------------------------------------------------------------------------------------------------------------------
static int c1_params(Conv1Params& p, const void* xv, int x_u8, int N, int C, int H, int W, int stride, float scale, int cout) {
  DTB_REQUIRE(C >= 1 && C <= C1_MAXC, DTB_ERR_UNSUPPORTED, "conv1 implicit GEMM: in_channels %d unsupported (1..%d)", C, C1_MAXC);
  DTB_REQUIRE(W % 4 == 0 && stride % 4 == 0 && H >= 8 && W >= 8, DTB_ERR_UNSUPPORTED, "conv1: needs W %% 4 == 0, stride %% 4 == 0");
  DTB_REQUIRE(cout == 16 || cout == 32, DTB_ERR_UNSUPPORTED, "conv1: out_channels %d unsupported (16, 32)", cout);
  DTB_REQUIRE((reinterpret_cast<uintptr_t>(xv) & 15) == 0, DTB_ERR_BAD_ARG, "conv1: input must be 16-byte aligned");
  memset(&p, 0, sizeof(p));
  if (x_u8) p.xu8 = reinterpret_cast<const uint8_t*>(xv); else p.x = reinterpret_cast<const float*>(xv);
  p.N = N; p.C = C; p.H = H; p.W = W; p.stride = stride; p.cout = cout; p.scale = scale;
  p.ho = (H - 8) / stride + 1; p.wo = (W - 8) / stride + 1;
  p.rows = (long long)N * p.ho * p.wo;
  DTB_REQUIRE(p.rows < (1LL << 31), DTB_ERR_UNSUPPORTED, "conv1: %lld output pixels exceed the 32-bit row index", p.rows);
  p.tiles = (int)((p.rows + 127) / 128);
  return DTB_OK;
}
bool conv1_implicit_supported(int C, int W, int stride, int cout) {
  return C >= 1 && C <= C1_MAXC && W % 4 == 0 && stride % 4 == 0 && (cout == 16 || cout == 32);
}

int conv1_fwd(cudaStream_t stream, const void* x, int x_u8, int N, int C, int H, int W, int stride, float scale, const bf16* w_staged,
              const float* bias, int cout, bf16* y) {
  Conv1Params p;
  DTB_TRY(c1_params(p, x, x_u8, N, C, H, W, stride, scale, cout));
  if (p.rows == 0) return DTB_OK;
  p.bias = bias; p.y = y;
  p.stages = C <= 2 ? 4 : 3;
  if (p.stages * C * C1_KB_BYTES > 192 * 1024) p.stages = 192 * 1024 / (C * C1_KB_BYTES);
  CUtensorMap tmW;
  DTB_TRY(make_tmap(&tmW, w_staged, (uint64_t)64 * C, (uint64_t)cout, (uint64_t)64 * C, 64, (uint32_t)cout));
  const int smem = 1024 + ((C * cout * 128 + 1023) & ~1023) + p.stages * C * C1_KB_BYTES + 256 + 512;
  static bool attr = false;
  if (!attr) { DTB_CUDA(cudaFuncSetAttribute(conv1_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 225 * 1024)); attr = true; }
  const int grid = p.tiles < num_sms() ? p.tiles : num_sms();
  ProfScope prof(stream, PF_CONV1_FWD, (x_u8 ? 1.0 : 4.0) * N * C * H * W + 2.0 * (double)p.rows * cout);
  DTB_LAUNCH(conv1_fwd_kernel, grid, C1_THREADS, smem, stream, tmW, p);
  DTB_CHECK_LAUNCH();
  return DTB_OK;
}

size_t conv1_wgrad_partial_floats(int C, int cout) { return (size_t)DTB_NUM_SMS_FALLBACK * cout * 64 * (C + 1); }

int conv1_wgrad(cudaStream_t stream, const void* x, int x_u8, int N, int C, int H, int W, int stride, float scale, const bf16* dz, int cout,
                float* partial, float* dw, float* db) {
  Conv1Params p;
  DTB_TRY(c1_params(p, x, x_u8, N, C, H, W, stride, scale, cout));
  if (p.rows == 0) {
    DTB_CUDA(cudaMemsetAsync(dw, 0, (size_t)cout * 64 * C * sizeof(float), stream));
    if (db) DTB_CUDA(cudaMemsetAsync(db, 0, (size_t)cout * sizeof(float), stream));
    return DTB_OK;
  }
  p.partial = partial;
  p.stages = 2;
  CUtensorMap tmZ;
  DTB_TRY(make_tmap(&tmZ, dz, (uint64_t)cout, (uint64_t)p.rows, (uint64_t)cout, 64, 128));
  const int smem = 1024 + p.stages * (C + 1) * C1_KB_BYTES + 2 * C1_KB_BYTES + 256 + 512;
  static bool attr = false;
  if (!attr) { DTB_CUDA(cudaFuncSetAttribute(conv1_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 225 * 1024)); attr = true; }
  int grid = p.tiles < num_sms() ? p.tiles : num_sms();
  if (grid > DTB_NUM_SMS_FALLBACK) grid = DTB_NUM_SMS_FALLBACK;
  ProfScope prof(stream, PF_CONV1_WGRAD, (x_u8 ? 1.0 : 4.0) * N * C * H * W + 2.0 * (double)p.rows * cout);
  DTB_LAUNCH(conv1_wgrad_kernel, grid, C1_THREADS, smem, stream, tmZ, p);
  DTB_CHECK_LAUNCH();
  prof.~ProfScope();
  prof.slot = -1;
  const int total = cout * (64 * C + 1);
  DTB_LAUNCH(conv1_wgrad_reduce_kernel, (total + 31) / 32, dim3(32, 8), 0, stream, partial, grid, cout, C, dw, db);
  DTB_CHECK_LAUNCH();
  return DTB_OK;
}

}  // namespace dtb
