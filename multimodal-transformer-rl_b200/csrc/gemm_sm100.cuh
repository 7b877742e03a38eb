// tcgen05 / TMEM / TMA GEMM for sm_100a — the tensor-core core of the Decision-Transformer path.
//
//   D[M,N] = epilogue( sum_k A(m,k) * B(n,k) )      bf16 operands, fp32 accumulate in TMEM
//
// Operand majors are runtime switches (they only change TMA boxes + UMMA descriptors):
//   K-major  A: memory [M rows][K cols] (K contiguous)          -> forward / dgrad activations
//   MN-major A: memory [K rows][M cols] (M contiguous)          -> wgrad (A = X^T read straight from X)
//   K-major  B: memory [N rows][K cols]                         -> W^T copy (fwd) or W itself (dgrad)
//   MN-major B: memory [K rows][N cols]                         -> wgrad (B = dY read straight from dY)
// Replaces the cuBLAS calls behind nn.Linear / Conv1D in the reference stack (SURVEY.md §2.3 K3,K5-K7,
// HF modeling_decision_transformer.py:179,217,234,236) and their autograd backward.
#pragma once
#include "common.cuh"
#include "kernels.h"

namespace dtb {

constexpr int GEMM_BM = 128;       // UMMA M (cta_group::1)
constexpr int GEMM_BK = 64;        // 64 bf16 = one 128-byte swizzle row
constexpr int GEMM_STAGES = 4;
constexpr int GEMM_MAX_BN = 256;
constexpr int GEMM_THREADS = 256;  // warp0 TMA, warp1 MMA, warp2 TMEM alloc, warp3 idle, warps4-7 epilogue
constexpr int GEMM_A_STAGE_BYTES = GEMM_BM * GEMM_BK * 2;      // 16 KB
constexpr int GEMM_B_STAGE_BYTES = GEMM_MAX_BN * GEMM_BK * 2;  // 32 KB
constexpr int GEMM_SMEM_BYTES = GEMM_STAGES * (GEMM_A_STAGE_BYTES + GEMM_B_STAGE_BYTES) + 1024 /*align*/ + 256 /*barriers*/;

__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const GemmParams p) {
  extern __shared__ uint8_t smem_raw[];
  // 1024-byte alignment is required by the 128B swizzle atoms.
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smA = smem;
  uint8_t* smB = smem + GEMM_STAGES * GEMM_A_STAGE_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + GEMM_STAGES * (GEMM_A_STAGE_BYTES + GEMM_B_STAGE_BYTES));
  uint64_t* full_bar = bars;                     // [STAGES]  TMA -> MMA
  uint64_t* empty_bar = bars + GEMM_STAGES;      // [STAGES]  MMA -> TMA
  uint64_t* tfull_bar = bars + 2 * GEMM_STAGES;  // [2]       MMA -> epilogue
  uint64_t* tempty_bar = tfull_bar + 2;          // [2]       epilogue -> MMA
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < GEMM_STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tfull_bar[s], 1);
      mbar_init(&tempty_bar[s], 4);  // one arrive per epilogue warp
    }
    fence_mbar_init();
  }
  if (warp == 2) {
    tmem_alloc(tmem_slot, 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int total_tiles = p.m_tiles * p.n_tiles * p.split_k;
  const int kblocks_total = (p.K + GEMM_BK - 1) / GEMM_BK;
  const int BN = p.BN;
  const uint32_t a_bytes = GEMM_A_STAGE_BYTES;
  const uint32_t b_bytes = (uint32_t)BN * GEMM_BK * 2;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      uint32_t stage = 0, phase = 0;
      for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
        const int split = t / (p.m_tiles * p.n_tiles);
        const int mn = t - split * (p.m_tiles * p.n_tiles);
        const int m0 = (mn / p.n_tiles) * GEMM_BM;
        const int n0 = (mn % p.n_tiles) * BN;
        const int kb0 = split * p.kb_per_split;
        const int kb1 = min(kb0 + p.kb_per_split, kblocks_total);
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          mbar_arrive_expect_tx(&full_bar[stage], a_bytes + b_bytes);
          uint8_t* a_dst = smA + stage * GEMM_A_STAGE_BYTES;
          uint8_t* b_dst = smB + stage * GEMM_B_STAGE_BYTES;
          const int k0 = kb * GEMM_BK;
          if (!p.a_mn_major) {
            tma_load_2d(a_dst, &tmA, &full_bar[stage], k0, m0);  // box [64 k][128 m]
          } else {
            tma_load_2d(a_dst, &tmA, &full_bar[stage], m0, k0);  // box [64 m][64 k] x2
            tma_load_2d(a_dst + GEMM_BK * 128, &tmA, &full_bar[stage], m0 + 64, k0);
          }
          if (!p.b_mn_major) {
            tma_load_2d(b_dst, &tmB, &full_bar[stage], k0, n0);  // box [64 k][BN n]
          } else {
            for (int j = 0; j < BN / 64; ++j)
              tma_load_2d(b_dst + j * (GEMM_BK * 128), &tmB, &full_bar[stage], n0 + 64 * j, k0);
          }
          if (++stage == GEMM_STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (single thread) =====================
    if (lane == 0) {
      const uint32_t idesc = make_idesc_bf16(GEMM_BM, BN, p.a_mn_major, p.b_mn_major);
      // descriptor constants per major
      const uint32_t a_lbo = p.a_mn_major ? GEMM_BK * 128 : 16;
      const uint32_t b_lbo = p.b_mn_major ? GEMM_BK * 128 : 16;
      const uint32_t a_kstep = p.a_mn_major ? 16 * 128 : 32;  // bytes per UMMA_K=16 step
      const uint32_t b_kstep = p.b_mn_major ? 16 * 128 : 32;
      uint32_t stage = 0, phase = 0;
      uint32_t acc = 0, acc_phase = 0;
      for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
        const int split = t / (p.m_tiles * p.n_tiles);
        const int kb0 = split * p.kb_per_split;
        const int kb1 = min(kb0 + p.kb_per_split, kblocks_total);
        mbar_wait(&tempty_bar[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * GEMM_MAX_BN;
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t a_addr = smem_u32(smA + stage * GEMM_A_STAGE_BYTES);
          const uint32_t b_addr = smem_u32(smB + stage * GEMM_B_STAGE_BYTES);
#pragma unroll
          for (int kk = 0; kk < GEMM_BK / 16; ++kk) {
            const uint64_t adesc = make_smem_desc_sw128(a_addr + kk * a_kstep, a_lbo, 1024);
            const uint64_t bdesc = make_smem_desc_sw128(b_addr + kk * b_kstep, b_lbo, 1024);
            umma_bf16(d_tmem, adesc, bdesc, idesc, (kb > kb0 || kk > 0) ? 1u : 0u);
          }
          umma_commit(&empty_bar[stage]);  // frees the smem slot when these MMAs retire
          if (++stage == GEMM_STAGES) { stage = 0; phase ^= 1; }
        }
        umma_commit(&tfull_bar[acc]);  // accumulator ready for the epilogue
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
  } else if (warp >= 4) {
    // ===================== epilogue: TMEM -> registers -> global =====================
    const int q = warp & 3;  // TMEM lane quadrant accessible to this warp
    const GemmEpilogue& ep = p.ep;
    uint32_t acc = 0, acc_phase = 0;
    for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
      const int split = t / (p.m_tiles * p.n_tiles);
      const int mn = t - split * (p.m_tiles * p.n_tiles);
      const int m0 = (mn / p.n_tiles) * GEMM_BM;
      const int n0 = (mn % p.n_tiles) * BN;
      const int row = m0 + q * 32 + lane;
      const bool row_ok = row < p.M;
      mbar_wait(&tfull_bar[acc], acc_phase);
      tc_fence_after();
      for (int c = 0; c < BN; c += 32) {
        uint32_t v[32];
        tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + acc * GEMM_MAX_BN + c, v);
        tmem_ld_wait();
        const int col0 = n0 + c;
        if (row_ok && col0 < p.N) {
          const int ncols = min(32, p.N - col0);  // N % 8 == 0 guaranteed by host
          if (ep.partial) {
            float* dst = ep.partial + ((size_t)split * p.M + row) * p.N + col0;
#pragma unroll
            for (int j = 0; j < 32; j += 4)
              if (j < ncols)
                *reinterpret_cast<float4*>(dst + j) = make_float4(__uint_as_float(v[j]), __uint_as_float(v[j + 1]),
                                                                  __uint_as_float(v[j + 2]), __uint_as_float(v[j + 3]));
          } else {
            float f[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]) * ep.alpha;
            if (ep.bias) {
#pragma unroll
              for (int j = 0; j < 32; j += 4)
                if (j < ncols) {
                  float4 b = ldg_f4(ep.bias + col0 + j);
                  f[j] += b.x; f[j + 1] += b.y; f[j + 2] += b.z; f[j + 3] += b.w;
                }
            }
            if (ep.out_preact) {
              bf16* dst = ep.out_preact + (size_t)row * ep.ldo_b + col0;
#pragma unroll
              for (int j = 0; j < 32; j += 8)
                if (j < ncols)
                  *reinterpret_cast<uint4*>(dst + j) = make_uint4(pack_bf16x2(f[j], f[j + 1]), pack_bf16x2(f[j + 2], f[j + 3]),
                                                                  pack_bf16x2(f[j + 4], f[j + 5]), pack_bf16x2(f[j + 6], f[j + 7]));
            }
            if (ep.act == 1) {
#pragma unroll
              for (int j = 0; j < 32; ++j) f[j] = fmaxf(f[j], 0.f);
            } else if (ep.act == 2) {
#pragma unroll
              for (int j = 0; j < 32; ++j) f[j] = gelu_new_f(f[j]);
            }
            if (ep.dact) {
              const bf16* a = ep.aux + (size_t)row * ep.ldaux + col0;
#pragma unroll
              for (int j = 0; j < 32; j += 8)
                if (j < ncols) {
                  uint4 u = ldg_u4(a + j);
                  uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
                  for (int e = 0; e < 4; ++e) {
                    float2 x = unpack_bf16x2(w[e]);
                    if (ep.dact == 1) {
                      f[j + 2 * e] = x.x > 0.f ? f[j + 2 * e] : 0.f;
                      f[j + 2 * e + 1] = x.y > 0.f ? f[j + 2 * e + 1] : 0.f;
                    } else {
                      f[j + 2 * e] *= gelu_new_grad_f(x.x);
                      f[j + 2 * e + 1] *= gelu_new_grad_f(x.y);
                    }
                  }
                }
            }
            if (ep.residual) {
              const float* r = ep.residual + (size_t)row * ep.ldr + col0;
#pragma unroll
              for (int j = 0; j < 32; j += 4)
                if (j < ncols) {
                  float4 b = ldg_f4(r + j);
                  f[j] += b.x; f[j + 1] += b.y; f[j + 2] += b.z; f[j + 3] += b.w;
                }
            }
            if (ep.out_f32) {
              float* dst = ep.out_f32 + (size_t)row * ep.ldo_f + col0;
#pragma unroll
              for (int j = 0; j < 32; j += 4)
                if (j < ncols) *reinterpret_cast<float4*>(dst + j) = make_float4(f[j], f[j + 1], f[j + 2], f[j + 3]);
            }
            if (ep.out_bf16) {
              bf16* dst = ep.out_bf16 + (size_t)row * ep.ldo_b + col0;
#pragma unroll
              for (int j = 0; j < 32; j += 8)
                if (j < ncols)
                  *reinterpret_cast<uint4*>(dst + j) = make_uint4(pack_bf16x2(f[j], f[j + 1]), pack_bf16x2(f[j + 2], f[j + 3]),
                                                                  pack_bf16x2(f[j + 4], f[j + 5]), pack_bf16x2(f[j + 6], f[j + 7]));
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty_bar[acc]);
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace dtb
