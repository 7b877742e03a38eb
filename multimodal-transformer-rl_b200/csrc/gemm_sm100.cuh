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
constexpr int GEMM_EPI_STAGE_BYTES = 4 * 32 * 36 * 4;  // 4 epilogue warps x [32 rows x 36 words] transpose staging (EPI_WARP_WORDS)
constexpr int GEMM_SMEM_BYTES = GEMM_STAGES * (GEMM_A_STAGE_BYTES + GEMM_B_STAGE_BYTES) + 1024 /*align*/ + 256 /*barriers*/ + GEMM_EPI_STAGE_BYTES;

// One 32-column chunk of one accumulator row: bias / activation / act' / residual / stores (or the split-K partial store).
__device__ __forceinline__ void gemm_epilogue_chunk(const GemmParams& p, const GemmEpilogue& ep, const uint32_t (&v)[32], int row, bool row_ok,
                                                    int col0, int split) {
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
            if (ep.drop.state) {  // resid_dropout(c_proj(.)) before the residual add (HF modeling_gpt2 :217-218, :236-237)
              const uint32_t key = drop_site_key(ep.drop) + (uint32_t)row * (uint32_t)p.N + (uint32_t)col0;
#pragma unroll
              for (int j = 0; j < 32; ++j) f[j] = drop_apply(f[j], key, (uint32_t)j, ep.drop.thr, ep.drop.scale);
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

// ---- staged epilogue (both kernels): every global access of the epilogue is warp-coalesced --------------------------------------
// A warp owns 32 accumulator rows (TMEM lanes) x 32 columns per chunk, one row per thread.  Row-per-thread global accesses touch 32
// different 128-byte lines per instruction, which made the epilogue (not the tensor pipe) the bound of the forward / dgrad GEMMs.
// Here every tile that enters or leaves the thread-owns-a-row layout is transposed through a per-warp shared-memory buffer
// (32 rows x 36 words: 16-byte aligned rows, conflict-free for 128-bit accesses in both directions), so a load / store instruction
// covers 4 whole fp32 lines (or 8 half lines of bf16).
constexpr int EPI_LD = 36;                        // staging row stride in 32-bit words
constexpr int EPI_WARP_WORDS = 32 * EPI_LD;       // 4608 B per warp
constexpr int EPI_LDH = 20;                       // bf16 staging: 32 columns = 16 words (+4 pad) per row

__device__ __forceinline__ void sts128(uint32_t* p, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  *reinterpret_cast<uint4*>(p) = make_uint4(a, b, c, d);
}
// thread-owns-row fp32 [32] -> staging -> coalesced fp32 global rows
__device__ __forceinline__ void epi_write_f32(uint32_t* st, int lane, const float (&f)[32], float* dst, int ld, int row0, int M, int col0,
                                              int ncols) {
#pragma unroll
  for (int j = 0; j < 32; j += 4)
    sts128(st + lane * EPI_LD + j, __float_as_uint(f[j]), __float_as_uint(f[j + 1]), __float_as_uint(f[j + 2]), __float_as_uint(f[j + 3]));
  __syncwarp();
  const int cc = 4 * (lane & 7);
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int rr = i * 4 + (lane >> 3);
    const uint4 v = *reinterpret_cast<const uint4*>(st + rr * EPI_LD + cc);
    if (row0 + rr < M && cc < ncols) *reinterpret_cast<uint4*>(dst + (size_t)(row0 + rr) * ld + col0 + cc) = v;
  }
  __syncwarp();
}
// thread-owns-row fp32 [32] -> bf16 -> staging -> coalesced bf16 global rows
__device__ __forceinline__ void epi_write_bf16(uint32_t* st, int lane, const float (&f)[32], bf16* dst, int ld, int row0, int M, int col0,
                                               int ncols) {
#pragma unroll
  for (int j = 0; j < 32; j += 8)
    sts128(st + lane * EPI_LDH + j / 2, pack_bf16x2(f[j], f[j + 1]), pack_bf16x2(f[j + 2], f[j + 3]), pack_bf16x2(f[j + 4], f[j + 5]),
           pack_bf16x2(f[j + 6], f[j + 7]));
  __syncwarp();
  const int piece = lane & 3;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int rr = i * 8 + (lane >> 2);
    const uint4 v = *reinterpret_cast<const uint4*>(st + rr * EPI_LDH + piece * 4);
    if (row0 + rr < M && piece * 8 < ncols) *reinterpret_cast<uint4*>(dst + (size_t)(row0 + rr) * ld + col0 + piece * 8) = v;
  }
  __syncwarp();
}
// coalesced global loads of a [32 rows x 32 cols] block, issued early (registers), landed in the thread-owns-row layout later
struct EpiLoadF32 { uint4 v[8]; };
struct EpiLoadBf16 { uint4 v[4]; };
__device__ __forceinline__ void epi_issue_f32(EpiLoadF32& L, int lane, const float* src, int ld, int row0, int M, int col0, int ncols) {
  const int cc = 4 * (lane & 7);
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int rr = i * 4 + (lane >> 3);
    L.v[i] = (row0 + rr < M && cc < ncols) ? ldg_u4(src + (size_t)(row0 + rr) * ld + col0 + cc) : make_uint4(0, 0, 0, 0);
  }
}
__device__ __forceinline__ void epi_land_f32(const EpiLoadF32& L, uint32_t* st, int lane, float (&r)[32]) {
  const int cc = 4 * (lane & 7);
#pragma unroll
  for (int i = 0; i < 8; ++i) *reinterpret_cast<uint4*>(st + (i * 4 + (lane >> 3)) * EPI_LD + cc) = L.v[i];
  __syncwarp();
#pragma unroll
  for (int j = 0; j < 32; j += 4) {
    const uint4 v = *reinterpret_cast<const uint4*>(st + lane * EPI_LD + j);
    r[j] = __uint_as_float(v.x); r[j + 1] = __uint_as_float(v.y); r[j + 2] = __uint_as_float(v.z); r[j + 3] = __uint_as_float(v.w);
  }
  __syncwarp();
}
__device__ __forceinline__ void epi_issue_bf16(EpiLoadBf16& L, int lane, const bf16* src, int ld, int row0, int M, int col0, int ncols) {
  const int piece = lane & 3;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int rr = i * 8 + (lane >> 2);
    L.v[i] = (row0 + rr < M && piece * 8 < ncols) ? ldg_u4(src + (size_t)(row0 + rr) * ld + col0 + piece * 8) : make_uint4(0, 0, 0, 0);
  }
}
__device__ __forceinline__ void epi_land_bf16(const EpiLoadBf16& L, uint32_t* st, int lane, float (&r)[32]) {
  const int piece = lane & 3;
#pragma unroll
  for (int i = 0; i < 4; ++i) *reinterpret_cast<uint4*>(st + (i * 8 + (lane >> 2)) * EPI_LDH + piece * 4) = L.v[i];
  __syncwarp();
#pragma unroll
  for (int j = 0; j < 32; j += 8) {
    const uint4 v = *reinterpret_cast<const uint4*>(st + lane * EPI_LDH + j / 2);
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float2 x = unpack_bf16x2(w[e]);
      r[j + 2 * e] = x.x; r[j + 2 * e + 1] = x.y;
    }
  }
  __syncwarp();
}


// One 32-row x 32-column chunk of a warp, staged: `v` = this thread's accumulator row, row0 = first row of the warp's 32 rows.
// `bias_lane` = bias[col0 + lane] (loaded early by the caller), `Lr` / `La` = the residual / aux blocks issued early.
__device__ __forceinline__ void gemm_epilogue_chunk_staged(const GemmParams& p, const GemmEpilogue& ep, const uint32_t (&v)[32], uint32_t* st,
                                                           int lane, int row0, int col0, int split, float bias_lane, const EpiLoadF32& Lr,
                                                           const EpiLoadBf16& La, float (&f)[32]) {
  if (col0 >= p.N) return;  // warp-uniform
  const int ncols = min(32, p.N - col0);
  if (ep.partial) {
#pragma unroll
    for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]);
    epi_write_f32(st, lane, f, ep.partial + (size_t)split * p.M * p.N, p.N, row0, p.M, col0, ncols);
    return;
  }
#pragma unroll
  for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]) * ep.alpha;
  if (ep.bias) {
#pragma unroll
    for (int j = 0; j < 32; ++j) f[j] += __shfl_sync(0xffffffffu, bias_lane, j);
  }
  if (ep.out_preact) epi_write_bf16(st, lane, f, ep.out_preact, ep.ldo_b, row0, p.M, col0, ncols);
  if (ep.act == 1) {
#pragma unroll
    for (int j = 0; j < 32; ++j) f[j] = fmaxf(f[j], 0.f);
  } else if (ep.act == 2) {
#pragma unroll
    for (int j = 0; j < 32; ++j) f[j] = gelu_new_f(f[j]);
  }
  if (ep.dact) {
    float a[32];
    epi_land_bf16(La, st, lane, a);
    if (ep.dact == 1) {
#pragma unroll
      for (int j = 0; j < 32; ++j) f[j] = a[j] > 0.f ? f[j] : 0.f;
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j) f[j] *= gelu_new_grad_f(a[j]);
    }
  }
  if (ep.drop.state) {  // same bits as gemm_epilogue_chunk: idx = row * N + col
    const uint32_t key = drop_site_key(ep.drop) + (uint32_t)(row0 + lane) * (uint32_t)p.N + (uint32_t)col0;
#pragma unroll
    for (int j = 0; j < 32; ++j) f[j] = drop_apply(f[j], key, (uint32_t)j, ep.drop.thr, ep.drop.scale);
  }
  if (ep.residual) {
    float r[32];
    epi_land_f32(Lr, st, lane, r);
#pragma unroll
    for (int j = 0; j < 32; ++j) f[j] += r[j];
  }
  if (ep.out_f32) epi_write_f32(st, lane, f, ep.out_f32, ep.ldo_f, row0, p.M, col0, ncols);
  if (ep.out_bf16) epi_write_bf16(st, lane, f, ep.out_bf16, ep.ldo_b, row0, p.M, col0, ncols);
}

// Split-K, last arrival reduces: called by one epilogue warp after it stored its partial rows [row0, row0 + 32) x the column chunks
// c = c_begin, c_begin + c_step, ... of the tile at n0.  The warp that takes the last of the split_k tickets of its counter sums the
// slices in index order (the same order for whichever CTA happens to be last: deterministic) and writes sk_out; it leaves the counter
// at zero for the next launch.  Slices are read with ld.global.cg (L2): the other CTAs' stores were fenced before their tickets.
__device__ __forceinline__ void gemm_splitk_finish(const GemmParams& p, const GemmEpilogue& ep, unsigned int* counter, int lane, int row0,
                                                   int n0, int c_begin, int c_step, int c_end) {
  __threadfence();
  __syncwarp();
  unsigned int ticket = 0;
  if (lane == 0) ticket = atomicAdd(counter, 1u);
  ticket = __shfl_sync(0xffffffffu, ticket, 0);
  if (ticket != (unsigned int)p.split_k - 1u) return;  // warp-uniform
  __threadfence();
  const size_t slice = (size_t)p.M * p.N;
  const int cc = 4 * (lane & 7);
  for (int c = c_begin; c < c_end; c += c_step) {
    const int col0 = n0 + c;
    if (col0 >= p.N) break;
    const int ncols = min(32, p.N - col0);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int row = row0 + i * 4 + (lane >> 3);
      if (row < p.M && cc < ncols) {
        const float* src = ep.partial + (size_t)row * p.N + col0 + cc;
        float4 s = __ldcg(reinterpret_cast<const float4*>(src));
        for (int k = 1; k < p.split_k; ++k) {
          const float4 t = __ldcg(reinterpret_cast<const float4*>(src + (size_t)k * slice));
          s.x += t.x; s.y += t.y; s.z += t.z; s.w += t.w;
        }
        *reinterpret_cast<float4*>(ep.sk_out + (size_t)row * ep.sk_ldo + col0 + cc) = s;
      }
    }
  }
  if (lane == 0) *counter = 0u;
}

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
  pdl_trigger();  // prologue done: the next kernel of the stream may start its own
  pdl_wait();     // the previous kernel's results are complete and visible from here on

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
    uint32_t* st = reinterpret_cast<uint32_t*>(smem + GEMM_STAGES * (GEMM_A_STAGE_BYTES + GEMM_B_STAGE_BYTES) + 256) + q * EPI_WARP_WORDS;
    uint32_t acc = 0, acc_phase = 0;
    for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
      const int split = t / (p.m_tiles * p.n_tiles);
      const int mn = t - split * (p.m_tiles * p.n_tiles);
      const int m0 = (mn / p.n_tiles) * GEMM_BM;
      const int n0 = (mn % p.n_tiles) * BN;
      const int row0 = m0 + q * 32;
      // the first chunk's bias / residual / aux blocks are fetched (coalesced) before the accumulator is ready
      EpiLoadF32 Lr;
      EpiLoadBf16 La;
      float bias_lane = 0.f;
      auto prefetch = [&](int col0) {
        if (ep.partial || col0 >= p.N) return;
        const int ncols = min(32, p.N - col0);
        if (ep.bias) bias_lane = lane < ncols ? __ldg(ep.bias + col0 + lane) : 0.f;
        if (ep.residual) epi_issue_f32(Lr, lane, ep.residual, ep.ldr, row0, p.M, col0, ncols);
        if (ep.dact) epi_issue_bf16(La, lane, ep.aux, ep.ldaux, row0, p.M, col0, ncols);
      };
      prefetch(n0);
      mbar_wait(&tfull_bar[acc], acc_phase);
      tc_fence_after();
      const uint32_t t_acc = tmem_base + ((uint32_t)(q * 32) << 16) + acc * GEMM_MAX_BN;
      float rsum = 0.f;
      for (int c = 0; c < BN; c += 32) {
        uint32_t v[32];
        tmem_ld_32x32(t_acc + c, v);
        tmem_ld_wait();
        const EpiLoadF32 Lr_cur = Lr;
        const EpiLoadBf16 La_cur = La;
        const float bias_cur = bias_lane;
        if (c + 32 < BN) prefetch(n0 + c + 32);  // next chunk's operands fly while this one is processed
        float f[32];
        gemm_epilogue_chunk_staged(p, ep, v, st, lane, row0, n0 + c, split, bias_cur, Lr_cur, La_cur, f);
        if (ep.ln_out) {  // fused LayerNorm, pass 1: the finished fp32 row values go back into the accumulator columns; row sum
#pragma unroll
          for (int j = 0; j < 32; ++j) { rsum += f[j]; v[j] = __float_as_uint(f[j]); }
          tmem_st_32x32(t_acc + c, v);
        }
      }
      if (ep.ln_out) {  // host guarantees BN == N, N % 32 == 0, one n-tile, no split-K: this thread owns the whole row
        tmem_st_wait();
        const float inv_n = 1.f / (float)p.N;
        const float mean = rsum * inv_n;
        float q2 = 0.f;
        for (int c = 0; c < BN; c += 32) {  // pass 2: centred sum of squares (two-pass variance, as nn.LayerNorm computes it)
          uint32_t v[32];
          tmem_ld_32x32(t_acc + c, v);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 32; ++j) { const float d = __uint_as_float(v[j]) - mean; q2 = fmaf(d, d, q2); }
        }
        const float rstd = rsqrtf(q2 * inv_n + ep.ln_eps);
        for (int c = 0; c < BN; c += 32) {  // pass 3: normalise, scale / shift, coalesced bf16 rows
          uint32_t v[32];
          tmem_ld_32x32(t_acc + c, v);
          tmem_ld_wait();
          const float g_lane = __ldg(ep.ln_gamma + c + lane), b_lane = __ldg(ep.ln_beta + c + lane);
          float y[32];
#pragma unroll
          for (int j = 0; j < 32; ++j)
            y[j] = fmaf((__uint_as_float(v[j]) - mean) * rstd, __shfl_sync(0xffffffffu, g_lane, j), __shfl_sync(0xffffffffu, b_lane, j));
          epi_write_bf16(st, lane, y, ep.ln_out, ep.ld_ln, row0, p.M, c, 32);
        }
        if (row0 + lane < p.M) {
          if (ep.ln_mean) ep.ln_mean[row0 + lane] = mean;
          if (ep.ln_rstd) ep.ln_rstd[row0 + lane] = rstd;
        }
      }
      if (ep.partial && ep.sk_counters) gemm_splitk_finish(p, ep, ep.sk_counters + (size_t)mn * 4 + q, lane, row0, n0, 0, 32, BN);
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



// ================================================================================================================================
// CTA-pair variant (cta_group::2): two CTAs of a cluster (the two SMs of a TPC) compute one 256 x 256 tile.  Each CTA stages its
// own 128 rows of A and its own 128-column half of B (32 KB per k-block instead of 48 KB), the leader CTA's single thread issues
// tcgen05.mma.cta_group::2 (M = 256: rows 0-127 accumulate in the leader's TMEM, rows 128-255 in the peer's), and both tensor cores
// read both halves of B.  Halves the shared-memory fill per flop, which is what bounds the single-CTA kernel on large GEMMs.
//   full[s]   : leader only; expect_tx = both CTAs' bytes (the peer's TMA signals the leader's barrier)
//   empty[s]  : one per CTA, arrived by the leader's tcgen05.commit multicast to both CTAs
//   tfull[a]  : one per CTA, same multicast commit;   tempty[a] : leader only, 8 arrivals (4 epilogue warps x 2 CTAs, remote for the peer)
// ================================================================================================================================
constexpr int GEMM2_STAGES = 5;
constexpr int GEMM2_BN = 256;
constexpr int GEMM2_THREADS = 384;  // warp0 TMA, warp1 MMA, warp2 TMEM alloc, warp3 idle, warps 4-11 epilogue (2 per TMEM lane quadrant)
constexpr int GEMM2_EPI_WARPS = 8;
constexpr int GEMM2_A_BYTES = 128 * GEMM_BK * 2;  // 16 KB
constexpr int GEMM2_B_BYTES = 128 * GEMM_BK * 2;  // 16 KB: this CTA's half of the 256 columns
constexpr int GEMM2_SMEM_BYTES = GEMM2_STAGES * (GEMM2_A_BYTES + GEMM2_B_BYTES) + 1024 + 256 + GEMM2_EPI_WARPS * EPI_WARP_WORDS * 4;

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t mapa_rank(uint32_t saddr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(saddr), "r"(rank));
  return r;
}
// TMA load issued by either CTA of the pair; the transaction bytes are credited to the LEADER's mbarrier (peer bit cleared).
__device__ __forceinline__ void tma_load_2d_pair(void* smem_dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
          smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(mapa_rank(smem_u32(bar), 0)), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void umma_bf16_pair(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive (once all prior MMAs of this thread retire) on the barrier at this smem offset in BOTH CTAs of the pair
__device__ __forceinline__ void umma_commit_pair(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(smem_u32(bar)),
               "h"((uint16_t)3)
               : "memory");
}
// arrive on the barrier at this smem offset in CTA `rank` of the cluster.  Default (.release.cta) semantics on purpose: the only thing
// ordered through these barriers is TMEM / smem traffic already fenced by tcgen05.wait / tcgen05.fence; `.release.cluster` compiled to
// an ERRBAR that drained every outstanding global store of the epilogue warp before the accumulator could be handed back (ncu: 7 % of
// the kernel's stall samples).
__device__ __forceinline__ void mbar_arrive_cluster(uint64_t* bar, uint32_t rank) {
  asm volatile(
      "{\n\t.reg .b32 ra;\n\t"
      "mapa.shared::cluster.u32 ra, %0, %1;\n\t"
      "mbarrier.arrive.shared::cluster.b64 _, [ra];\n\t}" ::"r"(smem_u32(bar)),
      "r"(rank)
      : "memory");
}
__device__ __forceinline__ void tmem_alloc_pair(uint32_t* dst_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tmem_relinquish_pair() {
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_pair(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
// wait on a barrier that the peer CTA's threads / tensor core / TMA arrive on.  Default (.acquire.cta) scope on purpose: what these
// barriers order is TMEM and async-proxy smem traffic, never generic global memory; `.acquire.cluster` compiled to a CCTL.IVALL (whole-L1
// invalidate) after every successful wait, i.e. per k-block in the producer / MMA loops and per tile in the epilogue warps.
__device__ __forceinline__ void mbar_wait_cluster(uint64_t* bar, uint32_t parity) {
  long long t0 = 0;
  for (;;) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    if (ok) return;
    if (t0 == 0) t0 = clock64();
    else if (clock64() - t0 > 4000000000LL) {
      printf("dtb200: cluster mbarrier wait timeout (block %d thread %d)\n", blockIdx.x, threadIdx.x);
      __trap();
    }
  }
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(GEMM2_THREADS, 1)
gemm_tcgen05_pair_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const GemmParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smA = smem;
  uint8_t* smB = smem + GEMM2_STAGES * GEMM2_A_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + GEMM2_STAGES * (GEMM2_A_BYTES + GEMM2_B_BYTES));
  uint64_t* full_bar = bars;                      // [STAGES]
  uint64_t* empty_bar = bars + GEMM2_STAGES;      // [STAGES]
  uint64_t* tfull_bar = bars + 2 * GEMM2_STAGES;  // [2]
  uint64_t* tempty_bar = tfull_bar + 2;           // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);
  uint32_t* epi_stage = reinterpret_cast<uint32_t*>(smem + GEMM2_STAGES * (GEMM2_A_BYTES + GEMM2_B_BYTES) + 256);  // [8 warps][32 x 36]

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < GEMM2_STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tfull_bar[s], 1);
      mbar_init(&tempty_bar[s], 2 * GEMM2_EPI_WARPS);  // epilogue warps of both CTAs
    }
    fence_mbar_init();
  }
  if (warp == 2) {
    tmem_alloc_pair(tmem_slot, 512);
    tmem_relinquish_pair();
  }
  tc_fence_before();
  cluster_sync_all();  // both CTAs' barriers are initialised before any remote arrive / TMA credit
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_trigger();  // prologue done: the next kernel of the stream may start its own
  pdl_wait();     // the previous kernel's results are complete and visible from here on

  // pair tiles: 256 x 256; both CTAs of a pair walk the same tile sequence
  const int pair = blockIdx.x >> 1, npairs = gridDim.x >> 1;
  const int total_tiles = p.m_tiles * p.n_tiles * p.split_k;  // m_tiles counts 256-row tiles here
  const int kblocks_total = (p.K + GEMM_BK - 1) / GEMM_BK;
  const uint32_t stage_tx = 2u * (GEMM2_A_BYTES + GEMM2_B_BYTES);

  if (warp == 0) {
    // ===================== TMA producer (one per CTA) =====================
    if (lane == 0) {
      uint32_t stage = 0, phase = 0;
      for (int t = pair; t < total_tiles; t += npairs) {
        const int split = t / (p.m_tiles * p.n_tiles);
        const int mn = t - split * (p.m_tiles * p.n_tiles);
        const int m0 = (mn / p.n_tiles) * 256 + (int)rank * 128;
        const int n0 = (mn % p.n_tiles) * GEMM2_BN + (int)rank * 128;
        const int kb0 = split * p.kb_per_split;
        const int kb1 = min(kb0 + p.kb_per_split, kblocks_total);
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait_cluster(&empty_bar[stage], phase ^ 1);
          if (leader) mbar_arrive_expect_tx(&full_bar[stage], stage_tx);
          uint8_t* a_dst = smA + stage * GEMM2_A_BYTES;
          uint8_t* b_dst = smB + stage * GEMM2_B_BYTES;
          const int k0 = kb * GEMM_BK;
          if (!p.a_mn_major) {
            tma_load_2d_pair(a_dst, &tmA, &full_bar[stage], k0, m0);
          } else {
            tma_load_2d_pair(a_dst, &tmA, &full_bar[stage], m0, k0);
            tma_load_2d_pair(a_dst + GEMM_BK * 128, &tmA, &full_bar[stage], m0 + 64, k0);
          }
          if (!p.b_mn_major) {
            tma_load_2d_pair(b_dst, &tmB, &full_bar[stage], k0, n0);
          } else {
            tma_load_2d_pair(b_dst, &tmB, &full_bar[stage], n0, k0);
            tma_load_2d_pair(b_dst + GEMM_BK * 128, &tmB, &full_bar[stage], n0 + 64, k0);
          }
          if (++stage == GEMM2_STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer: the leader CTA's single thread drives both tensor cores =====================
    if (leader && lane == 0) {
      const uint32_t idesc = make_idesc_bf16(256, GEMM2_BN, p.a_mn_major, p.b_mn_major);
      const uint32_t a_lbo = p.a_mn_major ? GEMM_BK * 128 : 16;
      const uint32_t b_lbo = p.b_mn_major ? GEMM_BK * 128 : 16;
      const uint32_t a_kstep = p.a_mn_major ? 16 * 128 : 32;
      const uint32_t b_kstep = p.b_mn_major ? 16 * 128 : 32;
      uint32_t stage = 0, phase = 0;
      uint32_t acc = 0, acc_phase = 0;
      for (int t = pair; t < total_tiles; t += npairs) {
        const int split = t / (p.m_tiles * p.n_tiles);
        const int kb0 = split * p.kb_per_split;
        const int kb1 = min(kb0 + p.kb_per_split, kblocks_total);
        mbar_wait_cluster(&tempty_bar[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * GEMM_MAX_BN;
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait_cluster(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t a_addr = smem_u32(smA + stage * GEMM2_A_BYTES);
          const uint32_t b_addr = smem_u32(smB + stage * GEMM2_B_BYTES);
#pragma unroll
          for (int kk = 0; kk < GEMM_BK / 16; ++kk) {
            const uint64_t adesc = make_smem_desc_sw128(a_addr + kk * a_kstep, a_lbo, 1024);
            const uint64_t bdesc = make_smem_desc_sw128(b_addr + kk * b_kstep, b_lbo, 1024);
            umma_bf16_pair(d_tmem, adesc, bdesc, idesc, (kb > kb0 || kk > 0) ? 1u : 0u);
          }
          umma_commit_pair(&empty_bar[stage]);
          if (++stage == GEMM2_STAGES) { stage = 0; phase ^= 1; }
        }
        umma_commit_pair(&tfull_bar[acc]);
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
  } else if (warp >= 4) {
    // ===================== epilogue: each CTA drains its own 128 rows; two warps per TMEM lane quadrant split the columns ==========
    const int q = warp & 3, half = (warp - 4) >> 2;
    const GemmEpilogue& ep = p.ep;
    uint32_t* st = epi_stage + (warp - 4) * EPI_WARP_WORDS;
    uint32_t acc = 0, acc_phase = 0;
    for (int t = pair; t < total_tiles; t += npairs) {
      const int split = t / (p.m_tiles * p.n_tiles);
      const int mn = t - split * (p.m_tiles * p.n_tiles);
      const int m0 = (mn / p.n_tiles) * 256 + (int)rank * 128;
      const int n0 = (mn % p.n_tiles) * GEMM2_BN;
      const int row0 = m0 + q * 32;
      // chunk i of this warp covers columns n0 + 32 * (half + 2 i): the first chunk's operands are fetched before the accumulator is ready
      EpiLoadF32 Lr;
      EpiLoadBf16 La;
      float bias_lane = 0.f;
      auto prefetch = [&](int col0) {
        if (ep.partial || col0 >= p.N) return;
        const int ncols = min(32, p.N - col0);
        if (ep.bias) bias_lane = lane < ncols ? __ldg(ep.bias + col0 + lane) : 0.f;
        if (ep.residual) epi_issue_f32(Lr, lane, ep.residual, ep.ldr, row0, p.M, col0, ncols);
        if (ep.dact) epi_issue_bf16(La, lane, ep.aux, ep.ldaux, row0, p.M, col0, ncols);
      };
      prefetch(n0 + 32 * half);
      mbar_wait_cluster(&tfull_bar[acc], acc_phase);
      tc_fence_after();
      for (int c = 32 * half; c < GEMM2_BN; c += 64) {
        uint32_t v[32];
        tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + acc * GEMM_MAX_BN + c, v);
        tmem_ld_wait();
        const EpiLoadF32 Lr_cur = Lr;
        const EpiLoadBf16 La_cur = La;
        const float bias_cur = bias_lane;
        if (c + 64 < GEMM2_BN) prefetch(n0 + c + 64);  // next chunk's operands fly while this one is processed
        float f[32];
        gemm_epilogue_chunk_staged(p, ep, v, st, lane, row0, n0 + c, split, bias_cur, Lr_cur, La_cur, f);
      }
      if (ep.partial && ep.sk_counters)
        gemm_splitk_finish(p, ep, ep.sk_counters + ((size_t)mn * 2 + rank) * GEMM2_EPI_WARPS + (warp - 4), lane, row0, n0, 32 * half, 64, GEMM2_BN);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(&tempty_bar[acc], 0);
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  }

  tc_fence_before();
  cluster_sync_all();  // neither CTA may exit (or free TMEM) while the other can still touch its smem / barriers / TMEM
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc_pair(tmem_base, 512);
  }
}

}  // namespace dtb
