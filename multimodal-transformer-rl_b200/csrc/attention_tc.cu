// Causal self-attention with padding mask on tcgen05 tensor cores (head_dim 32 / 64 / 128) — forward and backward.
// Replaces eager_attention_forward (HF:47-65) + the materialised [B,1,T,T] mask (HF:408-414, 631-635) and its autograd:
//   allowed(i, j) = (j <= i) and attention_mask[b, j / 3] > 0        (integer-exact index rule)
// Same tensors and conventions as the mma.sync kernels in attention.cu / attention_bwd.cu (qkv bf16 [B,T,3H] = q | k | v, head h at
// h*d; lse in the base-2 domain of s * scale_log2; fully masked rows -> out 0, lse 0; dS = P (dP - delta) / sqrt(d)).
//
// Structure (all three kernels): one CTA per (128-row tile, head, batch); TMEM lane = tile row, so each of the 128 softmax threads
// owns one row and needs no cross-lane reduction.  Warps 0-3 = softmax / epilogue (TMEM lane quadrant = warp), warp 4 = single-thread
// tcgen05.mma issuer (+ TMEM allocation), warp 5 = TMA producer.  64-wide blocks of the other sequence dimension stream through a
// 2-stage TMA ring; S (and dP) live in TMEM, P / dS go back to the tensor core through 128B-swizzled shared memory written by the
// row-owning threads; accumulators (O, dQ, dK, dV) stay in TMEM for the whole CTA.  Two CTAs per SM overlap each other's
// softmax (MUFU-bound at d = 64) with MMA and TMA.
#include "kernels.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

namespace dtb {

// Per-head_dim geometry.  Row tiles are stored as CH chunks of [rows x 64 bf16] (one 128-byte swizzle row per tile row and chunk), each
// filled by its own TMA box.  D = 128: two chunks; a contraction over head_dim takes KS = 8 UMMA k-steps (4 per chunk) and the head_dim-wide
// accumulators (O, dQ, dK, dV) are NO = 128 TMEM columns (MN-major B operand = the two chunks, LBO = chunk stride).  D = 32: the boxes stay
// 64 columns wide (the upper half is the next head / the next of q|k|v / zero fill beyond the tensor: never used as a contraction index,
// and as an output column it lands in accumulator columns 32..63 that are never read), contractions take KS = 2 k-steps, OC = 1 output chunk.
template <int D>
struct AtCfg {
  static_assert(D == 32 || D == 64 || D == 128, "head_dim 32 / 64 / 128");
  static constexpr int CH = D == 128 ? 2 : 1;
  static constexpr int KS = D / 16;
  static constexpr int NO = D == 128 ? 128 : 64;
  static constexpr int OC = D / 32;
  static constexpr int T128 = CH * 128 * 128;  // bytes of a 128-row tile
  static constexpr int T64 = CH * 64 * 128;    // bytes of a 64-row tile
  static constexpr int CTAS = D == 128 ? 1 : 2;  // CTAs per SM (shared memory / TMEM columns)
};
constexpr int AT_THREADS = 192;      // forward: warps 0-3 softmax, 4 MMA, 5 TMA
// backward: CW compute warps (4 or 8; warps w and w + 4 share TMEM lane quadrant w & 3), then one MMA warp and one TMA warp
constexpr int AT_TILE128 = 128 * 128;  // bytes: 128 rows x 128 B (64 bf16)
constexpr int AT_TILE64 = 64 * 128;    // bytes: 64 rows x 128 B
constexpr int AT_KV_STAGES = 3;        // TMA ring depth for the streamed 64-row blocks (prefetch distance 2 blocks ~ TMA latency)

struct AttnTcParams {
  int T, nh, Kmask;
  int ntile_b;          // batch size (forward: the persistent tile list spans query blocks x heads x batch)
  float scale, scale_log2;
  const float* mask;
  const float* lse_in;
  const bf16* out_in;   // forward output (backward: delta = rowsum(dO * O))
  const bf16* dout;
  bf16* out;
  float* lse;
  bf16* dqkv;
  float* delta;
  DropSite drop;        // attn_dropout on the softmax weights (group = b * nh + h, idx = i * T + j); state == nullptr: off
};

// this CTA's dropout key (0 threshold = off: callers branch on it, warp-uniformly)
__device__ __forceinline__ uint32_t at_drop_key(const AttnTcParams& p, int b, int h) {
  return p.drop.state ? drop_group_key(drop_site_key(p.drop), (uint32_t)(b * p.nh + h)) : 0u;
}

// key-validity bits of the 64-key block starting at k0: bit l of word c <=> key k0 + 32c + l is in range and not padding
__device__ __forceinline__ void at_valid_bits(const AttnTcParams& p, int b, int k0, int lane, uint32_t (&bits)[2]) {
#pragma unroll
  for (int c = 0; c < 2; ++c) {
    const int kj = k0 + 32 * c + lane;
    const bool ok = kj < p.T && (p.mask == nullptr || __ldg(p.mask + (size_t)b * p.Kmask + kj / 3) > 0.f);
    bits[c] = __ballot_sync(0xffffffffu, ok);
  }
}

// the same words for chunks 0..nchunks-1 (32 keys each), written to shared memory once per CTA by all of its warps
constexpr int AT_MAX_CHUNKS = 128;  // T <= 4096
__device__ __forceinline__ void at_fill_valid_bits(const AttnTcParams& p, int b, int nchunks, uint32_t* sBits) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarp = blockDim.x >> 5;
  for (int c = warp; c < nchunks; c += nwarp) {
    const int kj = 32 * c + lane;
    const bool ok = kj < p.T && (p.mask == nullptr || __ldg(p.mask + (size_t)b * p.Kmask + kj / 3) > 0.f);
    const uint32_t w = __ballot_sync(0xffffffffu, ok);
    if (lane == 0) sBits[c] = w;
  }
}

// causal word of a 32-key chunk: bit e set <=> key (chunk_first + e) <= query, given n = query - chunk_first
__device__ __forceinline__ uint32_t at_causal_word(int n) { return n >= 31 ? 0xffffffffu : (n < 0 ? 0u : ((2u << n) - 1u)); }

// row r of a [rows x 64] bf16 K-major tile (128 B per row, 128B swizzle): 16-byte unit u lives at r*128 + ((u ^ (r & 7)) << 4)
__device__ __forceinline__ void at_store_row_unit(uint32_t tile_base, int r, int u, uint32_t w0, uint32_t w1, uint32_t w2, uint32_t w3) {
  const uint32_t addr = tile_base + r * 128 + ((u ^ (r & 7)) << 4);
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(w0), "r"(w1), "r"(w2), "r"(w3) : "memory");
}

// ================================================================================================================================
// forward
// ================================================================================================================================
// smem: Q [128x64] | K[3] [64x64] | V[3] [64x64] | P[2] [128x64]  (all bf16, swizzled)            = 16 + 24 + 24 + 32 KB
// TMEM (256 columns allocated): S fp32 [128 x 64] at column 0, O fp32 [128 x 64] double-buffered at columns 64 / 128
// Persistent over tiles: a CTA walks tiles blockIdx.x, + gridDim.x, ... (heaviest query blocks first).  All three roles keep the
// same running block counter g, so the K/V ring, the S hand-shake and the P buffers run straight through tile boundaries: the next
// tile's Q / K / V loads and its first S = Q K^T are in flight while the softmax threads still normalise and store the current tile's O
// (which is why O is double-buffered in TMEM).  With gridDim.x = number of tiles it degenerates to one tile per CTA.
template <int D, bool DROP>  // DROP = false compiles every dropout branch away (dthr is a constant 0)
__global__ void __launch_bounds__(AT_THREADS, AtCfg<D>::CTAS)
attn_fwd_tc_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmKV, const AttnTcParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  using Cfg = AtCfg<D>;
  constexpr int T128 = Cfg::T128, T64 = Cfg::T64, KS = Cfg::KS, NO = Cfg::NO;
  uint8_t* sQ = smem;
  uint8_t* sK = sQ + T128;
  uint8_t* sV = sK + AT_KV_STAGES * T64;
  uint8_t* sP = sV + AT_KV_STAGES * T64;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sP + 2 * AT_TILE128);
  uint64_t* q_full = bars;          // [1]
  uint64_t* k_full = bars + 1;      // [4]
  uint64_t* k_free = bars + 5;      // [4]
  uint64_t* v_full = bars + 9;      // [4]
  uint64_t* v_free = bars + 13;     // [4]
  uint64_t* s_full = bars + 17;     // [1] MMA -> softmax
  uint64_t* s_free = bars + 18;     // [1] softmax (128) -> MMA
  uint64_t* p_full = bars + 19;     // [2] softmax (128) -> MMA
  uint64_t* pv_done = bars + 21;    // [2] MMA -> softmax
  uint64_t* q_free = bars + 23;     // [1] MMA (last S MMA of a tile retired) -> TMA
  uint64_t* o_free = bars + 24;     // [2] softmax (128: O buffer read out) -> MMA
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 26);
  uint32_t* sBits = reinterpret_cast<uint32_t*>(bars + 27);  // [AT_MAX_CHUNKS] key-validity words, 32 keys each

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int T = p.T;
  const int nkb_max = (T + 63) / 64, ntq = (T + 127) / 128, nhb = p.nh * p.ntile_b, ntiles = ntq * nhb;
  // tile t -> (head-batch pair t / ntq, query block descending within it): the tiles of one (b, h) — which share K / V — are adjacent
  // in the list, i.e. run at the same time on neighbouring CTAs (L2 reuse: DRAM reads 349 -> 177 MB vs a block-major order), while each
  // CTA's own list (stride gridDim.x) still cycles through light and heavy query blocks
  auto tile_qb = [&](int t) { return ntq - 1 - t % ntq; };
  auto tile_nkb = [&](int t) { const int n = (tile_qb(t) * 128 + 128) / 64; return n < nkb_max ? n : nkb_max; };

  if (threadIdx.x == 0) {
    mbar_init(q_full, 1);
    for (int s = 0; s < AT_KV_STAGES; ++s) {
      mbar_init(&k_full[s], 1); mbar_init(&k_free[s], 1); mbar_init(&v_full[s], 1); mbar_init(&v_free[s], 1);
    }
    for (int s = 0; s < 2; ++s) { mbar_init(&p_full[s], 128); mbar_init(&pv_done[s], 1); mbar_init(&o_free[s], 128); }
    mbar_init(s_full, 1);
    mbar_init(s_free, 128);
    mbar_init(q_free, 1);
    fence_mbar_init();
  }
  constexpr uint32_t TMEM_COLS = 64 + 2 * NO <= 256 ? 256 : 512;
  if (warp == 4) { tmem_alloc(tmem_slot, TMEM_COLS); tmem_relinquish(); }
  pdl_trigger();  // prologue done: the next kernel of the stream may start its own
  pdl_wait();     // the previous kernel's results are complete and visible from here on
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tS = tmem_base, tO0 = tmem_base + 64;
  const int H = p.nh * D;

  if (warp == 5) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      tma_prefetch_desc(&tmQ);
      tma_prefetch_desc(&tmKV);
      int g = 0, tl = 0;
      for (int t = blockIdx.x; t < ntiles; t += gridDim.x, ++tl) {
        const int hb = t / ntq, h = hb % p.nh, b = hb / p.nh, q0 = tile_qb(t) * 128, nkb = tile_nkb(t);
        if (tl > 0) mbar_wait(q_free, (tl - 1) & 1);  // the previous tile's last S MMA has read Q
        mbar_arrive_expect_tx(q_full, T128);
#pragma unroll
        for (int c = 0; c < Cfg::CH; ++c) tma_load_3d(sQ + c * AT_TILE128, &tmQ, q_full, h * D + 64 * c, q0, b);
        for (int j = 0; j < nkb; ++j, ++g) {
          const int s = g % AT_KV_STAGES, u = g / AT_KV_STAGES;
          if (u > 0) mbar_wait(&k_free[s], (u - 1) & 1);
          mbar_arrive_expect_tx(&k_full[s], T64);
#pragma unroll
          for (int c = 0; c < Cfg::CH; ++c) tma_load_3d(sK + s * T64 + c * AT_TILE64, &tmKV, &k_full[s], H + h * D + 64 * c, j * 64, b);
          if (u > 0) mbar_wait(&v_free[s], (u - 1) & 1);
          mbar_arrive_expect_tx(&v_full[s], T64);
#pragma unroll
          for (int c = 0; c < Cfg::CH; ++c) tma_load_3d(sV + s * T64 + c * AT_TILE64, &tmKV, &v_full[s], 2 * H + h * D + 64 * c, j * 64, b);
        }
      }
    }
  } else if (warp == 4) {
    // ===================== MMA issuer (single thread) =====================
    if (lane == 0) {
      const uint32_t idesc_qk = make_idesc_bf16(128, 64, 0, 0);
      const uint32_t idesc_pv = make_idesc_bf16(128, NO, 0, 1);
      const uint32_t q_addr = smem_u32(sQ);
      // S(g) = Q K_g^T; `last` = last block of its tile: its retirement frees the Q buffer for the next tile's load
      auto issue_qk = [&](int g, bool last) {
        const int s = g % AT_KV_STAGES;
        mbar_wait(&k_full[s], (g / AT_KV_STAGES) & 1);
        if (g > 0) mbar_wait(s_free, (g - 1) & 1);
        tc_fence_after();
        const uint32_t k_addr = smem_u32(sK + s * T64);
#pragma unroll
        for (int kk = 0; kk < KS; ++kk)  // contraction over head_dim: 4 k-steps per 64-column chunk
          umma_bf16(tS, make_smem_desc_sw128(q_addr + (kk >> 2) * AT_TILE128 + (kk & 3) * 32, 16, 1024),
                    make_smem_desc_sw128(k_addr + (kk >> 2) * AT_TILE64 + (kk & 3) * 32, 16, 1024), idesc_qk, kk > 0 ? 1u : 0u);
        umma_commit(&k_free[s]);
        if (last) umma_commit(q_free);
        umma_commit(s_full);
      };
      int g = 0, tl = 0;
      if (blockIdx.x < ntiles) {
        mbar_wait(q_full, 0);
        issue_qk(0, tile_nkb(blockIdx.x) == 1);
      }
      for (int t = blockIdx.x; t < ntiles; t += gridDim.x, ++tl) {
        const int nkb = tile_nkb(t), tn = t + gridDim.x;
        const uint32_t tO = tO0 + NO * (tl & 1);
        for (int j = 0; j < nkb; ++j, ++g) {
          // look-ahead S: the next block of this tile, or (crossing the tile boundary) the first block of the CTA's next tile
          if (j + 1 < nkb) {
            issue_qk(g + 1, j + 2 == nkb);
          } else if (tn < ntiles) {
            mbar_wait(q_full, (tl + 1) & 1);
            issue_qk(g + 1, tile_nkb(tn) == 1);
          }
          const int s = g % AT_KV_STAGES, pb = g & 1;
          mbar_wait(&v_full[s], (g / AT_KV_STAGES) & 1);
          mbar_wait(&p_full[pb], (g >> 1) & 1);
          if (j == 0 && tl >= 2) mbar_wait(&o_free[tl & 1], ((tl - 2) >> 1) & 1);  // this O buffer was read out (tile tl - 2)
          tc_fence_after();
          const uint32_t p_addr = smem_u32(sP + pb * AT_TILE128), v_addr = smem_u32(sV + s * T64);
#pragma unroll
          for (int kk = 0; kk < 4; ++kk)  // contraction over the 64 keys; V is the MN-major B operand (N = NO: LBO = chunk stride)
            umma_bf16(tO, make_smem_desc_sw128(p_addr + kk * 32, 16, 1024), make_smem_desc_sw128(v_addr + kk * 2048, AT_TILE64, 1024), idesc_pv,
                      (j > 0 || kk > 0) ? 1u : 0u);
          umma_commit(&v_free[s]);
          umma_commit(&pv_done[pb]);
        }
      }
    }
  } else {
    // ===================== softmax: one thread per query row =====================
    const int r = warp * 32 + lane;
    const uint32_t lane_addr = (uint32_t)(warp * 32) << 16;
    const uint32_t dthr = DROP ? p.drop.thr : 0u;
    int g = 0, tl = 0;
    for (int t = blockIdx.x; t < ntiles; t += gridDim.x, ++tl) {
    const int hb = t / ntq, h = hb % p.nh, b = hb / p.nh, q0 = tile_qb(t) * 128, nkb = tile_nkb(t);
    const int qi = q0 + r;
    const uint32_t tO = tO0 + NO * (tl & 1);
    const uint32_t drow = DROP ? at_drop_key(p, b, h) + (uint32_t)qi * (uint32_t)T : 0u;
    // key-validity words of this tile's batch row (the four softmax warps only: named barrier 1)
    asm volatile("bar.sync 1, 128;" ::: "memory");
    for (int c = warp; c < 2 * nkb; c += 4) {
      const int kj = 32 * c + lane;
      const bool ok = kj < T && (p.mask == nullptr || __ldg(p.mask + (size_t)b * p.Kmask + kj / 3) > 0.f);
      const uint32_t w = __ballot_sync(0xffffffffu, ok);
      if (lane == 0) sBits[c] = w;
    }
    asm volatile("bar.sync 1, 128;" ::: "memory");
    float m_use = -INFINITY, l = 0.f;
    for (int j = 0; j < nkb; ++j, ++g) {
      const int k0 = j * 64;
      const uint32_t bits[2] = {sBits[2 * j], sBits[2 * j + 1]};
      const bool diag = k0 + 63 > q0;  // some key of the block is beyond some query of the tile
      // chunk c of this warp is entirely above the diagonal when its first key exceeds the warp's last row
      const bool act0 = !diag || (k0 <= q0 + warp * 32 + 31), act1 = !diag || (k0 + 32 <= q0 + warp * 32 + 31);
      mbar_wait(s_full, g & 1);
      tc_fence_after();
      uint32_t v0[32], v1[32];
      if (act0) tmem_ld_32x32(tS + lane_addr, v0);
      if (act1) tmem_ld_32x32(tS + lane_addr + 32, v1);
      tmem_ld_wait();
      tc_fence_before();
      mbar_arrive(s_free);
      // ---- mask (branch-free: one allowed-key word per 32-key chunk and row) + row max (raw domain; scale > 0) ----
      uint32_t m0 = 0u, m1 = 0u;
      if (act0) m0 = bits[0] & (diag ? at_causal_word(qi - k0) : 0xffffffffu);
      if (act1) m1 = bits[1] & (diag ? at_causal_word(qi - k0 - 32) : 0xffffffffu);
      const bool all_open = __all_sync(0xffffffffu, (m0 & m1) == 0xffffffffu);  // warp-uniform fast path: nothing to mask
      float mxa = -INFINITY, mxb = -INFINITY, mxc = -INFINITY, mxd = -INFINITY;
      if (!all_open) {
        if (act0) {
#pragma unroll
          for (int e = 0; e < 32; ++e) v0[e] = (m0 & (1u << e)) ? v0[e] : 0xff800000u;  // -inf
        }
        if (act1) {
#pragma unroll
          for (int e = 0; e < 32; ++e) v1[e] = (m1 & (1u << e)) ? v1[e] : 0xff800000u;
        }
      }
      if (act0) {
#pragma unroll
        for (int e = 0; e < 32; e += 4) {
          mxa = fmaxf(mxa, __uint_as_float(v0[e])); mxb = fmaxf(mxb, __uint_as_float(v0[e + 1]));
          mxc = fmaxf(mxc, __uint_as_float(v0[e + 2])); mxd = fmaxf(mxd, __uint_as_float(v0[e + 3]));
        }
      }
      if (act1) {
#pragma unroll
        for (int e = 0; e < 32; e += 4) {
          mxa = fmaxf(mxa, __uint_as_float(v1[e])); mxb = fmaxf(mxb, __uint_as_float(v1[e + 1]));
          mxc = fmaxf(mxc, __uint_as_float(v1[e + 2])); mxd = fmaxf(mxd, __uint_as_float(v1[e + 3]));
        }
      }
      const float mxs = fmaxf(fmaxf(mxa, mxb), fmaxf(mxc, mxd)) * p.scale_log2;
      // lazy rescale: the running maximum only moves when it grows by more than 8 (p <= 2^8 stays exact enough in bf16 / fp32)
      float alpha = 1.f;
      bool moved = false;
      if (j == 0) {
        m_use = mxs;
      } else if (mxs > m_use + 8.f) {
        alpha = fast_exp2(m_use - mxs);  // m_use = -inf -> 0
        m_use = mxs;
        moved = true;
      }
      const float nms = m_use == -INFINITY ? 0.f : -m_use;
      // ---- P = exp2(s * scale_log2 - m), row sum, bf16 pack ----
      const int s = g & 1;
      if (g >= 2) mbar_wait(&pv_done[s], ((g - 2) >> 1) & 1);  // P buffer s was last read by PV(g - 2) (possibly the previous tile's)
      const uint32_t p_tile = smem_u32(sP + s * AT_TILE128);
      float rsa = 0.f, rsb = 0.f, rsc = 0.f, rsd = 0.f;
      if (act0) {
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          float f[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) f[e] = fast_exp2(fmaf(__uint_as_float(v0[u * 8 + e]), p.scale_log2, nms));
          rsa += f[0] + f[4]; rsb += f[1] + f[5]; rsc += f[2] + f[6]; rsd += f[3] + f[7];
          if (dthr) {  // dropped weights leave the PV product, not the normaliser; 1 / (1 - p) is applied to O at the end
#pragma unroll
            for (int e = 0; e < 8; ++e) f[e] = drop_keep(drow, (uint32_t)(k0 + u * 8 + e), dthr) ? f[e] : 0.f;
          }
          at_store_row_unit(p_tile, r, u, pack_bf16x2(f[0], f[1]), pack_bf16x2(f[2], f[3]), pack_bf16x2(f[4], f[5]), pack_bf16x2(f[6], f[7]));
        }
      } else {
#pragma unroll
        for (int u = 0; u < 4; ++u) at_store_row_unit(p_tile, r, u, 0u, 0u, 0u, 0u);
      }
      if (act1) {
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          float f[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) f[e] = fast_exp2(fmaf(__uint_as_float(v1[u * 8 + e]), p.scale_log2, nms));
          rsa += f[0] + f[4]; rsb += f[1] + f[5]; rsc += f[2] + f[6]; rsd += f[3] + f[7];
          if (dthr) {
#pragma unroll
            for (int e = 0; e < 8; ++e) f[e] = drop_keep(drow, (uint32_t)(k0 + 32 + u * 8 + e), dthr) ? f[e] : 0.f;
          }
          at_store_row_unit(p_tile, r, 4 + u, pack_bf16x2(f[0], f[1]), pack_bf16x2(f[2], f[3]), pack_bf16x2(f[4], f[5]), pack_bf16x2(f[6], f[7]));
        }
      } else {
#pragma unroll
        for (int u = 0; u < 4; ++u) at_store_row_unit(p_tile, r, 4 + u, 0u, 0u, 0u, 0u);
      }
      const float rs = (rsa + rsb) + (rsc + rsd);
      l = l * alpha + rs;
      // ---- rescale O when some row of this warp moved its maximum (PV(j-1) must have retired) ----
      if (__any_sync(0xffffffffu, moved)) {
        mbar_wait(&pv_done[(g - 1) & 1], ((g - 1) >> 1) & 1);  // moved implies j > 0: PV(g - 1) belongs to this tile
        tc_fence_after();
#pragma unroll
        for (int c = 0; c < Cfg::OC; ++c) {
          uint32_t o[32];
          tmem_ld_32x32(tO + lane_addr + c * 32, o);
          tmem_ld_wait();
#pragma unroll
          for (int e = 0; e < 32; ++e) o[e] = __float_as_uint(__uint_as_float(o[e]) * alpha);
          tmem_st_32x32(tO + lane_addr + c * 32, o);
        }
        tmem_st_wait();
      }
      fence_proxy_async();  // P stores (generic proxy) -> visible to the tensor core
      tc_fence_before();
      mbar_arrive(&p_full[s]);
    }
    // ---- epilogue: O / l -> bf16, lse ----
    mbar_wait(&pv_done[(g - 1) & 1], ((g - 1) >> 1) & 1);
    tc_fence_after();
    const float inv = l > 0.f ? (dthr ? p.drop.scale : 1.f) / l : 0.f;
    bf16* orow = p.out + ((size_t)b * T + qi) * H + h * D;
#pragma unroll
    for (int c = 0; c < Cfg::OC; ++c) {
      uint32_t o[32];
      tmem_ld_32x32(tO + lane_addr + c * 32, o);
      tmem_ld_wait();
      if (qi < T) {
#pragma unroll
        for (int e = 0; e < 32; e += 8)
          *reinterpret_cast<uint4*>(orow + c * 32 + e) =
              make_uint4(pack_bf16x2(__uint_as_float(o[e]) * inv, __uint_as_float(o[e + 1]) * inv),
                         pack_bf16x2(__uint_as_float(o[e + 2]) * inv, __uint_as_float(o[e + 3]) * inv),
                         pack_bf16x2(__uint_as_float(o[e + 4]) * inv, __uint_as_float(o[e + 5]) * inv),
                         pack_bf16x2(__uint_as_float(o[e + 6]) * inv, __uint_as_float(o[e + 7]) * inv));
      }
    }
    if (qi < T) p.lse[((size_t)b * p.nh + h) * T + qi] = l > 0.f ? m_use + log2f(l) : 0.f;
    tc_fence_before();
    mbar_arrive(&o_free[tl & 1]);  // this thread's rows of the O buffer are in registers / stored: tile tl + 2 may overwrite them
    }  // tile loop
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 4) { tc_fence_after(); tmem_dealloc(tmem_base, TMEM_COLS); }
}

// ================================================================================================================================
// backward, pass 1: delta = rowsum(dO * O) and dQ = sum_j dS_ij K_j        (one CTA per 128-query tile)
// ================================================================================================================================
// smem: Q [128x64] | dO [128x64] | K[3] [64x64] | V[2] [64x64] | dS[2] [128x64]                     = 16 + 16 + 24 + 16 + 32 KB
// TMEM (256 columns): S at 0, dP at 64, dQ at 128 (fp32 [128 x 64] each)
constexpr int AT_DQ_KST = 3, AT_DQ_VST = 2;

// one 32-column half of a row: P = exp2(s * scale_log2 - lse), dS = P (dP - delta) / sqrt(d) -> bf16 units u0..u0+3 of the dS tile row
// dropout (dthr != 0): dP = D o (dO V^T) with D = keep / (1 - p) recomputed from dbase + column
__device__ __forceinline__ void at_ds_half(const uint32_t (&sv)[32], const uint32_t (&dv)[32], uint32_t word, float scale_log2, float nlse,
                                           float delta, float scale, uint32_t tile, int r, int u0, uint32_t dbase, uint32_t dthr,
                                           float dscale) {
#pragma unroll
  for (int u = 0; u < 4; ++u) {
    float f[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int c = u * 8 + e;
      float pe = fast_exp2(fmaf(__uint_as_float(sv[c]), scale_log2, nlse));
      pe = (word & (1u << c)) ? pe : 0.f;
      float g = __uint_as_float(dv[c]);
      if (dthr) g = drop_keep(dbase, (uint32_t)c, dthr) ? g * dscale : 0.f;
      f[e] = pe * (g - delta) * scale;
    }
    at_store_row_unit(tile, r, u0 + u, pack_bf16x2(f[0], f[1]), pack_bf16x2(f[2], f[3]), pack_bf16x2(f[4], f[5]), pack_bf16x2(f[6], f[7]));
  }
}

template <int D, bool DROP, int CW>  // DROP = false compiles every dropout branch away; CW = compute warps (4: both halves per thread, 8: one)
__global__ void __launch_bounds__(32 * (CW + 2), AtCfg<D>::CTAS)
attn_bwd_dq_tc_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmKV,
                      const __grid_constant__ CUtensorMap tmDO, const AttnTcParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  using Cfg = AtCfg<D>;
  constexpr int T128 = Cfg::T128, T64 = Cfg::T64, KS = Cfg::KS, NO = Cfg::NO;
  uint8_t* sQ = smem;
  uint8_t* sDO = sQ + T128;
  uint8_t* sK = sDO + T128;
  uint8_t* sV = sK + AT_DQ_KST * T64;
  uint8_t* sDS = sV + AT_DQ_VST * T64;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sDS + 2 * AT_TILE128);
  uint64_t* qdo_full = bars;        // [1]
  uint64_t* k_full = bars + 1;      // [3]
  uint64_t* k_free = bars + 4;      // [3]
  uint64_t* v_full = bars + 7;      // [2]
  uint64_t* v_free = bars + 9;      // [2]
  uint64_t* sdp_full = bars + 11;   // [1] MMA -> threads
  uint64_t* sdp_free = bars + 12;   // [1] threads (256) -> MMA
  uint64_t* ds_full = bars + 13;    // [2] threads (256) -> MMA
  uint64_t* ds_done = bars + 15;    // [2] MMA -> threads
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 17);
  uint32_t* sBits = reinterpret_cast<uint32_t*>(bars + 18);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int qb = gridDim.x - 1 - blockIdx.x;
  const int h = blockIdx.y, b = blockIdx.z;
  const int q0 = qb * 128;
  const int T = p.T;
  int nkb = (q0 + 128) / 64;
  const int nkb_max = (T + 63) / 64;
  if (nkb > nkb_max) nkb = nkb_max;

  if (threadIdx.x == 0) {
    mbar_init(qdo_full, 1);
    for (int s = 0; s < AT_DQ_KST; ++s) { mbar_init(&k_full[s], 1); mbar_init(&k_free[s], 1); }
    for (int s = 0; s < AT_DQ_VST; ++s) { mbar_init(&v_full[s], 1); mbar_init(&v_free[s], 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(&ds_full[s], 32 * CW); mbar_init(&ds_done[s], 1); }
    mbar_init(sdp_full, 1);
    mbar_init(sdp_free, 32 * CW);
    fence_mbar_init();
  }
  if (warp == CW) { tmem_alloc(tmem_slot, 256); tmem_relinquish(); }
  pdl_trigger();  // prologue done: the next kernel of the stream may start its own
  pdl_wait();     // the previous kernel's results are complete and visible from here on
  at_fill_valid_bits(p, b, 2 * nkb, sBits);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tS = tmem_base, tDP = tmem_base + 64, tDQ = tmem_base + 128;  // 128 + NO <= 256 columns
  const int H = p.nh * D;

  if (warp == CW + 1) {
    if (lane == 0) {
      tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmKV); tma_prefetch_desc(&tmDO);
      mbar_arrive_expect_tx(qdo_full, 2 * T128);
#pragma unroll
      for (int c = 0; c < Cfg::CH; ++c) {
        tma_load_3d(sQ + c * AT_TILE128, &tmQ, qdo_full, h * D + 64 * c, q0, b);
        tma_load_3d(sDO + c * AT_TILE128, &tmDO, qdo_full, h * D + 64 * c, q0, b);
      }
      for (int j = 0; j < nkb; ++j) {
        const int sk = j % AT_DQ_KST, uk = j / AT_DQ_KST, sv = j % AT_DQ_VST, uv = j / AT_DQ_VST;
        if (uk > 0) mbar_wait(&k_free[sk], (uk - 1) & 1);
        mbar_arrive_expect_tx(&k_full[sk], T64);
#pragma unroll
        for (int c = 0; c < Cfg::CH; ++c) tma_load_3d(sK + sk * T64 + c * AT_TILE64, &tmKV, &k_full[sk], H + h * D + 64 * c, j * 64, b);
        if (uv > 0) mbar_wait(&v_free[sv], (uv - 1) & 1);
        mbar_arrive_expect_tx(&v_full[sv], T64);
#pragma unroll
        for (int c = 0; c < Cfg::CH; ++c) tma_load_3d(sV + sv * T64 + c * AT_TILE64, &tmKV, &v_full[sv], 2 * H + h * D + 64 * c, j * 64, b);
      }
    }
  } else if (warp == CW) {
    if (lane == 0) {
      const uint32_t idesc_kk = make_idesc_bf16(128, 64, 0, 0), idesc_kn = make_idesc_bf16(128, NO, 0, 1);
      const uint32_t q_addr = smem_u32(sQ), do_addr = smem_u32(sDO);
      auto issue_sdp = [&](int j) {
        const int sk = j % AT_DQ_KST, sv = j % AT_DQ_VST;
        mbar_wait(&k_full[sk], (j / AT_DQ_KST) & 1);
        mbar_wait(&v_full[sv], (j / AT_DQ_VST) & 1);
        if (j > 0) mbar_wait(sdp_free, (j - 1) & 1);
        tc_fence_after();
        const uint32_t k_addr = smem_u32(sK + sk * T64), v_addr = smem_u32(sV + sv * T64);
#pragma unroll
        for (int kk = 0; kk < KS; ++kk)
          umma_bf16(tS, make_smem_desc_sw128(q_addr + (kk >> 2) * AT_TILE128 + (kk & 3) * 32, 16, 1024),
                    make_smem_desc_sw128(k_addr + (kk >> 2) * AT_TILE64 + (kk & 3) * 32, 16, 1024), idesc_kk, kk > 0 ? 1u : 0u);
#pragma unroll
        for (int kk = 0; kk < KS; ++kk)
          umma_bf16(tDP, make_smem_desc_sw128(do_addr + (kk >> 2) * AT_TILE128 + (kk & 3) * 32, 16, 1024),
                    make_smem_desc_sw128(v_addr + (kk >> 2) * AT_TILE64 + (kk & 3) * 32, 16, 1024), idesc_kk, kk > 0 ? 1u : 0u);
        umma_commit(&v_free[sv]);
        umma_commit(sdp_full);
      };
      mbar_wait(qdo_full, 0);
      issue_sdp(0);
      for (int j = 0; j < nkb; ++j) {
        if (j + 1 < nkb) issue_sdp(j + 1);
        const int sk = j % AT_DQ_KST, db = j & 1;
        mbar_wait(&ds_full[db], (j >> 1) & 1);
        tc_fence_after();
        const uint32_t ds_addr = smem_u32(sDS + db * AT_TILE128), k_addr = smem_u32(sK + sk * T64);
#pragma unroll
        for (int kk = 0; kk < 4; ++kk)
          umma_bf16(tDQ, make_smem_desc_sw128(ds_addr + kk * 32, 16, 1024), make_smem_desc_sw128(k_addr + kk * 2048, AT_TILE64, 1024), idesc_kn,
                    (j > 0 || kk > 0) ? 1u : 0u);
        umma_commit(&k_free[sk]);
        umma_commit(&ds_done[db]);
      }
    }
  } else {
    // compute warps: TMEM lane quadrant qd = w & 3 (rows 32 qd .. +31, one per thread).  CW = 8: warp w owns the 32-column half w >> 2 of
    // every 64-key block; CW = 4: every thread walks both halves.
    const int qd = warp & 3, h_lo = CW == 8 ? (warp >> 2) : 0, h_hi = CW == 8 ? h_lo + 1 : 2;
    const int r = qd * 32 + lane;
    const int qi = q0 + r;
    const uint32_t lane_base = (uint32_t)(qd * 32) << 16;
    const size_t row_off = ((size_t)b * p.nh + h) * T;
    // delta = rowsum(dO * O) for this thread's row, straight from global (128 B each)
    float delta = 0.f, nlse = 0.f;
    const uint32_t dthr = DROP ? p.drop.thr : 0u;
    const uint32_t drow = DROP ? at_drop_key(p, b, h) + (uint32_t)qi * (uint32_t)T : 0u;
    if (qi < T) {
      const uint4* orow = reinterpret_cast<const uint4*>(p.out_in + ((size_t)b * T + qi) * H + h * D);
      const uint4* drow_g = reinterpret_cast<const uint4*>(p.dout + ((size_t)b * T + qi) * H + h * D);
#pragma unroll
      for (int u = 0; u < D / 8; ++u) {
        const uint4 a = __ldg(orow + u), g = __ldg(drow_g + u);
        const uint32_t aw[4] = {a.x, a.y, a.z, a.w}, gw[4] = {g.x, g.y, g.z, g.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float2 x = unpack_bf16x2(aw[e]), y = unpack_bf16x2(gw[e]);
          delta += x.x * y.x + x.y * y.y;
        }
      }
      if (h_lo == 0) p.delta[row_off + qi] = delta;
      nlse = -__ldg(p.lse_in + row_off + qi);
    }
    for (int j = 0; j < nkb; ++j) {
      const bool diag = j * 64 + 63 > q0;
      const int db = j & 1;
      const uint32_t tile = smem_u32(sDS + db * AT_TILE128);
      mbar_wait(sdp_full, j & 1);
      tc_fence_after();
#pragma unroll
      for (int half = h_lo; half < h_hi; ++half) {
        const int k0 = j * 64 + 32 * half;  // first key of the 32-key chunk
        const bool act = !diag || (k0 <= q0 + qd * 32 + 31);
        uint32_t m = 0u;
        if (act) m = sBits[2 * j + half] & (diag ? at_causal_word(qi - k0) : 0xffffffffu);
        uint32_t sv[32], dv[32];
        if (act) {
          tmem_ld_32x32(tS + lane_base + 32 * half, sv);
          tmem_ld_32x32(tDP + lane_base + 32 * half, dv);
          tmem_ld_wait();
        }
        if (half == h_hi - 1) {  // S / dP of this block are in registers: the next block's MMAs may overwrite them
          tc_fence_before();
          mbar_arrive(sdp_free);
        }
        if (half == h_lo && j >= 2) mbar_wait(&ds_done[db], ((j - 2) >> 1) & 1);  // dS buffer db was last read by the dQ MMAs of block j - 2
        if (act) {
          at_ds_half(sv, dv, m, p.scale_log2, nlse, delta, p.scale, tile, r, 4 * half, drow + (uint32_t)k0, dthr, p.drop.scale);
        } else {
#pragma unroll
          for (int u = 0; u < 4; ++u) at_store_row_unit(tile, r, 4 * half + u, 0u, 0u, 0u, 0u);
        }
      }
      fence_proxy_async();
      mbar_arrive(&ds_full[db]);
    }
    mbar_wait(&ds_done[(nkb - 1) & 1], ((nkb - 1) >> 1) & 1);
    tc_fence_after();
    // dQ read-out: 32-column chunks of the head; CW = 8: the two warps of a lane quadrant take alternate chunks
#pragma unroll
    for (int half = h_lo; half < Cfg::OC; half += (CW == 8 ? 2 : 1)) {
      bf16* grow = p.dqkv + ((size_t)b * T + qi) * 3 * H + h * D + 32 * half;
      uint32_t o[32];
      tmem_ld_32x32(tDQ + lane_base + 32 * half, o);
      tmem_ld_wait();
      if (qi < T) {
#pragma unroll
        for (int e = 0; e < 32; e += 8)
          *reinterpret_cast<uint4*>(grow + e) =
              make_uint4(pack_bf16x2(__uint_as_float(o[e]), __uint_as_float(o[e + 1])), pack_bf16x2(__uint_as_float(o[e + 2]), __uint_as_float(o[e + 3])),
                         pack_bf16x2(__uint_as_float(o[e + 4]), __uint_as_float(o[e + 5])), pack_bf16x2(__uint_as_float(o[e + 6]), __uint_as_float(o[e + 7])));
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == CW) { tc_fence_after(); tmem_dealloc(tmem_base, 256); }
}

// ================================================================================================================================
// backward, pass 2: dV = P^T dO, dK = dS^T Q        (one CTA per 128-key tile; TMEM lane = key, columns = queries)
// ================================================================================================================================
// smem: K [128x64] | V [128x64] | Q[2] [64x64] | dO[2] [64x64] | P^T [128x64] | dS^T [128x64] | lse[Tp] | delta[Tp]
// TMEM (256 columns): S^T at 0, dP^T at 64, dV at 128, dK at 192
constexpr int AT_MAX_T_BWD = 1024;

template <int D, bool DROP, int CW>
__global__ void __launch_bounds__(32 * (CW + 2), AtCfg<D>::CTAS)
attn_bwd_dkv_tc_kernel(const __grid_constant__ CUtensorMap tmKV, const __grid_constant__ CUtensorMap tmQ64,
                       const __grid_constant__ CUtensorMap tmDO64, const AttnTcParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  using Cfg = AtCfg<D>;
  constexpr int T128 = Cfg::T128, T64 = Cfg::T64, KS = Cfg::KS, NO = Cfg::NO;
  uint8_t* sK = smem;
  uint8_t* sV = sK + T128;
  uint8_t* sQ = sV + T128;
  uint8_t* sDO = sQ + 2 * T64;
  uint8_t* sPT = sDO + 2 * T64;
  uint8_t* sDST = sPT + AT_TILE128;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sDST + AT_TILE128);
  uint64_t* kv_full = bars;         // [1]
  uint64_t* qdo_full = bars + 1;    // [2]
  uint64_t* qdo_free = bars + 3;    // [2]
  uint64_t* sdp_full = bars + 5;    // [1]
  uint64_t* sdp_free = bars + 6;    // [1] (256)
  uint64_t* pds_full = bars + 7;    // [1] (256)
  uint64_t* pds_done = bars + 8;    // [1]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 9);
  float* sLse = reinterpret_cast<float*>(bars + 10);
  float* sDelta = sLse + AT_MAX_T_BWD;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int kb = blockIdx.x, h = blockIdx.y, b = blockIdx.z;  // key tile 0 has the most query blocks: heaviest first
  const int k0 = kb * 128;
  const int T = p.T;
  const int i0 = k0 / 64, nqb = (T + 63) / 64, n = nqb - i0;
  const size_t row_off = ((size_t)b * p.nh + h) * T;

  if (threadIdx.x == 0) {
    mbar_init(kv_full, 1);
    for (int s = 0; s < 2; ++s) { mbar_init(&qdo_full[s], 1); mbar_init(&qdo_free[s], 1); }
    mbar_init(sdp_full, 1);
    mbar_init(sdp_free, 32 * CW);
    mbar_init(pds_full, 32 * CW);
    mbar_init(pds_done, 1);
    fence_mbar_init();
  }
  constexpr uint32_t TMEM_COLS = 128 + 2 * NO <= 256 ? 256 : 512;
  if (warp == CW) { tmem_alloc(tmem_slot, TMEM_COLS); tmem_relinquish(); }
  pdl_trigger();  // prologue done: the next kernel of the stream may start its own
  pdl_wait();     // the previous kernel's results are complete and visible from here on
  for (int t = i0 * 64 + threadIdx.x; t < nqb * 64; t += blockDim.x) {
    sLse[t] = t < T ? -__ldg(p.lse_in + row_off + t) : 0.f;   // negated: exponent = s * scale_log2 + (-lse)
    sDelta[t] = t < T ? __ldg(p.delta + row_off + t) : 0.f;
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tST = tmem_base, tDPT = tmem_base + 64, tDV = tmem_base + 128, tDK = tmem_base + 128 + NO;
  const int H = p.nh * D;

  if (warp == CW + 1) {
    if (lane == 0) {
      tma_prefetch_desc(&tmKV); tma_prefetch_desc(&tmQ64); tma_prefetch_desc(&tmDO64);
      mbar_arrive_expect_tx(kv_full, 2 * T128);
#pragma unroll
      for (int c = 0; c < Cfg::CH; ++c) {
        tma_load_3d(sK + c * AT_TILE128, &tmKV, kv_full, H + h * D + 64 * c, k0, b);
        tma_load_3d(sV + c * AT_TILE128, &tmKV, kv_full, 2 * H + h * D + 64 * c, k0, b);
      }
      for (int t = 0; t < n; ++t) {
        const int s = t & 1, u = t >> 1;
        if (u > 0) mbar_wait(&qdo_free[s], (u - 1) & 1);
        mbar_arrive_expect_tx(&qdo_full[s], 2 * T64);
#pragma unroll
        for (int c = 0; c < Cfg::CH; ++c) {
          tma_load_3d(sQ + s * T64 + c * AT_TILE64, &tmQ64, &qdo_full[s], h * D + 64 * c, (i0 + t) * 64, b);
          tma_load_3d(sDO + s * T64 + c * AT_TILE64, &tmDO64, &qdo_full[s], h * D + 64 * c, (i0 + t) * 64, b);
        }
      }
    }
  } else if (warp == CW) {
    if (lane == 0) {
      const uint32_t idesc_kk = make_idesc_bf16(128, 64, 0, 0), idesc_kn = make_idesc_bf16(128, NO, 0, 1);
      const uint32_t k_addr = smem_u32(sK), v_addr = smem_u32(sV), pt_addr = smem_u32(sPT), dst_addr = smem_u32(sDST);
      auto issue_sdp = [&](int t) {
        const int s = t & 1;
        mbar_wait(&qdo_full[s], (t >> 1) & 1);
        if (t > 0) mbar_wait(sdp_free, (t - 1) & 1);
        tc_fence_after();
        const uint32_t q_addr = smem_u32(sQ + s * T64), do_addr = smem_u32(sDO + s * T64);
#pragma unroll
        for (int kk = 0; kk < KS; ++kk)
          umma_bf16(tST, make_smem_desc_sw128(k_addr + (kk >> 2) * AT_TILE128 + (kk & 3) * 32, 16, 1024),
                    make_smem_desc_sw128(q_addr + (kk >> 2) * AT_TILE64 + (kk & 3) * 32, 16, 1024), idesc_kk, kk > 0 ? 1u : 0u);
#pragma unroll
        for (int kk = 0; kk < KS; ++kk)
          umma_bf16(tDPT, make_smem_desc_sw128(v_addr + (kk >> 2) * AT_TILE128 + (kk & 3) * 32, 16, 1024),
                    make_smem_desc_sw128(do_addr + (kk >> 2) * AT_TILE64 + (kk & 3) * 32, 16, 1024), idesc_kk, kk > 0 ? 1u : 0u);
        umma_commit(sdp_full);
      };
      mbar_wait(kv_full, 0);
      issue_sdp(0);
      for (int t = 0; t < n; ++t) {
        if (t + 1 < n) issue_sdp(t + 1);
        const int s = t & 1;
        mbar_wait(pds_full, t & 1);
        tc_fence_after();
        const uint32_t q_addr = smem_u32(sQ + s * T64), do_addr = smem_u32(sDO + s * T64);
#pragma unroll
        for (int kk = 0; kk < 4; ++kk)
          umma_bf16(tDV, make_smem_desc_sw128(pt_addr + kk * 32, 16, 1024), make_smem_desc_sw128(do_addr + kk * 2048, AT_TILE64, 1024), idesc_kn,
                    (t > 0 || kk > 0) ? 1u : 0u);
#pragma unroll
        for (int kk = 0; kk < 4; ++kk)
          umma_bf16(tDK, make_smem_desc_sw128(dst_addr + kk * 32, 16, 1024), make_smem_desc_sw128(q_addr + kk * 2048, AT_TILE64, 1024), idesc_kn,
                    (t > 0 || kk > 0) ? 1u : 0u);
        umma_commit(&qdo_free[s]);
        umma_commit(pds_done);
      }
    }
  } else {
    // compute warps: TMEM lane quadrant qd = w & 3 (keys 32 qd .. +31, one per thread); CW = 8: warp w owns the 32-query half w >> 2 of
    // every 64-query block, CW = 4: every thread walks both halves
    const int qd = warp & 3, h_lo = CW == 8 ? (warp >> 2) : 0, h_hi = CW == 8 ? h_lo + 1 : 2;
    const int r = qd * 32 + lane;
    const int kj = k0 + r;
    const uint32_t lane_base = (uint32_t)(qd * 32) << 16;
    const bool key_ok = kj < T && (p.mask == nullptr || __ldg(p.mask + (size_t)b * p.Kmask + kj / 3) > 0.f);
    const uint32_t dthr = DROP ? p.drop.thr : 0u;
    const uint32_t dcol = DROP ? at_drop_key(p, b, h) + (uint32_t)kj : 0u;
    const uint32_t pt_tile = smem_u32(sPT), dst_tile = smem_u32(sDST);
    for (int t = 0; t < n; ++t) {
      mbar_wait(sdp_full, t & 1);
      tc_fence_after();
#pragma unroll
      for (int half = h_lo; half < h_hi; ++half) {
        const int qc = (i0 + t) * 64 + 32 * half;             // first query of the 32-wide chunk
        const bool act = qc + 31 >= k0 + qd * 32 && qc < T;   // some (query, key) pair of this warp's rows is on or below the diagonal
        uint32_t sv[32], dv[32];
        if (act) {
          tmem_ld_32x32(tST + lane_base + 32 * half, sv);
          tmem_ld_32x32(tDPT + lane_base + 32 * half, dv);
          tmem_ld_wait();
        }
        if (half == h_hi - 1) {
          tc_fence_before();
          mbar_arrive(sdp_free);
        }
        if (half == h_lo && t > 0) mbar_wait(pds_done, (t - 1) & 1);  // P^T / dS^T were read by the previous block's dV / dK MMAs
        if (act) {
          // allowed queries of the chunk for this key: kj <= q < T
          const int lo = kj - qc, hi = T - qc;
          uint32_t word = lo <= 0 ? 0xffffffffu : (lo >= 32 ? 0u : ~((1u << lo) - 1u));
          word &= hi >= 32 ? 0xffffffffu : ((1u << hi) - 1u);
          if (!key_ok) word = 0u;
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            float pf[8], df[8];
            const float4 l0 = *reinterpret_cast<const float4*>(sLse + qc + u * 8), l1 = *reinterpret_cast<const float4*>(sLse + qc + u * 8 + 4);
            const float4 d0 = *reinterpret_cast<const float4*>(sDelta + qc + u * 8), d1 = *reinterpret_cast<const float4*>(sDelta + qc + u * 8 + 4);
            const float nl[8] = {l0.x, l0.y, l0.z, l0.w, l1.x, l1.y, l1.z, l1.w}, dl[8] = {d0.x, d0.y, d0.z, d0.w, d1.x, d1.y, d1.z, d1.w};
#pragma unroll
            for (int e = 0; e < 8; ++e) {
              const int col = u * 8 + e;
              float pe = fast_exp2(fmaf(__uint_as_float(sv[col]), p.scale_log2, nl[e]));
              pe = (word & (1u << col)) ? pe : 0.f;
              float g = __uint_as_float(dv[col]);
              bool keep = true;
              if (dthr) {  // bit of (query qc + col, key kj); dV's 1 / (1 - p) is applied once in the epilogue
                keep = drop_keep(dcol, (uint32_t)(qc + col) * (uint32_t)T, dthr);
                g = keep ? g * p.drop.scale : 0.f;
              }
              pf[e] = keep ? pe : 0.f;
              df[e] = pe * (g - dl[e]) * p.scale;
            }
            at_store_row_unit(pt_tile, r, 4 * half + u, pack_bf16x2(pf[0], pf[1]), pack_bf16x2(pf[2], pf[3]), pack_bf16x2(pf[4], pf[5]),
                              pack_bf16x2(pf[6], pf[7]));
            at_store_row_unit(dst_tile, r, 4 * half + u, pack_bf16x2(df[0], df[1]), pack_bf16x2(df[2], df[3]), pack_bf16x2(df[4], df[5]),
                              pack_bf16x2(df[6], df[7]));
          }
        } else {
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            at_store_row_unit(pt_tile, r, 4 * half + u, 0u, 0u, 0u, 0u);
            at_store_row_unit(dst_tile, r, 4 * half + u, 0u, 0u, 0u, 0u);
          }
        }
      }
      fence_proxy_async();
      mbar_arrive(pds_full);
    }
    mbar_wait(pds_done, (n - 1) & 1);
    tc_fence_after();
#pragma unroll
    for (int which = 0; which < 2; ++which) {  // 0: dV -> columns 2H.., 1: dK -> columns H..
#pragma unroll
      for (int half = h_lo; half < Cfg::OC; half += (CW == 8 ? 2 : 1)) {  // 32-column chunks of the head (CW = 8: alternate chunks per warp)
        uint32_t o[32];
        tmem_ld_32x32((which == 0 ? tDV : tDK) + lane_base + 32 * half, o);
        tmem_ld_wait();
        if (which == 0 && dthr) {
#pragma unroll
          for (int e = 0; e < 32; ++e) o[e] = __float_as_uint(__uint_as_float(o[e]) * p.drop.scale);
        }
        if (kj < T) {
          bf16* dst = p.dqkv + ((size_t)b * T + kj) * 3 * H + h * D + 32 * half + (which == 0 ? 2 * H : H);
#pragma unroll
          for (int e = 0; e < 32; e += 8)
            *reinterpret_cast<uint4*>(dst + e) =
                make_uint4(pack_bf16x2(__uint_as_float(o[e]), __uint_as_float(o[e + 1])), pack_bf16x2(__uint_as_float(o[e + 2]), __uint_as_float(o[e + 3])),
                           pack_bf16x2(__uint_as_float(o[e + 4]), __uint_as_float(o[e + 5])), pack_bf16x2(__uint_as_float(o[e + 6]), __uint_as_float(o[e + 7])));
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == CW) { tc_fence_after(); tmem_dealloc(tmem_base, TMEM_COLS); }
}

template <int D> constexpr int at_fwd_smem() { return 1024 + AtCfg<D>::T128 + 2 * AT_KV_STAGES * AtCfg<D>::T64 + 2 * AT_TILE128 + 256 + AT_MAX_CHUNKS * 4; }
template <int D> constexpr int at_dq_smem() {
  return 1024 + 2 * AtCfg<D>::T128 + (AT_DQ_KST + AT_DQ_VST) * AtCfg<D>::T64 + 2 * AT_TILE128 + 256 + AT_MAX_CHUNKS * 4;
}
template <int D> constexpr int at_dkv_smem() { return 1024 + 2 * AtCfg<D>::T128 + 4 * AtCfg<D>::T64 + 2 * AT_TILE128 + 256 + 2 * AT_MAX_T_BWD * 4; }

bool attn_tc_supported(int T, int d) {
  static int legacy = -1;
  if (legacy < 0) {
    const char* e = getenv("DTB_ATTN_LEGACY");
    legacy = (e && e[0] == '1') ? 1 : 0;
  }
  return !legacy && (d == 32 || d == 64 || d == 128) && T >= 1 && T <= AT_MAX_T_BWD;
}

// 64-column boxes of 128 / 64 rows over a [B, T, cols] bf16 tensor (columns beyond `cols` read as zeros: the d = 32 boxes of the last head)
static int at_maps(const bf16* x, int B, int T, int cols, CUtensorMap* tm128, CUtensorMap* tm64) {
  DTB_TRY(make_tmap3(tm128, x, (uint64_t)cols, (uint64_t)T, (uint64_t)B, (uint64_t)cols, (uint64_t)T * cols, 64, 128));
  DTB_TRY(make_tmap3(tm64, x, (uint64_t)cols, (uint64_t)T, (uint64_t)B, (uint64_t)cols, (uint64_t)T * cols, 64, 64));
  return DTB_OK;
}

template <int D>
static int at_fwd_launch(cudaStream_t stream, const CUtensorMap& tmQ, const CUtensorMap& tmKV, const AttnTcParams& p, long long ntiles,
                         bool drop) {
  static bool attr = false;
  if (!attr) {
    DTB_CUDA(cudaFuncSetAttribute(attn_fwd_tc_kernel<D, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, at_fwd_smem<D>()));
    DTB_CUDA(cudaFuncSetAttribute(attn_fwd_tc_kernel<D, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, at_fwd_smem<D>()));
    attr = true;
  }
  // persistent: AtCfg::CTAS CTAs per SM (shared memory / TMEM columns each) walk the tile list; DTB_ATTN_PERSIST=0 -> one tile per CTA
  static int persist = -1;
  if (persist < 0) {
    const char* e = getenv("DTB_ATTN_PERSIST");
    persist = (e && e[0] == '0') ? 0 : 1;
  }
  const long long cap = persist ? (long long)AtCfg<D>::CTAS * num_sms() : ntiles;
  dim3 grid((unsigned)(ntiles < cap ? ntiles : cap));
  if (drop) DTB_LAUNCH((attn_fwd_tc_kernel<D, true>), grid, AT_THREADS, at_fwd_smem<D>(), stream, tmQ, tmKV, p);
  else DTB_LAUNCH((attn_fwd_tc_kernel<D, false>), grid, AT_THREADS, at_fwd_smem<D>(), stream, tmQ, tmKV, p);
  return DTB_OK;
}

int attn_fwd_tc(cudaStream_t stream, const bf16* qkv, const float* mask, int B, int T, int nh, int d, bf16* out, float* lse,
                const DropSite& drop) {
  DTB_REQUIRE(d == 32 || d == 64 || d == 128, DTB_ERR_UNSUPPORTED, "tcgen05 attention: head_dim %d unsupported (32, 64, 128)", d);
  DTB_REQUIRE((nh * d) % 8 == 0, DTB_ERR_BAD_ARG, "tcgen05 attention: hidden size must be a multiple of 8");
  const int H = nh * d;
  CUtensorMap tmQ, tmKV;
  DTB_TRY(at_maps(qkv, B, T, 3 * H, &tmQ, &tmKV));
  AttnTcParams p;
  memset(&p, 0, sizeof(p));
  p.T = T; p.nh = nh; p.Kmask = T / 3;
  p.scale = 1.f / sqrtf((float)d);
  p.scale_log2 = 1.4426950408889634f * p.scale;
  p.mask = mask; p.out = out; p.lse = lse; p.drop = drop;
  p.ntile_b = B;
  const long long ntiles = (long long)((T + 127) / 128) * nh * B;
  DTB_REQUIRE(ntiles < (1LL << 31), DTB_ERR_UNSUPPORTED, "tcgen05 attention: too many tiles");
  ProfScope prof(stream, PF_ATTN_FWD, 2.0 * B * nh * (double)T * T * d);
  const bool dr = drop.state != nullptr;
  if (d == 32) DTB_TRY(at_fwd_launch<32>(stream, tmQ, tmKV, p, ntiles, dr));
  else if (d == 64) DTB_TRY(at_fwd_launch<64>(stream, tmQ, tmKV, p, ntiles, dr));
  else DTB_TRY(at_fwd_launch<128>(stream, tmQ, tmKV, p, ntiles, dr));
  DTB_CHECK_LAUNCH();
  return DTB_OK;
}

template <int D, bool DROP, int CW>
static void at_bwd_launch(dim3 grid, cudaStream_t stream, const CUtensorMap& tm128, const CUtensorMap& tm64, const CUtensorMap& tmDO128,
                          const CUtensorMap& tmDO64, const AttnTcParams& p) {
  static bool attr = false;
  if (!attr) {
    cudaFuncSetAttribute(attn_bwd_dq_tc_kernel<D, DROP, CW>, cudaFuncAttributeMaxDynamicSharedMemorySize, at_dq_smem<D>());
    cudaFuncSetAttribute(attn_bwd_dkv_tc_kernel<D, DROP, CW>, cudaFuncAttributeMaxDynamicSharedMemorySize, at_dkv_smem<D>());
    attr = true;
  }
  DTB_LAUNCH((attn_bwd_dq_tc_kernel<D, DROP, CW>), grid, 32 * (CW + 2), at_dq_smem<D>(), stream, tm128, tm64, tmDO128, p);
  DTB_LAUNCH((attn_bwd_dkv_tc_kernel<D, DROP, CW>), grid, 32 * (CW + 2), at_dkv_smem<D>(), stream, tm128, tm64, tmDO64, p);
}
template <int D>
static void at_bwd_dispatch(bool drop, bool w8, dim3 grid, cudaStream_t stream, const CUtensorMap& tm128, const CUtensorMap& tm64,
                            const CUtensorMap& tmDO128, const CUtensorMap& tmDO64, const AttnTcParams& p) {
  if (drop) { if (w8) at_bwd_launch<D, true, 8>(grid, stream, tm128, tm64, tmDO128, tmDO64, p); else at_bwd_launch<D, true, 4>(grid, stream, tm128, tm64, tmDO128, tmDO64, p); }
  else      { if (w8) at_bwd_launch<D, false, 8>(grid, stream, tm128, tm64, tmDO128, tmDO64, p); else at_bwd_launch<D, false, 4>(grid, stream, tm128, tm64, tmDO128, tmDO64, p); }
}
// compute warps of the backward kernels: 4 (each thread walks both 32-column halves) or 8 (one half per warp); env DTB_ATTN_BWD_WARPS
static int at_bwd_warps(bool dropout) {
  static int forced = -1;
  if (forced < 0) {
    const char* e = getenv("DTB_ATTN_BWD_WARPS");
    forced = e ? atoi(e) : 0;
  }
  if (forced == 4 || forced == 8) return forced;
  return dropout ? 8 : 4;
}

int attn_bwd_tc(cudaStream_t stream, const bf16* qkv, const float* mask, const bf16* out, const bf16* dout, const float* lse, int B, int T,
                int nh, int d, bf16* dqkv, float* delta, const DropSite& drop) {
  DTB_REQUIRE((d == 32 || d == 64 || d == 128) && T <= AT_MAX_T_BWD, DTB_ERR_UNSUPPORTED, "tcgen05 attention bwd: head_dim %d / T %d unsupported", d, T);
  const int H = nh * d;
  CUtensorMap tm128, tm64, tmDO128, tmDO64;
  DTB_TRY(at_maps(qkv, B, T, 3 * H, &tm128, &tm64));
  DTB_TRY(at_maps(dout, B, T, H, &tmDO128, &tmDO64));
  AttnTcParams p;
  memset(&p, 0, sizeof(p));
  p.T = T; p.nh = nh; p.Kmask = T / 3;
  p.scale = 1.f / sqrtf((float)d);
  p.scale_log2 = 1.4426950408889634f * p.scale;
  p.mask = mask; p.lse_in = lse; p.out_in = out; p.dout = dout; p.dqkv = dqkv; p.delta = delta; p.drop = drop;
  dim3 grid((T + 127) / 128, nh, B);
  ProfScope prof(stream, PF_ATTN_BWD, 5.0 * B * nh * (double)T * T * d);
  const bool dr = drop.state != nullptr, w8 = at_bwd_warps(dr) == 8;
  if (d == 32) at_bwd_dispatch<32>(dr, w8, grid, stream, tm128, tm64, tmDO128, tmDO64, p);
  else if (d == 64) at_bwd_dispatch<64>(dr, w8, grid, stream, tm128, tm64, tmDO128, tmDO64, p);
  else at_bwd_dispatch<128>(dr, w8, grid, stream, tm128, tm64, tmDO128, tmDO64, p);
  DTB_CHECK_LAUNCH();
  return DTB_OK;
}

}  // namespace dtb
