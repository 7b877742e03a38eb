// Host side of the tcgen05 GEMM: TMA descriptor encoding (driver entry point fetched through
// cudart so the library has no link-time dependency on libcuda), tile/split heuristics, launch.
#include "gemm_sm100.cuh"
#include "kernels.h"

#include <cudaTypedefs.h>
#include <mutex>
#include <unordered_map>
#include <string.h>
#include <stdlib.h>

namespace dtb {

static PFN_cuTensorMapEncodeTiled_v12000 g_encode = nullptr;

static int get_encode_fn() {
  if (g_encode) return DTB_OK;
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult qres;
  DTB_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
  DTB_REQUIRE(fn != nullptr && qres == cudaDriverEntryPointSuccess, DTB_ERR_CUDA, "cuTensorMapEncodeTiled entry point unavailable");
  g_encode = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(fn);
  return DTB_OK;
}

struct TmKey {
  const void* ptr;
  uint64_t d0, d1, stride;
  uint32_t b0, b1;
  bool operator==(const TmKey& o) const {
    return ptr == o.ptr && d0 == o.d0 && d1 == o.d1 && stride == o.stride && b0 == o.b0 && b1 == o.b1;
  }
};
struct TmKeyHash {
  size_t operator()(const TmKey& k) const {
    size_t h = reinterpret_cast<size_t>(k.ptr);
    h ^= k.d0 * 0x9E3779B97F4A7C15ull + (h << 6) + (h >> 2);
    h ^= k.d1 * 0xC2B2AE3D27D4EB4Full + (h << 6) + (h >> 2);
    h ^= k.stride * 0x165667B19E3779F9ull + (h << 6) + (h >> 2);
    h ^= ((size_t)k.b0 << 32 | k.b1) + (h << 6) + (h >> 2);
    return h;
  }
};
static std::unordered_map<TmKey, CUtensorMap, TmKeyHash> g_tm_cache;
static std::mutex g_tm_mutex;

// 2-D bf16 tensor map: dims {d0 (contiguous), d1}, row stride in elements, box {b0, b1}, 128B swizzle.
int make_tmap(CUtensorMap* out, const void* ptr, uint64_t d0, uint64_t d1, uint64_t stride_elems, uint32_t b0, uint32_t b1) {
  DTB_TRY(get_encode_fn());
  DTB_REQUIRE((reinterpret_cast<uintptr_t>(ptr) & 15) == 0, DTB_ERR_BAD_ARG, "GEMM operand pointer must be 16-byte aligned");
  DTB_REQUIRE((stride_elems * 2) % 16 == 0, DTB_ERR_BAD_ARG, "GEMM operand row stride must be a multiple of 8 elements (got %llu)",
              (unsigned long long)stride_elems);
  TmKey key{ptr, d0, d1, stride_elems, b0, b1};
  {
    std::lock_guard<std::mutex> g(g_tm_mutex);
    auto it = g_tm_cache.find(key);
    if (it != g_tm_cache.end()) {
      *out = it->second;
      return DTB_OK;
    }
  }
  cuuint64_t gdim[2] = {d0, d1};
  cuuint64_t gstride[1] = {stride_elems * 2};
  cuuint32_t box[2] = {b0, b1};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = g_encode(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), gdim, gstride, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  DTB_REQUIRE(r == CUDA_SUCCESS, DTB_ERR_CUDA, "cuTensorMapEncodeTiled failed (%d) dims=(%llu,%llu) stride=%llu box=(%u,%u)", (int)r,
              (unsigned long long)d0, (unsigned long long)d1, (unsigned long long)stride_elems, b0, b1);
  {
    std::lock_guard<std::mutex> g(g_tm_mutex);
    if (g_tm_cache.size() > 4096) g_tm_cache.clear();
    g_tm_cache.emplace(key, *out);
  }
  return DTB_OK;
}

// 3-D bf16 tensor map (uncached: callers keep it for the launch): dims {d0 contiguous, d1, d2}, strides of d1 / d2 in elements,
// box {b0, b1, 1}, 128B swizzle, out-of-bounds elements read as zero.
int make_tmap3(CUtensorMap* out, const void* ptr, uint64_t d0, uint64_t d1, uint64_t d2, uint64_t stride1, uint64_t stride2, uint32_t b0,
               uint32_t b1) {
  DTB_TRY(get_encode_fn());
  DTB_REQUIRE((reinterpret_cast<uintptr_t>(ptr) & 15) == 0 && (stride1 * 2) % 16 == 0 && (stride2 * 2) % 16 == 0, DTB_ERR_BAD_ARG,
              "tensor map: pointer / strides must be 16-byte aligned");
  cuuint64_t gdim[3] = {d0, d1, d2};
  cuuint64_t gstride[2] = {stride1 * 2, stride2 * 2};
  cuuint32_t box[3] = {b0, b1, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = g_encode(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(ptr), gdim, gstride, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  DTB_REQUIRE(r == CUDA_SUCCESS, DTB_ERR_CUDA, "cuTensorMapEncodeTiled(3d) failed (%d)", (int)r);
  return DTB_OK;
}

static bool g_attr_set = false;

// Ticket counters of the in-kernel split-K reduction (gemm_splitk_finish).  Zero at load, and every kernel leaves the counters it used
// at zero.  Launches take consecutive slices of the ring, so GEMMs that run concurrently on different streams (data-gradient chain,
// weight-gradient side streams of the trunk and of both towers) never share a counter; a slice is reused only SK_RING / (counters per
// launch) launches later.
constexpr unsigned int SK_RING = 1u << 18;
__device__ unsigned int g_sk_counters[SK_RING];
static unsigned int* g_sk_base = nullptr;
static unsigned int g_sk_cursor = 0;
static std::mutex g_sk_mutex;

int ticket_counters_take(unsigned int need, unsigned int** out) {
  *out = nullptr;
  if (need == 0 || need > SK_RING / 8) return DTB_OK;  // too many tiles: the caller falls back to the separate reduction kernel
  std::lock_guard<std::mutex> g(g_sk_mutex);
  if (!g_sk_base) DTB_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&g_sk_base), g_sk_counters));
  if (g_sk_cursor + need > SK_RING) g_sk_cursor = 0;
  *out = g_sk_base + g_sk_cursor;
  g_sk_cursor += need;
  return DTB_OK;
}
// Off by default: the slices of a tile are then summed by the few warps that own it (m_tiles x n_tiles x 4), a long chain of dependent
// L2 round trips at the small-M/N, long-K weight-gradient shapes of c1-c3 (measured: c2 step 1.16 -> 2.0 ms), whereas the separate
// splitk_reduce launch spreads the same sums over hundreds of blocks.  dtb_set_option("splitk_inkernel", 1) turns it on.
static bool sk_inkernel_enabled() { return opt_splitk_inkernel() != 0; }

size_t gemm_splitk_workspace_bytes(int M, int N, int K) {
  // upper bound used by callers that pre-size the workspace: split <= 32
  (void)K;
  return (size_t)32 * M * N * sizeof(float);
}

// CTA-pair kernel (256 x 256 tiles, cta_group::2) for problems big enough to fill the 74 SM pairs.
bool gemm_use_pair(int M, int N, int K) {
  static int off = -1;
  if (off < 0) {
    const char* e = getenv("DTB_GEMM_NO_PAIR");
    off = (e && e[0] == '1') ? 1 : 0;
  }
  if (off) return false;
  if (M < 512 || N < 256 || N % 128 != 0 || K < 128) return false;
  const long long tiles = (long long)((M + 255) / 256) * ((N + 255) / 256);
  const long long kblocks = (K + GEMM_BK - 1) / GEMM_BK;
  // enough (tile x possible split) work for the 74 pairs
  return tiles * (kblocks >= 16 ? kblocks / 8 : 1) >= (num_sms() / 2);
}

// Split-K factor by wave efficiency: work items (tiles x split) should fill whole waves of the `units` concurrent CTAs / pairs.
static int pick_split_eff(long long tiles, int kblocks, int units) {
  int max_split = kblocks / 8;
  if (max_split < 1) max_split = 1;
  if (max_split > 32) max_split = 32;
  int best = 1;
  double best_eff = 0.0;
  for (int s = 1; s <= max_split; ++s) {
    const long long items = tiles * s;
    const long long waves = (items + units - 1) / units;
    const double eff = (double)items / (double)(waves * units);
    if (eff > best_eff + 0.02) { best_eff = eff; best = s; }  // prefer the smaller split unless clearly better
    if (best_eff >= 0.95) break;
  }
  return best;
}

int gemm_pick_split(int M, int N, int K, int BN) {
  const int kblocks = (K + GEMM_BK - 1) / GEMM_BK;
  if (gemm_use_pair(M, N, K)) return pick_split_eff((long long)((M + 255) / 256) * ((N + 255) / 256), kblocks, num_sms() / 2);
  const int m_tiles = (M + GEMM_BM - 1) / GEMM_BM, n_tiles = (N + BN - 1) / BN;
  return pick_split_eff((long long)m_tiles * n_tiles, kblocks, num_sms());
}

int gemm_pick_bn(int M, int N, int b_mn_major) {
  const int m_tiles = (M + GEMM_BM - 1) / GEMM_BM;
  const int cands[4] = {256, 128, 64, 32};
  int best = -1;
  for (int i = 0; i < 4; ++i) {
    int bn = cands[i];
    if (b_mn_major && bn % 64) continue;
    if (N % bn) continue;
    if (best < 0) best = bn;  // largest divisor candidate
    if (m_tiles * (N / bn) >= num_sms()) return bn;
    best = bn;  // keep shrinking while the grid is under-filled
    if (bn == 64) break;  // do not go below 64 just to add CTAs
  }
  if (best > 0) return best;
  // N not divisible by any candidate: predicate the tail
  return b_mn_major ? 64 : (N >= 64 ? 64 : 32);
}

int gemm_bf16(cudaStream_t stream, const bf16* A, int lda, int a_mn_major, const bf16* B, int ldb, int b_mn_major, int M, int N, int K,
              const GemmEpilogue& ep_in, int split_k, int bn_override) {
  DTB_REQUIRE(M > 0 && N > 0 && K > 0, DTB_ERR_BAD_ARG, "gemm: empty problem %d x %d x %d", M, N, K);
  DTB_REQUIRE(N % 8 == 0, DTB_ERR_BAD_ARG, "gemm: N (%d) must be a multiple of 8", N);
  GemmParams p;
  memset(&p, 0, sizeof(p));
  p.M = M; p.N = N; p.K = K;
  p.a_mn_major = a_mn_major; p.b_mn_major = b_mn_major;
  // bn_override: 0 = automatic; 32..256 = single-CTA kernel with that N tile; 1000 = force the CTA-pair kernel
  const bool pair = bn_override == 1000 || (bn_override == 0 && gemm_use_pair(M, N, K));
  p.BN = pair ? GEMM2_BN : (bn_override > 0 ? bn_override : gemm_pick_bn(M, N, b_mn_major));
  DTB_REQUIRE(p.BN % 32 == 0 && p.BN >= 32 && p.BN <= 256 && (!b_mn_major || p.BN % 64 == 0), DTB_ERR_BAD_ARG, "gemm: bad BN %d", p.BN);
  p.m_tiles = pair ? (M + 255) / 256 : (M + GEMM_BM - 1) / GEMM_BM;
  p.n_tiles = (N + p.BN - 1) / p.BN;
  const int kblocks = (K + GEMM_BK - 1) / GEMM_BK;
  if (split_k < 1) split_k = 1;
  if (split_k > kblocks) split_k = kblocks;
  p.kb_per_split = (kblocks + split_k - 1) / split_k;
  p.split_k = (kblocks + p.kb_per_split - 1) / p.kb_per_split;  // no empty splits
  p.ep = ep_in;
  if (p.ep.alpha == 0.f) p.ep.alpha = 1.f;
  if (p.split_k > 1) DTB_REQUIRE(p.ep.partial != nullptr, DTB_ERR_BAD_ARG, "gemm: split-K needs a partial workspace");
  if (p.ep.ln_out) {
    DTB_REQUIRE(!pair && p.BN == N && N % 32 == 0 && N <= GEMM_MAX_BN && p.split_k == 1 && !p.ep.partial && p.ep.ln_gamma && p.ep.ln_beta,
                DTB_ERR_BAD_ARG, "gemm: fused LayerNorm needs the single-CTA kernel with BN == N (N %% 32 == 0, <= 256) and no split-K");
    DTB_REQUIRE(p.ep.ld_ln % 8 == 0 && (reinterpret_cast<uintptr_t>(p.ep.ln_out) & 15) == 0, DTB_ERR_BAD_ARG, "gemm: ln_out must be 16-byte aligned rows");
  }
  if (p.ep.sk_out) {  // in-kernel reduction requested: one counter per (tile, epilogue warp)
    p.ep.sk_counters = nullptr;
    if (p.ep.partial && p.split_k > 1 && sk_inkernel_enabled() && p.ep.sk_ldo % 4 == 0 && (reinterpret_cast<uintptr_t>(p.ep.sk_out) & 15) == 0) {
      const unsigned int need = (unsigned int)(p.m_tiles * p.n_tiles) * (pair ? 2u * GEMM2_EPI_WARPS : 4u);
      DTB_TRY(ticket_counters_take(need, &p.ep.sk_counters));
    }
    if (!p.ep.sk_counters) p.ep.sk_out = nullptr;
  }

  const uint32_t b_rows = pair ? 128u : (uint32_t)p.BN;  // a pair CTA stages its own 128-column half of B
  CUtensorMap tmA, tmB;
  if (!a_mn_major) DTB_TRY(make_tmap(&tmA, A, (uint64_t)K, (uint64_t)M, (uint64_t)lda, GEMM_BK, GEMM_BM));
  else             DTB_TRY(make_tmap(&tmA, A, (uint64_t)M, (uint64_t)K, (uint64_t)lda, 64, GEMM_BK));
  if (!b_mn_major) DTB_TRY(make_tmap(&tmB, B, (uint64_t)K, (uint64_t)N, (uint64_t)ldb, GEMM_BK, b_rows));
  else             DTB_TRY(make_tmap(&tmB, B, (uint64_t)N, (uint64_t)K, (uint64_t)ldb, 64, GEMM_BK));

  if (!g_attr_set) {
    DTB_CUDA(cudaFuncSetAttribute(gemm_tcgen05_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, GEMM_SMEM_BYTES));
    DTB_CUDA(cudaFuncSetAttribute(gemm_tcgen05_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, GEMM2_SMEM_BYTES));
    g_attr_set = true;
  }
  const int total = p.m_tiles * p.n_tiles * p.split_k;
  // roofline family: a shape whose arithmetic intensity is below the machine balance (sustained bf16 peak / copy bandwidth of
  // MEASURED_PEAKS.json, handed over by bench.py through dtb_profile_set_balance) is HBM-bound; its work is counted in algorithmic
  // bytes (operands once + every epilogue stream once)
  const double flops = 2.0 * M * N * K, mn = (double)M * N;
  const double bytes = 2.0 * ((double)M * K + (double)K * N) +
                       (p.ep.partial ? 4.0 * mn * p.split_k
                                   : mn * ((p.ep.out_bf16 ? 2.0 : 0.0) + (p.ep.out_f32 ? 4.0 : 0.0) + (p.ep.out_preact ? 2.0 : 0.0) +
                                           (p.ep.residual ? 4.0 : 0.0) + (p.ep.dact ? 2.0 : 0.0)));
  const bool mem_bound = flops < opt_balance() * bytes;
  ProfScope prof(stream, mem_bound ? PF_GEMM_MEM : PF_GEMM, mem_bound ? bytes : flops);
  if (pair) {
    const int npairs = total < num_sms() / 2 ? total : num_sms() / 2;
    DTB_LAUNCH(gemm_tcgen05_pair_kernel, 2 * npairs, GEMM2_THREADS, GEMM2_SMEM_BYTES, stream, tmA, tmB, p);
  } else {
    const int grid = total < num_sms() ? total : num_sms();
    DTB_LAUNCH(gemm_tcgen05_kernel, grid, GEMM_THREADS, GEMM_SMEM_BYTES, stream, tmA, tmB, p);
  }
  DTB_CHECK_LAUNCH();
  // > 0: number of partial slices the caller still has to reduce; 0: split-K partials were reduced inside the kernel into ep.sk_out
  return (p.ep.partial && p.ep.sk_counters) ? 0 : p.split_k;
}

// ---- split-K reduction: out[M, ldo] (+)= scale * sum_s partial[s][M][N], slices added in index order (deterministic) -------------
// 32 float4 columns x 8 split-lanes per block: lane l sums slices l, l+8, ... and the 8 lanes are combined in index order.
__global__ void __launch_bounds__(256) splitk_reduce_kernel(const float* __restrict__ partial, int splits, int M, int N,
                                                            float* __restrict__ out, int ldo, int accumulate, float scale) {
  pdl_trigger();
  pdl_wait();
  __shared__ float4 red[8][32];
  const size_t n4 = (size_t)M * N / 4;
  for (size_t base = (size_t)blockIdx.x * 32; base < n4; base += (size_t)gridDim.x * 32) {
    const size_t i = base + threadIdx.x;
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    if (i < n4)
      for (int k = threadIdx.y; k < splits; k += 8) {
        const float4 v = ldg_f4(partial + (size_t)k * M * N + i * 4);
        s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
      }
    red[threadIdx.y][threadIdx.x] = s;
    __syncthreads();
    if (threadIdx.y == 0 && i < n4) {
      for (int l = 1; l < 8; ++l) {
        const float4 v = red[l][threadIdx.x];
        s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
      }
      const size_t e = i * 4;
      float* o = out + (e / N) * (size_t)ldo + (e % N);
      const float4 prev = accumulate ? *reinterpret_cast<float4*>(o) : make_float4(0.f, 0.f, 0.f, 0.f);
      *reinterpret_cast<float4*>(o) = make_float4(prev.x + scale * s.x, prev.y + scale * s.y, prev.z + scale * s.z, prev.w + scale * s.w);
    }
    __syncthreads();
  }
}

int splitk_reduce(cudaStream_t stream, const float* partial, int splits, int M, int N, float* out, int ldo, int accumulate, float scale) {
  DTB_REQUIRE(N % 4 == 0 && ldo % 4 == 0, DTB_ERR_BAD_ARG, "splitk_reduce: N and ldo must be multiples of 4");
  const size_t n4 = (size_t)M * N / 4;
  size_t blocks = (n4 + 31) / 32;
  if (blocks > (size_t)num_sms() * 8) blocks = (size_t)num_sms() * 8;
  if (blocks < 1) blocks = 1;
  DTB_LAUNCH(splitk_reduce_kernel, (unsigned)blocks, dim3(32, 8), 0, stream, partial, splits, M, N, out, ldo, accumulate, scale);
  DTB_CHECK_LAUNCH();
  return DTB_OK;
}

// Weight gradient dW[Mout, Nout] = A^T B with both operands read in place (MN-major): A memory [Kred][Mout] (lda),
// B memory [Kred][Nout] (ldb).  Split-K over the token dimension with fp32 partials reduced in index order (deterministic).
int gemm_wgrad_split(int Mout, int Nout, int Kred, size_t partial_floats) {
  const int bn = gemm_use_pair(Mout, Nout, Kred) ? GEMM2_BN : gemm_pick_bn(Mout, Nout, 1);
  int split = gemm_pick_split(Mout, Nout, Kred, bn);
  const size_t per = (size_t)Mout * Nout;
  while (split > 1 && per * split > partial_floats) --split;
  return split;
}

int gemm_wgrad(cudaStream_t stream, const bf16* A, int lda, const bf16* B, int ldb, int Mout, int Nout, int Kred, float* out, int ldo,
               float* partial, size_t partial_floats) {
  const int split = gemm_wgrad_split(Mout, Nout, Kred, partial_floats);
  GemmEpilogue ep;
  memset(&ep, 0, sizeof(ep));
  ep.alpha = 1.f;
  if (split <= 1) {
    ep.out_f32 = out;
    ep.ldo_f = ldo;
    const int rc = gemm_bf16(stream, A, lda, 1, B, ldb, 1, Mout, Nout, Kred, ep, 1, 0);
    return rc < 0 ? rc : DTB_OK;
  }
  ep.partial = partial;
  ep.sk_out = out;  // reduced by the last-arriving CTA of every tile inside the GEMM (no separate reduction launch)
  ep.sk_ldo = ldo;
  const int rc = gemm_bf16(stream, A, lda, 1, B, ldb, 1, Mout, Nout, Kred, ep, split, 0);
  if (rc <= 0) return rc;
  return splitk_reduce(stream, partial, rc, Mout, Nout, out, ldo, 0, 1.f);
}

}  // namespace dtb
