// HBM-bound helpers: fp32 -> bf16 weight casts (with transpose), bias-gradient column sums.
#include "kernels.h"
#include <string.h>

namespace dtb {

// ---- cast + transpose: src fp32 [R, C] -> dst bf16 [R, C], dst_t bf16 [C, R] -------------------
// 32x32 tiles through shared memory so both the read and the two writes are coalesced.
__global__ void cast_transpose_kernel(const float* __restrict__ src, int R, int C, bf16* __restrict__ dst, bf16* __restrict__ dst_t) {
  pdl_trigger();
  pdl_wait();
  __shared__ float tile[32][33];
  const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int r = r0 + i, c = c0 + threadIdx.x;
    float v = (r < R && c < C) ? src[(size_t)r * C + c] : 0.f;
    tile[i][threadIdx.x] = v;
    if (dst && r < R && c < C) dst[(size_t)r * C + c] = __float2bfloat16(v);
  }
  __syncthreads();
  if (dst_t) {
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
      const int c = c0 + i, r = r0 + threadIdx.x;
      if (r < R && c < C) dst_t[(size_t)c * R + r] = __float2bfloat16(tile[threadIdx.x][i]);
    }
  }
}

int cast_transpose(cudaStream_t stream, const float* src, int R, int C, bf16* dst, bf16* dst_t) {
  dim3 grid((C + 31) / 32, (R + 31) / 32), block(32, 8);
  DTB_LAUNCH(cast_transpose_kernel, grid, block, 0, stream, src, R, C, dst, dst_t);
  DTB_CHECK_LAUNCH();
  return DTB_OK;
}

__global__ void cast_bf16_kernel(const float* __restrict__ src, size_t n, bf16* __restrict__ dst) {
  pdl_trigger();
  pdl_wait();
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  const size_t n4 = n / 4;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n4; i += stride) {
    float4 v = ldg_f4(src + 4 * i);
    *reinterpret_cast<uint2*>(dst + 4 * i) = make_uint2(pack_bf16x2(v.x, v.y), pack_bf16x2(v.z, v.w));
  }
  for (size_t i = n4 * 4 + blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += stride) dst[i] = __float2bfloat16(src[i]);
}

int cast_bf16(cudaStream_t stream, const float* src, size_t n, bf16* dst) {
  if (n == 0) return DTB_OK;
  DTB_REQUIRE((reinterpret_cast<uintptr_t>(src) & 15) == 0 && (reinterpret_cast<uintptr_t>(dst) & 7) == 0, DTB_ERR_BAD_ARG,
              "cast_bf16: misaligned pointers");
  int blocks = (int)((n / 4 + 255) / 256);
  if (blocks > num_sms() * 8) blocks = num_sms() * 8;
  if (blocks < 1) blocks = 1;
  DTB_LAUNCH(cast_bf16_kernel, blocks, 256, 0, stream, src, n, dst);
  DTB_CHECK_LAUNCH();
  return DTB_OK;
}

// ---- column sums (bias gradients): x bf16 [M, N] -> out fp32 [N] --------------------------------
// Stage 1: grid (column groups, row slabs); each thread owns 8 adjacent columns (one 128-bit load per row) and strides rows;
//          the row-lanes of the block are reduced through shared memory in index order -> scratch[slab][N].
// Stage 2: 32 columns x 8 row-lanes per block sum the slabs, lanes combined in index order (deterministic).
constexpr int CS_SLABS_MAX = 1024;        // narrow matrices (conv bias grads, N = 32) need many row slabs to fill the GPU
constexpr int CS_SCRATCH_MIN = 64 * 1024;  // floats

__global__ void __launch_bounds__(256) colsum_stage1(const bf16* __restrict__ x, int M, int N, int ld, float* __restrict__ scratch) {
  pdl_trigger();
  pdl_wait();
  __shared__ float red[256][9];
  const int cx = blockDim.x, ry = blockDim.y;
  const int col = (blockIdx.x * cx + threadIdx.x) * 8;
  const int slab = blockIdx.y, nslab = gridDim.y;
  float acc[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) acc[e] = 0.f;
  if (col < N) {
    for (int r = slab * ry + threadIdx.y; r < M; r += nslab * ry) {
      const uint4 u = ldg_u4(x + (size_t)r * ld + col);
      const float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), c = unpack_bf16x2(u.z), d = unpack_bf16x2(u.w);
      acc[0] += a.x; acc[1] += a.y; acc[2] += b.x; acc[3] += b.y; acc[4] += c.x; acc[5] += c.y; acc[6] += d.x; acc[7] += d.y;
    }
  }
  const int tid = threadIdx.y * cx + threadIdx.x;
#pragma unroll
  for (int e = 0; e < 8; ++e) red[tid][e] = acc[e];
  __syncthreads();
  if (threadIdx.y == 0 && col < N) {
    for (int i = 1; i < ry; ++i)
#pragma unroll
      for (int e = 0; e < 8; ++e) acc[e] += red[i * cx + threadIdx.x][e];
    float* dst = scratch + (size_t)slab * N + col;
    *reinterpret_cast<float4*>(dst) = make_float4(acc[0], acc[1], acc[2], acc[3]);
    *reinterpret_cast<float4*>(dst + 4) = make_float4(acc[4], acc[5], acc[6], acc[7]);
  }
}

__global__ void __launch_bounds__(256) colsum_stage2(const float* __restrict__ scratch, int nslab, int N, float* __restrict__ out,
                                                     int accumulate) {
  pdl_trigger();
  pdl_wait();
  __shared__ float red[8][33];
  const int c = blockIdx.x * 32 + threadIdx.x;
  float s = 0.f;
  if (c < N)
    for (int i = threadIdx.y; i < nslab; i += 8) s += scratch[(size_t)i * N + c];
  red[threadIdx.y][threadIdx.x] = s;
  __syncthreads();
  if (threadIdx.y == 0 && c < N) {
    for (int i = 1; i < 8; ++i) s += red[i][threadIdx.x];
    out[c] = accumulate ? out[c] + s : s;
  }
}

size_t colsum_scratch_floats(int N) { return (size_t)64 * N > (size_t)CS_SCRATCH_MIN ? (size_t)64 * N : (size_t)CS_SCRATCH_MIN; }

int colsum_bf16(cudaStream_t stream, const bf16* x, int M, int N, int ld, float* out, int accumulate, float* scratch) {
  DTB_REQUIRE(N % 8 == 0 && ld % 8 == 0 && (reinterpret_cast<uintptr_t>(x) & 15) == 0, DTB_ERR_BAD_ARG,
              "colsum: N and ld must be multiples of 8 and x 16-byte aligned (N=%d ld=%d)", N, ld);
  const int n8 = N / 8;
  int cx = 4;
  while (cx < 32 && cx < n8) cx *= 2;
  const int ry = 256 / cx;
  const int colgroups = (n8 + cx - 1) / cx;
  int nslab = (4 * num_sms() + colgroups - 1) / colgroups;
  const int max_slab = (M + 4 * ry - 1) / (4 * ry);
  if (nslab > max_slab) nslab = max_slab;
  if (nslab > CS_SLABS_MAX) nslab = CS_SLABS_MAX;
  if ((size_t)nslab * N > colsum_scratch_floats(N)) nslab = (int)(colsum_scratch_floats(N) / N);
  if (nslab < 1) nslab = 1;
  ProfScope prof(stream, PF_COLSUM, 2.0 * M * N);
  DTB_LAUNCH(colsum_stage1, dim3(colgroups, nslab), dim3(cx, ry), 0, stream, x, M, N, ld, scratch);
  DTB_CHECK_LAUNCH();
  DTB_LAUNCH(colsum_stage2, (N + 31) / 32, dim3(32, 8), 0, stream, scratch, nslab, N, out, accumulate);
  DTB_CHECK_LAUNCH();
  return DTB_OK;
}

// ---- allocation-free column sum (C ABI dtb_colsum_bf16_static; no counterpart in the reference) ----------------------------------
// Same two kernels as colsum_bf16, with the slab partials in a file-scope device array instead of caller scratch, for callers that
// cannot allocate (graph capture, serving loops).  The array costs CS_STATIC_FLOATS * 4 = 256 KB on every GPU that loads the library.
// Its host-side name is only a handle: the device address is looked up once per device with cudaGetSymbolAddress (that device
// current) and cached.  CS_STATIC_FLOATS == CS_SCRATCH_MIN == colsum_scratch_floats(N) for every N <= CS_STATIC_MAX_N, so the slab
// count — and therefore every bit of the result — equals dtb_colsum_bf16's.  Stage 1 overwrites every partial stage 2 reads: nothing
// has to start at zero.  All calls share the array: calls that overlap on the GPU (different streams, no ordering) race.
constexpr int CS_STATIC_FLOATS = CS_SCRATCH_MIN;
constexpr int CS_STATIC_MAX_N = CS_STATIC_FLOATS / 64;  // 1024 columns
constexpr int CS_STATIC_MAX_DEVICES = 64;
static __device__ __align__(256) float g_colsum_static[CS_STATIC_FLOATS];
static float* g_colsum_static_ptr[CS_STATIC_MAX_DEVICES] = {};

static_assert(CS_STATIC_MAX_N == DTB_COLSUM_STATIC_MAX_N, "include/dtb200.h states the limit");

int colsum_bf16_static(cudaStream_t stream, const bf16* x, int M, int N, int ld, float* out, int accumulate) {
  DTB_REQUIRE(M >= 1 && N >= 8 && N % 8 == 0 && ld % 8 == 0 && ld >= N && (reinterpret_cast<uintptr_t>(x) & 15) == 0, DTB_ERR_BAD_ARG,
              "colsum_static: M >= 1, N and ld multiples of 8, ld >= N and x 16-byte aligned required (M=%d N=%d ld=%d)", M, N, ld);
  DTB_REQUIRE(N <= CS_STATIC_MAX_N, DTB_ERR_BAD_ARG,
              "colsum_static: N=%d exceeds the static scratch limit of %d columns (use dtb_colsum_bf16 with caller scratch)", N,
              CS_STATIC_MAX_N);
  int dev = -1;
  DTB_CUDA(cudaGetDevice(&dev));
  DTB_REQUIRE(dev >= 0 && dev < CS_STATIC_MAX_DEVICES, DTB_ERR_BAD_ARG, "colsum_static: device ordinal %d out of range", dev);
  float* scratch = g_colsum_static_ptr[dev];
  if (!scratch) {  // set-up, once per device, with that device current
    DTB_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&scratch), g_colsum_static));
    DTB_REQUIRE(scratch != nullptr, DTB_ERR_CUDA, "colsum_static: cudaGetSymbolAddress returned a null device address");
    g_colsum_static_ptr[dev] = scratch;
  }
  const int n8 = N / 8;
  int cx = 4;
  while (cx < 32 && cx < n8) cx *= 2;
  const int ry = 256 / cx;
  const int colgroups = (n8 + cx - 1) / cx;
  int nslab = (4 * num_sms() + colgroups - 1) / colgroups;
  const int max_slab = (M + 4 * ry - 1) / (4 * ry);
  if (nslab > max_slab) nslab = max_slab;
  if (nslab > CS_SLABS_MAX) nslab = CS_SLABS_MAX;
  if ((size_t)nslab * N > (size_t)CS_STATIC_FLOATS) nslab = CS_STATIC_FLOATS / N;
  if (nslab < 1) nslab = 1;
  DTB_LAUNCH(colsum_stage1, dim3(colgroups, nslab), dim3(cx, ry), 0, stream, x, M, N, ld, scratch);
  DTB_CHECK_LAUNCH();
  DTB_LAUNCH(colsum_stage2, (N + 31) / 32, dim3(32, 8), 0, stream, (const float*)scratch, nslab, N, out, accumulate);
  DTB_CHECK_LAUNCH();
  return DTB_OK;
}

// ---- get_action history packing: keep the t most recent rows, LEFT-pad with zeros to K rows, mask = [0]*(K-t)+[1]*t ----
// One launch for all per-timestep tensors of the call (states / frames / audio / actions / rtg / timesteps): segment s has
// row_words[s] 32-bit words per timestep; blockIdx.y = segment.  Padding and indexing are integer-exact copies.
struct PadArgs {
  const uint32_t* src[8];
  uint32_t* dst[8];
  int row_words[8];
  int t, K;
  float* mask;
};
__global__ void __launch_bounds__(256) pad_left_kernel(const PadArgs a) {
  pdl_trigger();
  pdl_wait();
  const int s = blockIdx.y;
  const int rw = a.row_words[s];
  const long long total = (long long)a.K * rw, shift = (long long)(a.K - a.t) * rw;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x)
    a.dst[s][i] = i >= shift ? a.src[s][i - shift] : 0u;
  if (s == 0 && blockIdx.x == 0 && a.mask)
    for (int k = threadIdx.x; k < a.K; k += blockDim.x) a.mask[k] = k >= a.K - a.t ? 1.f : 0.f;
}

int pad_left(cudaStream_t stream, int nseg, const void* const* srcs, void* const* dsts, const int* row_bytes, int t, int K, float* mask) {
  DTB_REQUIRE(nseg >= 1 && nseg <= 8 && t >= 0 && t <= K && K >= 1, DTB_ERR_BAD_ARG, "pad_left: bad arguments (nseg=%d t=%d K=%d)", nseg, t, K);
  PadArgs a;
  memset(&a, 0, sizeof(a));
  long long maxw = 1;
  for (int s = 0; s < nseg; ++s) {
    DTB_REQUIRE(row_bytes[s] > 0 && row_bytes[s] % 4 == 0 && dsts[s] && (srcs[s] || t == 0), DTB_ERR_BAD_ARG,
                "pad_left: segment %d needs 4-byte-multiple rows and non-null pointers", s);
    a.src[s] = reinterpret_cast<const uint32_t*>(srcs[s]);
    a.dst[s] = reinterpret_cast<uint32_t*>(dsts[s]);
    a.row_words[s] = row_bytes[s] / 4;
    if ((long long)K * a.row_words[s] > maxw) maxw = (long long)K * a.row_words[s];
  }
  a.t = t; a.K = K; a.mask = mask;
  long long blocks = (maxw + 255) / 256;
  if (blocks > 64) blocks = 64;
  DTB_LAUNCH(pad_left_kernel, dim3((unsigned)blocks, nseg), 256, 0, stream, a);
  DTB_CHECK_LAUNCH();
  return DTB_OK;
}

// ---- dropout plumbing ------------------------------------------------------------------------------------------------------
__global__ void dropout_advance_kernel(uint32_t* state) {
  pdl_trigger();
  pdl_wait();
  if (threadIdx.x == 0 && blockIdx.x == 0) state[2] += 1u;
}
int dropout_advance(cudaStream_t stream, void* state) {
  DTB_REQUIRE(state != nullptr, DTB_ERR_BAD_ARG, "dropout_advance: null state");
  DTB_LAUNCH(dropout_advance_kernel, 1, 32, 0, stream, reinterpret_cast<uint32_t*>(state));
  DTB_CHECK_LAUNCH();
  return DTB_OK;
}
// keep bits of one site exactly as the fused kernels draw them: out[g * n + i] = keep(site, step, g, i)   (tests / oracle masks)
__global__ void __launch_bounds__(256) dropout_mask_kernel(const DropSite d, int groups, long long n, uint8_t* __restrict__ out) {
  const uint32_t key0 = drop_site_key(d);
  const long long total = (groups > 0 ? (long long)groups : 1LL) * n;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long g = i / n, e = i - g * n;
    const uint32_t key = groups > 0 ? drop_group_key(key0, (uint32_t)g) : key0;
    out[i] = drop_keep(key, (uint32_t)e, d.thr) ? 1 : 0;
  }
}
int dropout_mask(cudaStream_t stream, const DropSite& d, int groups, long long n, uint8_t* out) {
  DTB_REQUIRE(d.state && out && n >= 0 && groups >= 0, DTB_ERR_BAD_ARG, "dropout_mask: bad arguments");
  if (n == 0) return DTB_OK;
  dropout_mask_kernel<<<num_sms() * 4, 256, 0, stream>>>(d, groups, n, out);
  DTB_CHECK_LAUNCH();
  return DTB_OK;
}

}  // namespace dtb
