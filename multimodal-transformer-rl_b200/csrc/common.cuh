// Shared device/host helpers for the dtb200 sm_100a kernels.
// PTX wrappers for mbarrier, TMA (cp.async.bulk.tensor), tcgen05 (alloc / mma / commit / ld),
// plus small math + error plumbing.  Everything here is hand-written for sm_100a only.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include "../../include/dtb200.h"

#ifndef DTB_NUM_SMS_FALLBACK
#define DTB_NUM_SMS_FALLBACK 148
#endif

namespace dtb {

// ----------------------------------------------------------------------------------------------
// error plumbing (host)
// ----------------------------------------------------------------------------------------------
// error codes come from the public header (include/dtb200.h)

void set_error(const char* fmt, ...);
int cuda_fail(cudaError_t e, const char* what, const char* file, int line);

#define DTB_CUDA(expr)                                                        \
  do {                                                                        \
    cudaError_t _e = (expr);                                                  \
    if (_e != cudaSuccess) return ::dtb::cuda_fail(_e, #expr, __FILE__, __LINE__); \
  } while (0)

#define DTB_CHECK_LAUNCH()                                                    \
  do {                                                                        \
    ::dtb::count_launch();                                                    \
    cudaError_t _e = cudaGetLastError();                                      \
    if (_e != cudaSuccess) return ::dtb::cuda_fail(_e, "kernel launch", __FILE__, __LINE__); \
  } while (0)

#define DTB_REQUIRE(cond, code, ...)                                          \
  do {                                                                        \
    if (!(cond)) {                                                            \
      ::dtb::set_error(__VA_ARGS__);                                          \
      return (code);                                                          \
    }                                                                         \
  } while (0)

#define DTB_TRY(expr)                                                         \
  do {                                                                        \
    int _rc = (expr);                                                         \
    if (_rc != 0) return _rc;                                                 \
  } while (0)

int num_sms();
// runtime options (core.cu; C ABI: dtb_set_option / dtb_profile_set_balance)
int opt_splitk_inkernel();
int opt_ln_fused();
int opt_conv_implicit();
double opt_balance();
void opt_set_balance(double flop_per_byte);
int opt_set(const char* name, int value);
void count_launch();
long long launch_count();

// Optional live profiling (bench.py roofline): CUDA events recorded on the launching stream around each launch of a kernel family.
enum ProfFamily { PF_GEMM = 0, PF_CONV1_FWD, PF_CONV1_WGRAD, PF_ATTN_FWD, PF_ATTN_BWD, PF_LN_FWD, PF_LN_BWD, PF_EMBED_FWD, PF_COLSUM,
                  PF_IM2COL, PF_COL2IM, PF_GEMM_MEM /* GEMM shapes below the machine balance: work = algorithmic bytes */, PF_CONVN_FWD, PF_CONVN_DGRAD,
                  PF_CONVN_WGRAD /* implicit-GEMM conv2 / conv3: work = algorithmic bytes (input + output once) */, PF_COUNT };
bool prof_on();
int prof_begin(cudaStream_t stream, int family, double work);  // returns a slot (or -1)
void prof_end(cudaStream_t stream, int slot);
struct ProfScope {
  cudaStream_t st; int slot;
  ProfScope(cudaStream_t s, int family, double work) : st(s), slot(prof_on() ? prof_begin(s, family, work) : -1) {}
  ~ProfScope() { if (slot >= 0) prof_end(st, slot); }
};

// ---- kernel launch with programmatic dependent launch (PDL) ------------------------------------------------------------------
// Every kernel of the library starts with pdl_trigger() + pdl_wait(): launched with the programmatic-stream-serialization attribute,
// its CTAs may be scheduled (and run their prologue: barrier init, TMEM allocation, descriptor prefetch) while the previous kernel
// of the stream is still draining; griddepcontrol.wait then blocks until that kernel has completed and its writes are visible.
// In a captured CUDA graph the edge becomes a programmatic dependency.  DTB_NO_PDL=1 launches without the attribute (plain edges).
bool pdl_enabled();
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream, Args&&... args) {
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}
#define DTB_LAUNCH(kernel, grid, block, smem, stream, ...) \
  (void)::dtb::launch_pdl(kernel, dim3(grid), dim3(block), (size_t)(smem), stream, __VA_ARGS__)

typedef __nv_bfloat16 bf16;
typedef __nv_bfloat162 bf162;

// ----------------------------------------------------------------------------------------------
// device helpers
// ----------------------------------------------------------------------------------------------
#ifdef __CUDACC__

// PDL: let the next kernel of the stream start its prologue / wait for the previous kernel's results (no-ops without the attribute)
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// 128-bit streaming global accesses
__device__ __forceinline__ float4 ldg_f4(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }
__device__ __forceinline__ uint4 ldg_u4(const void* p) { return __ldg(reinterpret_cast<const uint4*>(p)); }

__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  bf162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ float2 unpack_bf16x2(uint32_t u) {
  bf162 v = *reinterpret_cast<bf162*>(&u);
  return __bfloat1622float2(v);
}

// activation functions (act: 0 none, 1 relu, 2 gelu_new = tanh approximation)
__device__ __forceinline__ float gelu_new_f(float x) {
  const float k0 = 0.7978845608028654f, k1 = 0.044715f;
  float u = k0 * (x + k1 * x * x * x);
  return 0.5f * x * (1.f + tanhf(u));
}
__device__ __forceinline__ float gelu_new_grad_f(float x) {
  const float k0 = 0.7978845608028654f, k1 = 0.044715f;
  float u = k0 * (x + k1 * x * x * x);
  float t = tanhf(u);
  float du = k0 * (1.f + 3.f * k1 * x * x);
  return 0.5f * (1.f + t) + 0.5f * x * (1.f - t * t) * du;
}

// ---- mbarrier -------------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.shared::cta.b64 st, [%0];\n\t}" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n\t}" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
// Bounded wait: a mis-programmed pipeline traps (launch error) instead of hanging the GPU box.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  if (mbar_try_wait(bar, parity)) return;
  long long t0 = clock64();
  while (!mbar_try_wait(bar, parity)) {
    if (clock64() - t0 > 4000000000LL) {  // ~2 s at 2 GHz
      printf("dtb200: mbarrier wait timeout (block %d thread %d)\n", blockIdx.x, threadIdx.x);
      __trap();
    }
  }
}

// ---- TMA ------------------------------------------------------------------------------------
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* m) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(m)) : "memory");
}
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
          smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}

// ---- tcgen05 --------------------------------------------------------------------------------
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

__device__ __forceinline__ void tmem_alloc(uint32_t* dst_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)), "r"(ncols)
               : "memory");
}
__device__ __forceinline__ void tmem_relinquish() {
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
// D[tmem] (+)= A[smem desc] * B[smem desc], bf16 inputs, fp32 accumulate. One thread issues.
__device__ __forceinline__ void umma_bf16(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// Arrive on an mbarrier once all previously issued tcgen05.mma of this thread have completed.
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// TMEM -> registers: this warp's 32 lanes x 32 consecutive fp32 columns.
__device__ __forceinline__ void tmem_ld_32x32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
        "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
        "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
}
// registers -> TMEM: this warp's 32 lanes x 32 consecutive fp32 columns.
__device__ __forceinline__ void tmem_st_32x32(uint32_t taddr, const uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};" ::"r"(taddr),
      "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]),
      "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]), "r"(v[16]), "r"(v[17]), "r"(v[18]), "r"(v[19]), "r"(v[20]), "r"(v[21]), "r"(v[22]),
      "r"(v[23]), "r"(v[24]), "r"(v[25]), "r"(v[26]), "r"(v[27]), "r"(v[28]), "r"(v[29]), "r"(v[30]), "r"(v[31])
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tma_load_3d(void* smem_dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(
          smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
__device__ __forceinline__ float fast_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// ---- dropout (counter-based, stateless) ----------------------------------------------------------------------------------------
// keep(site, step, group, idx) = bits32(key + idx) >= thr, thr = p * 2^32; key = f(seed, step, site[, group]).  The generator state
// {seed_lo, seed_hi, step} lives in device memory so that a captured CUDA graph draws a fresh mask on every replay (a 1-thread
// kernel advances `step`); forward and backward kernels of a site recompute the same bits, nothing is stored.  mix32 is the
// "lowbias32" integer finaliser (two multiply / xor-shift rounds, full avalanche).  Replaces nn.Dropout at the reference's three
// GPT-2 sites (HF modeling_gpt2: embd drop, attn_dropout on the softmax weights, resid_dropout after both c_proj).
struct DropSite {
  const uint32_t* state;  // device {seed_lo, seed_hi, step}; nullptr = dropout off
  uint32_t site;          // 0 embd, 1 + 3l attn, 2 + 3l resid (attn c_proj), 3 + 3l resid (mlp c_proj)
  uint32_t thr;           // drop iff mix32(key + idx) < thr
  float scale;            // 1 / (1 - p)
};
__host__ __device__ __forceinline__ uint32_t drop_mix(uint32_t x) {
  x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16;
  return x;
}
__host__ __device__ __forceinline__ uint32_t drop_key_from(uint32_t s0, uint32_t s1, uint32_t step, uint32_t site) {
  return drop_mix(drop_mix(s0 ^ (site * 0x9E3779B9U)) + drop_mix(s1 + step * 0x85EBCA6BU + 0x165667B1U));
}
__device__ __forceinline__ uint32_t drop_site_key(const DropSite& d) {
  return drop_key_from(__ldg(d.state), __ldg(d.state + 1), __ldg(d.state + 2), d.site);
}
__host__ __device__ __forceinline__ uint32_t drop_group_key(uint32_t key, uint32_t group) { return drop_mix(key ^ (group * 0xC2B2AE35U + 0x27D4EB2FU)); }
// per-element draw: the same finaliser without its last xor-shift (it only repairs the low bits; the threshold compare is decided
// by the high ones — lag / stride autocorrelation measured identical to the full finaliser)
__host__ __device__ __forceinline__ uint32_t drop_bits(uint32_t x) {
  x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU;
  return x;
}
__host__ __device__ __forceinline__ bool drop_keep(uint32_t key, uint32_t idx, uint32_t thr) { return drop_bits(key + idx) >= thr; }
__host__ __device__ __forceinline__ float drop_apply(float v, uint32_t key, uint32_t idx, uint32_t thr, float scale) {
  return drop_bits(key + idx) >= thr ? v * scale : 0.f;
}
static inline DropSite drop_none() { DropSite d; d.state = nullptr; d.site = 0; d.thr = 0; d.scale = 1.f; return d; }
static inline DropSite drop_make(const void* state, uint32_t site, float p) {
  DropSite d = drop_none();
  if (state && p > 0.f) {
    double t = (double)p * 4294967296.0;
    d.state = reinterpret_cast<const uint32_t*>(state);
    d.site = site;
    d.thr = t >= 4294967295.0 ? 4294967295u : (uint32_t)t;
    d.scale = (float)(1.0 / (1.0 - (double)d.thr / 4294967296.0));
  }
  return d;
}

// Shared-memory matrix descriptor for tcgen05.mma, 128-byte swizzle (SM100 descriptor version 1).
// K-major : rows of 128 B (64 bf16 along K), 8-row swizzle atoms 1024 B apart (SBO); LBO unused.
// MN-major: rows of 128 B (64 bf16 along M/N) indexed by k; 8-k atoms 1024 B apart (SBO);
//           the next 64-wide M/N chunk starts LBO bytes later.
__device__ __forceinline__ uint64_t make_smem_desc_sw128(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr & 0x3FFFF) >> 4);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;  // descriptor version (Blackwell)
  d |= (uint64_t)2 << 61;  // SWIZZLE_128B
  return d;
}
// Instruction descriptor: kind::f16, A=B=bf16, D=fp32, M x N tile.
__device__ __forceinline__ uint32_t make_idesc_bf16(uint32_t M, uint32_t N, uint32_t a_mn_major, uint32_t b_mn_major) {
  uint32_t d = 0;
  d |= 1u << 4;                  // D format fp32
  d |= 1u << 7;                  // A format bf16
  d |= 1u << 10;                 // B format bf16
  d |= (a_mn_major & 1u) << 15;  // A major
  d |= (b_mn_major & 1u) << 16;  // B major
  d |= (N >> 3) << 17;
  d |= (M >> 4) << 24;
  return d;
}

#endif  // __CUDACC__

}  // namespace dtb
