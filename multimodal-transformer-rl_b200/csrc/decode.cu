// KV-cached incremental get_action as ONE kernel: a thread-block cluster of 1-16 CTAs walks the whole
// B = 1 forward — embedding of the new tokens, every GPT-2 block, ln_f and the action head — for the 2-3 tokens a rollout step adds,
// attending to the keys / values of all earlier tokens kept in a per-layer cache.
//
// Why it is exact: the trunk has no positional embedding (timestep embeddings are absolute), attention is causal, so the hidden states
// of token j depend only on tokens <= j.  As long as the whole episode fits the window (t <= K: every Minecraft2d episode, Config.py:76)
// the left-padded recompute of the reference (A9) and "append to the cache" are the same function.  Once the window slides the host
// falls back to the recompute path (model.py).  One detail of the upstream call convention: the action of the newest timestep is a zero
// placeholder that becomes the real action on the next call, so the action token of timestep t-1 is (re)computed together with R_t, s_t.
//
// Two forms of the kernel (template parameter TC; dc_use_tc picks):
//  * TC = true, hidden 128 (c1 / c2 / c5), one CTA: the GEMVs run on the tensor core with the WEIGHTS as the A operand (128 output features
//    per tcgen05.mma, the <= 3 new tokens padded to 16 columns as the B operand) — M = 3 token rows as the A operand would waste > 97 % of
//    a 128-row tile, features-as-rows wastes nothing but columns.  Weights stream from L2 as pre-swizzled 16 KB blocks (see dt_gemv);
//    everything between the phases stays in shared memory.  Measured 72 us per step at c1 (the CUDA-core form: 127 us, issue / barrier bound).
//  * TC = false, any other width: every weight matrix is column-partitioned over the C CTAs of a cluster (each reads 1/C of it through a
//    cp.async.bulk tile ring, k split over the threads, partial sums combined in a fixed order), activations of the new tokens exchanged
//    through a tiny L2-resident scratch between cluster barriers (5 per layer).
// Both are deterministic and independent of how many tokens a call processes: a rollout that appends step by step and a single call over
// the same history produce the same bits.  Step mode (dtb_dt_decode_step): the step count comes from device memory, so one captured CUDA
// graph serves every step of a rollout.
#include "dt_handle.h"

#include <new>
#include <stdlib.h>
#include <string.h>

namespace dtb {
int require_sm100();

constexpr int DC_MAX_CTAS = 16;  // cluster size: 1 .. 16, chosen per model so that one CTA streams at most ~1.5 MB of weights per layer
constexpr int DC_THREADS = 512;  // 16 warps: the GEMV is issue / latency bound on few SMs (ncu: 28 % issue-active with 8 warps)
constexpr int DC_NT = 3;        // tokens per launch: a_{t-1}, R_t, s_t
constexpr int DC_MAXH = 1024;

constexpr int DC_RED_FLOATS = 4096 * DC_NT + 1536 * DC_NT;  // level 1: KG * NT * ncols <= 512 * 8 * NT; level 2: RP * n * ncols <= max(512, NT * 1536)

struct DecodeParams {
  int H, I, L, nh, d, sd, ad, E, act, tanh_act, sd_small;
  float eps;
  const float* P;   // fp32 master (biases, LayerNorms, embeddings, heads)
  const bf16* Pb;   // bf16 shadow (the GEMM weights, same offsets)
  const bf16* Wd;   // decode layout of the four weight matrices of every layer (dtb_dt_decode_stage)
  bf16* Wd_out;     // (staging kernel only)
  long long e_off[E_COUNT], l_off[L_COUNT], t_off[T_COUNT], layer_stride;
  // history (device pointers, one row per timestep)
  const float* states;      // [t, sd] fp32 (vector states) or null
  const bf16* states_bf16;  // [t, sd] bf16 (wide states = tower features) or null
  const float* actions;     // [t, ad]
  const float* rtg;         // [t]
  const long long* timesteps;
  int p0, n;                // tokens [p0, p0 + n) of the interleaved stream are processed by this launch
  int C;                    // CTAs of the cluster (the grid is one cluster)
  bf16* kc;                 // key cache   [L][Tmax][H]
  bf16* vc;                 // value cache [L][Tmax][H]
  int Tmax;
  float* scratch;           // exchange buffers, see dc_scratch_floats
  float* action_out;        // [ad]: prediction of the last processed token (written when it is a state token)
  int tc_slots, tc_red_floats;  // tensor-core path: ring slots that fit, length of `red`
  // step mode (dtb_dt_decode_step, CUDA-graph replayable): the timestep count t of this call is read from device memory (and advanced
  // by the kernel), the per-timestep histories are 2-row windows [t - 2, t - 1], the wide state of the newest timestep is state_new
  int* t_dev;
  const bf16* state_new;
  unsigned long long* stamps;   // developer probe (DTB_DECODE_STAMP=1): %globaltimer at the phase boundaries, written by thread 0
};

__device__ __forceinline__ void dc_stamp(const DecodeParams& p, int& k) {
  if (p.stamps && threadIdx.x == 0) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    p.stamps[k] = t;
  }
  ++k;
}

// scratch layout (floats): sp [H] | q [NT][H] | o [NT][H] | x1 [NT][H] | x2 [NT][H] | fc [NT][I]
__host__ __device__ inline size_t dc_scratch_floats(int H, int I) { return (size_t)H * (1 + 4 * DC_NT) + (size_t)DC_NT * I; }

__device__ __forceinline__ void dc_cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t dc_cluster_rank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ float dc_round_bf16(float v) { return __bfloat162float(__float2bfloat16(v)); }
__device__ __forceinline__ float dc_act(float v, int act) { return act == 1 ? fmaxf(v, 0.f) : gelu_new_f(v); }

// LayerNorm of the n token rows x[t][0..H) (shared memory) -> y[t][0..H) rounded to bf16 (what the tensor-core path feeds its GEMMs);
// warp t handles token t (two-pass statistics, as ln_fwd_kernel).
template <bool SMEM_PARAMS = false>
__device__ __forceinline__ void dc_layernorm(const float* x, float* y, int n, int H, const float* gamma, const float* beta, float eps) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp < n) {
    const float* xr = x + warp * H;
    float s = 0.f;
    for (int c = lane; c < H; c += 32) s += xr[c];
    const float mean = warp_sum(s) / H;
    float q = 0.f;
    for (int c = lane; c < H; c += 32) { const float dlt = xr[c] - mean; q += dlt * dlt; }
    const float rstd = rsqrtf(warp_sum(q) / H + eps);
    for (int c = lane; c < H; c += 32)
      y[warp * H + c] = dc_round_bf16((xr[c] - mean) * rstd * (SMEM_PARAMS ? gamma[c] : __ldg(gamma + c)) + (SMEM_PARAMS ? beta[c] : __ldg(beta + c)));
  }
  __syncthreads();
}

// ---- weight stream -------------------------------------------------------------------------------------------------------------------
// The kernel is bound by streaming this CTA's column slices of the four weight matrices of every layer from L2.  They are walked as ONE
// sequence of tiles (layer, matrix, k range): [rows][ncols] bf16, copied by TMA bulk copies (one per tile row, issued by warp 0) into
// a ring of DC_SLOTS shared-memory slots, DC_DEPTH tiles ahead of the tile being consumed.  The sequence runs straight through the phase
// boundaries of the forward, so the next matrix (and the next layer) is already landing while a phase waits at a cluster barrier, runs
// its LayerNorm or the attention.  (First version: __ldg in the GEMV loop — one dependent L2 round trip per 8 loads, 140 us at c1.)
constexpr int DC_SLOT_BYTES = 24 * 1024, DC_SLOTS = 4, DC_DEPTH = 3;

struct DcStream {
  int K[4], N[4], ncols[4], rows[4], tiles[4], lid[4];  // per matrix: q|k|v, attn.c_proj, mlp.c_fc, mlp.c_proj
  size_t moff[4], layer_elems;                           // decode layout: element offset of matrix m inside a layer, elements per layer
  int per_layer, total;
  int issued;  // tiles issued so far (identical in every thread)
};
__device__ __forceinline__ void dc_stream_init(DcStream& st, const DecodeParams& p) {
  const int H = p.H, I = p.I;
  const int K[4] = {H, H, H, I}, N[4] = {3 * H, H, I, H}, lid[4] = {L_WQKV, L_WO, L_W1, L_W2};
  st.per_layer = 0;
  st.layer_elems = 0;
  for (int m = 0; m < 4; ++m) {
    st.moff[m] = st.layer_elems;
    st.layer_elems += (size_t)K[m] * N[m];
    st.K[m] = K[m]; st.N[m] = N[m]; st.lid[m] = lid[m];
    st.ncols[m] = N[m] / p.C;
    int r = (DC_SLOT_BYTES / (st.ncols[m] * 2)) & ~7;
    st.rows[m] = r > K[m] ? K[m] : r;
    st.tiles[m] = (K[m] + st.rows[m] - 1) / st.rows[m];
    st.per_layer += st.tiles[m];
  }
  st.total = st.per_layer * p.L;
  st.issued = 0;
}
// Thread 0 issues tile g as ONE bulk copy (TMA, cp.async.bulk) completing on the slot's mbarrier: the weights are read from the decode
// layout (dtb_dt_decode_stage), where the [K][ncols] column slice of every (layer, matrix, CTA rank) is one contiguous block, so a tile —
// `rows` consecutive k of that block — is rows * ncols * 2 contiguous bytes.  (Earlier forms, measured: one 16-byte cp.async per thread
// and chunk = 24 % of all instructions in address arithmetic and an issue-bound kernel; one bulk copy per 100-800 byte row of the
// Conv1D-layout matrix = slower still, the copy engine pays per descriptor.)
__device__ __forceinline__ void dc_stream_issue(DcStream& st, const DecodeParams& p, int rank, uint8_t* slots, uint64_t* full) {
  const int g = st.issued++;
  if (g >= st.total || threadIdx.x != 0) return;
  const int l = g / st.per_layer;
  int r = g - l * st.per_layer, m = 0;
  while (r >= st.tiles[m]) { r -= st.tiles[m]; ++m; }
  const int k0 = r * st.rows[m], rows = min(st.rows[m], st.K[m] - k0), ncols = st.ncols[m];
  const bf16* W = p.Wd + (size_t)l * st.layer_elems + st.moff[m] + ((size_t)rank * st.K[m] + k0) * ncols;
  const int slot = g % DC_SLOTS;
  const uint32_t bytes = (uint32_t)rows * (uint32_t)ncols * 2u;
  mbar_arrive_expect_tx(&full[slot], bytes);
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(slots + (size_t)slot * DC_SLOT_BYTES)),
               "l"(W), "r"(bytes), "r"(smem_u32(&full[slot]))
               : "memory");
}

// Conv1D-layout bf16 weights (the shadow the tensor-core path reads) -> decode layout: for every layer, matrix (q|k|v, attn.c_proj,
// mlp.c_fc, mlp.c_proj) and CTA rank one contiguous [K][N / C] block.  One thread per 16-byte chunk.
__global__ void __launch_bounds__(256) dt_decode_stage_kernel(const DecodeParams p, size_t layer_elems) {
  pdl_trigger();
  pdl_wait();
  const int H = p.H, I = p.I;
  const int K[4] = {H, H, H, I}, N[4] = {3 * H, H, I, H}, lid[4] = {L_WQKV, L_WO, L_W1, L_W2};
  const size_t chunks_per_layer = layer_elems / 8, total = chunks_per_layer * p.L;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int l = (int)(i / chunks_per_layer);
    size_t e = (i - (size_t)l * chunks_per_layer) * 8;  // element offset inside the layer (decode layout)
    int m = 0;
    while (e >= (size_t)K[m] * N[m]) { e -= (size_t)K[m] * N[m]; ++m; }
    const int ncols = N[m] / p.C;
    const int r = (int)(e / ((size_t)K[m] * ncols));
    const size_t q = e - (size_t)r * K[m] * ncols;
    const int k = (int)(q / ncols), c = (int)(q - (size_t)k * ncols);
    const bf16* src = p.Pb + p.l_off[lid[m]] + (long long)l * p.layer_stride + (size_t)k * N[m] + r * ncols + c;
    *reinterpret_cast<uint4*>(p.Wd_out + (size_t)l * layer_elems + (i - (size_t)l * chunks_per_layer) * 8) = __ldg(reinterpret_cast<const uint4*>(src));
  }
}

// Column slice of  out[t][c] = sum_k in[t][k] W[k][c]  for the n tokens, W = matrix m of the current layer, read from the tile ring.
// Thread (kg, cg): 8 adjacent columns (one 16-byte shared-memory load per k), k = kg, kg + KG, ... inside every tile; the KG partial
// sums of a column are added in a fixed order (two levels).  `consumed` = tiles consumed so far (the stream position of this gemv).
// Result left in red[t * ncols + j].
__device__ __forceinline__ void dc_gemv(DcStream& st, const DecodeParams& p, int rank, uint8_t* slots, uint64_t* full, int& consumed, int m,
                                        const float* in, int n, float* red) {
  const int K = st.K[m], ncols = st.ncols[m];
  const int CG = ncols >> 3;                 // 8-column groups of the slice
  const int KG = DC_THREADS / CG;            // k-lanes
  const int cg = threadIdx.x % CG, kg = threadIdx.x / CG;
  float acc[DC_NT][8];
#pragma unroll
  for (int t = 0; t < DC_NT; ++t)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[t][j] = 0.f;
  for (int tile = 0; tile < st.tiles[m]; ++tile, ++consumed) {
    __syncthreads();                           // everybody is done with tile consumed - 1: its slot is free for tile consumed + DEPTH
    dc_stream_issue(st, p, rank, slots, full);
    mbar_wait(&full[consumed % DC_SLOTS], (uint32_t)(consumed / DC_SLOTS) & 1u);  // tile `consumed` has landed
    const uint8_t* slot = slots + (size_t)(consumed % DC_SLOTS) * DC_SLOT_BYTES;
    const int k0 = tile * st.rows[m], rows = min(st.rows[m], K - k0);
    if (kg < KG) {
      // all DC_NT accumulator rows are always computed (rows >= n read whatever the input buffer holds and are never used): no branches,
      // pointers stepped by constants — the loop is 1 LDS.128 + 3 LDS + 8 unpack + 24 FFMA
      const uint8_t* wp = slot + (size_t)(kg * ncols + cg * 8) * 2;
      const float* ip = in + k0 + kg;
      const int wstep = KG * ncols * 2;
      for (int r = kg; r < rows; r += KG, wp += wstep, ip += KG) {
        const uint4 w = *reinterpret_cast<const uint4*>(wp);
        const uint32_t ww[4] = {w.x, w.y, w.z, w.w};
        const float a0 = ip[0], a1 = ip[K], a2 = ip[2 * K];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float lo = __uint_as_float(ww[e] << 16), hi = __uint_as_float(ww[e] & 0xffff0000u);
          acc[0][2 * e] = fmaf(a0, lo, acc[0][2 * e]); acc[0][2 * e + 1] = fmaf(a0, hi, acc[0][2 * e + 1]);
          acc[1][2 * e] = fmaf(a1, lo, acc[1][2 * e]); acc[1][2 * e + 1] = fmaf(a1, hi, acc[1][2 * e + 1]);
          acc[2][2 * e] = fmaf(a2, lo, acc[2][2 * e]); acc[2][2 * e + 1] = fmaf(a2, hi, acc[2][2 * e + 1]);
        }
      }
    }
  }
  if (kg < KG) {
#pragma unroll
    for (int t = 0; t < DC_NT; ++t)
#pragma unroll
      for (int j = 0; j < 8; ++j) red[((size_t)kg * DC_NT + t) * ncols + cg * 8 + j] = acc[t][j];
  }
  __syncthreads();
  // the KG partial sums of every output are added in kg order, in two levels (fixed grouping: deterministic): RP threads per output sum
  // contiguous kg ranges, then one thread adds the RP range sums
  const int nout = n * ncols;
  int RP = DC_THREADS / nout;
  RP = RP < 1 ? 1 : (RP > KG ? KG : RP);
  float* red2 = red + (size_t)KG * DC_NT * ncols;  // [RP][nout]
  for (int i = threadIdx.x; i < nout * RP; i += DC_THREADS) {
    const int r = i / nout, o = i - r * nout, t = o / ncols, j = o - t * ncols;
    const int g0 = (int)((long long)r * KG / RP), g1 = (int)((long long)(r + 1) * KG / RP);
    float s = 0.f;
    for (int g = g0; g < g1; ++g) s += red[((size_t)g * DC_NT + t) * ncols + j];
    red2[(size_t)r * nout + o] = s;
  }
  __syncthreads();
  for (int o = threadIdx.x; o < nout; o += DC_THREADS) {
    float s = red2[o];
    for (int r = 1; r < RP; ++r) s += red2[(size_t)r * nout + o];
    red[o] = s;  // result: red[t * ncols + j]
  }
  __syncthreads();
}

// ---- tensor-core GEMV (hidden % 128 == 0) -----------------------------------------------------------------------------------------------
// The FFMA GEMV above is bound by instruction issue and by its per-tile block barriers on the one SM a small model uses (c1: 147 us for
// 1.2 MB of weights).  Here the WEIGHTS are the A operand of tcgen05.mma (M = 128 output features per instruction, MN-major: the Conv1D
// layout [k][feature] is read as it is) and the <= 3 new tokens are the B operand (N = 16 token slots, K-major, written by the threads into
// the 128B-swizzled image): D[feature][token] accumulates in TMEM, one thread issues everything, and no thread ever touches a weight.
// The weights are staged once per parameter update as a sequence of 16 KB blocks — [128 features][64 k], already in the swizzled
// shared-memory image — so one plain bulk copy (cp.async.bulk) per block lands MMA-ready data; thread 0 keeps `depth` blocks in flight,
// straight through the phase boundaries of the forward.  Columns of D are independent, so the bits of a token do not depend on which
// slot it occupies or on how a history is chunked into launches.
constexpr int DT_BLOCK_BYTES = 16384, DT_BLOCK_ELEMS = 8192, DT_MAX_SLOTS = 8, DT_TMEM_COLS = 64, DT_NTOK = 16;

struct DcTc {
  int K[4], ncols[4], ntile[4], nkb[4];  // per matrix: k extent, this CTA's columns, its 128-feature tiles, 64-wide k-blocks
  size_t moff[4], layer_elems;
  int per_layer, total;   // blocks per layer / in the whole forward (per CTA)
  int slots, depth, sshift;  // ring slots (a power of two), blocks in flight, log2(slots)
  // thread 0 only: stream position
  int issued, consumed, is_l, is_m, is_b;
  uint32_t acc_phase;     // every thread: parity of the accumulator barrier
  uint32_t tmem;
  uint8_t *ring, *bop;
  uint64_t *full, *empty, *acc;
};

__device__ __forceinline__ void dt_init(DcTc& tc, const DecodeParams& p) {
  const int H = p.H, I = p.I;
  const int K[4] = {H, H, H, I}, N[4] = {3 * H, H, I, H};
  tc.per_layer = 0;
  tc.layer_elems = 0;
  for (int m = 0; m < 4; ++m) {
    tc.moff[m] = tc.layer_elems;
    tc.layer_elems += (size_t)K[m] * N[m];
    tc.K[m] = K[m];
    tc.ncols[m] = N[m] / p.C;
    tc.ntile[m] = tc.ncols[m] / 128;
    tc.nkb[m] = K[m] / 64;
    tc.per_layer += tc.ntile[m] * tc.nkb[m];
  }
  tc.total = tc.per_layer * p.L;
  tc.issued = tc.consumed = tc.is_l = tc.is_m = tc.is_b = 0;
  tc.acc_phase = 0;
}
// thread 0: issue the next weight block of this CTA's stream (if any) into its ring slot
__device__ __forceinline__ void dt_issue(DcTc& tc, const DecodeParams& p, int rank) {
  if (tc.issued >= tc.total) return;
  const int g = tc.issued++;
  const int slot = g % tc.slots;
  if (g >= tc.slots) mbar_wait(&tc.empty[slot], (uint32_t)(g / tc.slots - 1) & 1u);  // the MMAs that read this slot have retired
  const int m = tc.is_m;
  const bf16* src = p.Wd + (size_t)tc.is_l * tc.layer_elems + tc.moff[m] + (size_t)rank * tc.K[m] * tc.ncols[m] + (size_t)tc.is_b * DT_BLOCK_ELEMS;
  mbar_arrive_expect_tx(&tc.full[slot], DT_BLOCK_BYTES);
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(tc.ring + (size_t)slot * DT_BLOCK_BYTES)),
               "l"(src), "r"((uint32_t)DT_BLOCK_BYTES), "r"(smem_u32(&tc.full[slot]))
               : "memory");
  if (++tc.is_b == tc.ntile[m] * tc.nkb[m]) {
    tc.is_b = 0;
    if (++tc.is_m == 4) { tc.is_m = 0; ++tc.is_l; }
  }
}
__device__ __forceinline__ void dt_tmem_ld4(uint32_t taddr, uint32_t (&v)[4]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]) : "r"(taddr) : "memory");
}

// Conv1D-layout bf16 weights -> the tensor-core decode layout: per layer, matrix and CTA rank the blocks (feature tile, k-block) in stream
// order; inside a block two 64-feature halves of [64 k][64 features] rows (128 B each), 16-byte units XOR-swizzled by (k & 7) — exactly
// what a TMA box {64 features, 64 k} with SWIZZLE_128B would have written.  One thread per 16-byte unit.
__global__ void __launch_bounds__(256) dt_decode_stage_tc_kernel(const DecodeParams p, size_t layer_elems) {
  pdl_trigger();
  pdl_wait();
  const int H = p.H, I = p.I;
  const int K[4] = {H, H, H, I}, N[4] = {3 * H, H, I, H}, lid[4] = {L_WQKV, L_WO, L_W1, L_W2};
  const size_t chunks_per_layer = layer_elems / 8, total = chunks_per_layer * p.L;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int l = (int)(i / chunks_per_layer);
    size_t e = (i - (size_t)l * chunks_per_layer) * 8;  // element offset inside the layer (decode layout)
    int m = 0;
    while (e >= (size_t)K[m] * N[m]) { e -= (size_t)K[m] * N[m]; ++m; }
    const int ncols = N[m] / p.C, nkb = K[m] / 64;
    const size_t slice = (size_t)K[m] * ncols;
    const int r = (int)(e / slice);
    const size_t q = e - (size_t)r * slice;
    const int blk = (int)(q / DT_BLOCK_ELEMS), w = (int)(q - (size_t)blk * DT_BLOCK_ELEMS);
    const int ft = blk / nkb, kb = blk - ft * nkb;
    const int j = w >> 12, rr = (w & 4095) >> 6, up = (w & 63) >> 3;
    const int row = kb * 64 + rr, col = r * ncols + ft * 128 + j * 64 + ((up ^ (rr & 7)) << 3);
    const bf16* src = p.Pb + p.l_off[lid[m]] + (long long)l * p.layer_stride + (size_t)row * N[m] + col;
    *reinterpret_cast<uint4*>(p.Wd_out + (size_t)l * layer_elems + (i - (size_t)l * chunks_per_layer) * 8) = __ldg(reinterpret_cast<const uint4*>(src));
  }
}

// out[t][c] = sum_k in[t][k] W[k][c] for this CTA's columns of matrix m, on the tensor core; result left in red[t * ncols + c].
// epi(f, o): called by the thread that owns output feature f of this CTA's slice with the DC_NT token values o[] (bias, activation, residual
// and the store happen there: the accumulators never take a detour through shared memory).
template <typename Epi>
__device__ __forceinline__ void dt_gemv(DcTc& tc, const DecodeParams& p, int rank, int m, const float* in, int n, Epi epi) {
  const int K = tc.K[m], ku_per_row = K >> 3;
  // B operand: token rows (bf16, the inputs are already bf16-rounded values) in the K-major 128B-swizzled image, one 2 KB region per k-block
  for (int idx = threadIdx.x; idx < DC_NT * ku_per_row; idx += DC_THREADS) {
    const int t = idx / ku_per_row, ku = idx - t * ku_per_row;
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    if (t < n) {
      const float4 a = *reinterpret_cast<const float4*>(in + (size_t)t * K + ku * 8), b = *reinterpret_cast<const float4*>(in + (size_t)t * K + ku * 8 + 4);
      v = make_uint4(pack_bf16x2(a.x, a.y), pack_bf16x2(a.z, a.w), pack_bf16x2(b.x, b.y), pack_bf16x2(b.z, b.w));
    }
    const uint32_t addr = smem_u32(tc.bop) + (uint32_t)(ku >> 3) * 2048u + (uint32_t)t * 128u + (uint32_t)(((ku & 7) ^ (t & 7)) << 4);
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
  }
  fence_proxy_async();  // generic-proxy stores -> visible to the tensor core (async proxy)
  __syncthreads();
  // Jobs: the matrix is cut into <= 4 independent accumulations — its 128-feature tiles, and when there are fewer than 4 tiles also
  // halves / quarters of k (partial sums in separate TMEM columns, added below in a fixed order: deterministic).  Lane 0 of warps 0-3
  // (one per SM sub-partition) each walks the blocks of one job: wait for the block, 4 MMAs (K = 16 each), release the slot on
  // retirement, and issue the load `depth` blocks further down the stream (every block is consumed by exactly one thread, so every
  // load is issued exactly once).  The scalar chain per block (~0.5 us of waits, descriptor arithmetic and async-proxy instructions)
  // is what paces this kernel, hence four chains side by side.  Job order = stream order (tile-major, then k).
  const int ntile = tc.ntile[m], nkb = tc.nkb[m];
  int ksplit = 1;
  while (ntile * ksplit * 2 <= 4 && nkb % (ksplit * 2) == 0) ksplit *= 2;
  const int njobs = ntile * ksplit, bpj = nkb / ksplit;
  const uint32_t g_base = (uint32_t)tc.consumed;
  if ((threadIdx.x & 31) == 0 && threadIdx.x < 128) {
    const int j = threadIdx.x >> 5;
    if (j < njobs) {
      tc_fence_after();
      const uint32_t idesc = make_idesc_bf16(128, DT_NTOK, 1, 0);
      const uint32_t smask = (uint32_t)tc.slots - 1u, sshift = (uint32_t)tc.sshift, depth = (uint32_t)tc.depth, total = (uint32_t)tc.total;
      const uint64_t adesc0 = make_smem_desc_sw128(smem_u32(tc.ring), 8192, 1024);  // MN-major: 16 k rows of 128 B per step, halves 8 KB apart
      const uint64_t bdesc0 = make_smem_desc_sw128(smem_u32(tc.bop), 16, 1024);     // K-major: 32 B per step inside the 128 B row
      const uint32_t ring0 = smem_u32(tc.ring), full0 = smem_u32(tc.full);
      const uint32_t d_tmem = tc.tmem + (uint32_t)j * DT_NTOK;
      uint32_t g = g_base + (uint32_t)(j * bpj);
      int kb = (j % ksplit) * bpj;
      for (int b = 0; b < bpj; ++b, ++g, ++kb) {
        const uint32_t slot = g & smask;
        mbar_wait(&tc.full[slot], (g >> sshift) & 1u);
        tc_fence_after();
        const uint64_t ad = adesc0 + (uint64_t)(slot * (DT_BLOCK_BYTES >> 4)), bd = bdesc0 + (uint64_t)(kb * (2048 >> 4));
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) umma_bf16(d_tmem, ad + kk * (2048 >> 4), bd + kk * (32 >> 4), idesc, (b > 0 || kk > 0) ? 1u : 0u);
        umma_commit(&tc.empty[slot]);  // frees the ring slot when these MMAs retire
        const uint32_t g2 = g + depth;
        if (g2 < total) {
          const uint32_t s2 = g2 & smask;
          if (g2 > smask) mbar_wait(&tc.empty[s2], ((g2 >> sshift) - 1u) & 1u);  // its previous block (g2 - slots) has retired
          const bf16* src = p.Wd + (size_t)g2 * DT_BLOCK_ELEMS;                   // C = 1: the whole stream is one contiguous array
          mbar_arrive_expect_tx(&tc.full[s2], DT_BLOCK_BYTES);
          asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(ring0 + s2 * (uint32_t)DT_BLOCK_BYTES),
                       "l"(src), "r"((uint32_t)DT_BLOCK_BYTES), "r"(full0 + s2 * 8u)
                       : "memory");
        }
      }
      umma_commit(tc.acc);  // this job's accumulator is final
    } else {
      mbar_arrive(tc.acc);
    }
  }
  tc.consumed += ntile * nkb;
  mbar_wait(tc.acc, tc.acc_phase);
  tc.acc_phase ^= 1u;
  tc_fence_after();
  {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, q = warp & 3;
    for (int ft = warp >> 2; ft < ntile; ft += DC_THREADS / 128) {
      float o[DC_NT];
#pragma unroll
      for (int t = 0; t < DC_NT; ++t) o[t] = 0.f;
      for (int ks = 0; ks < ksplit; ++ks) {  // partial sums of the k ranges, in k order
        uint32_t v[4];
        dt_tmem_ld4(tc.tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)(ft * ksplit + ks) * DT_NTOK, v);
        tmem_ld_wait();
#pragma unroll
        for (int t = 0; t < DC_NT; ++t) o[t] += __uint_as_float(v[t]);
      }
      epi(ft * 128 + q * 32 + lane, o);
    }
  }
  tc_fence_before();
  __syncthreads();
}

template <bool TC>
__global__ void __launch_bounds__(DC_THREADS, 1) dt_decode_kernel(const DecodeParams p) {
  extern __shared__ __align__(16) float dc_smem[];
  const int H = p.H, I = p.I, d = p.d;
  int p0 = p.p0, n = p.n, hbase = 0;
  const int KI = H > I ? H : I;
  float* x = dc_smem;                 // [NT][H]  residual stream of the new tokens (replicated in every CTA)
  float* hb = x + DC_NT * H;          // [NT][KI] GEMV input (LayerNorm output / attention output / MLP hidden), bf16-rounded values
  float* red = hb + DC_NT * KI;       // DC_RED_FLOATS: GEMV partial sums (two levels), attention merge
  uint8_t* slots = reinterpret_cast<uint8_t*>(red + DC_RED_FLOATS);  // weight tile ring (16-byte aligned: every region above is a multiple of 16 B)
  uint64_t* full = reinterpret_cast<uint64_t*>(slots + (size_t)DC_SLOTS * DC_SLOT_BYTES);  // [DC_SLOTS] tile-landed barriers
  DcTc tc;
  float *lx = nullptr, *pl = nullptr;
  if (TC) {  // tensor-core path (one CTA): red is tc_red_floats long, then the exchange buffers and this layer's small parameters (both in
             // shared memory: no L2 round trip between the phases), then (1 KB aligned) the B-operand image, the block ring, the barriers
    lx = red + p.tc_red_floats;          // sp [H] | q, o, x1, x2 [NT][H] | fc [NT][I]
    pl = lx + 13 * H + DC_NT * I;        // ln_1.w, ln_1.b [H] | b_qkv [3H] | b_o [H] | ln_2.w, ln_2.b [H] | b_fc [I] | b_proj [H]
    uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(pl + 9 * H + I) + 1023) & ~uintptr_t(1023));
    tc.bop = base;
    tc.ring = base + (size_t)(KI / 64) * 2048;
    tc.slots = p.tc_slots;
    tc.sshift = 31 - __clz(p.tc_slots);
    tc.depth = p.tc_slots > 4 ? p.tc_slots - 3 : p.tc_slots - 1;  // a refilled slot was consumed >= 3 blocks ago: its MMAs have retired, thread 0 never waits for them
    tc.full = reinterpret_cast<uint64_t*>(tc.ring + (size_t)tc.slots * DT_BLOCK_BYTES);
    tc.empty = tc.full + DT_MAX_SLOTS;
    tc.acc = tc.empty + DT_MAX_SLOTS;
  }
  const int DC_CTAS = p.C;
  const int rank = (int)dc_cluster_rank();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const float* P = p.P;
  const bf16* Pb = p.Pb;
  float* s_sp = TC ? lx : p.scratch;        // [H]      wide-state projection
  float* s_q = s_sp + H;                    // [NT][H]
  float* s_o = TC ? hb : s_q + DC_NT * H;   // [NT][H]  (tensor-core path: the attention output IS the next GEMV input)
  float* s_x1 = s_o + DC_NT * H;            // [NT][H]
  float* s_x2 = s_x1 + DC_NT * H;           // [NT][H]
  float* s_fc = s_x2 + DC_NT * H;           // [NT][I]
  pdl_trigger();
  pdl_wait();
  int sk = 0;
  dc_stamp(p, sk);  // 0: start
  if (p.t_dev) {  // step mode: tokens a_{t-2}, R_{t-1}, s_{t-1} (0-based timesteps) of the t-step history; rows of the 2-row windows = k - (t - 2)
    const int t_now = *reinterpret_cast<volatile int*>(p.t_dev);
    p0 = t_now > 1 ? 3 * (t_now - 1) - 1 : 0;
    n = 3 * t_now - 1 - p0;
    hbase = t_now - 2;
  }
  // start the weight stream: the first DEPTH tiles fly while the tokens are embedded
  DcStream st;
  int consumed = 0;
  if (TC) {
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tc.acc + 1);
    if (threadIdx.x == 0) {
      for (int i = 0; i < DT_MAX_SLOTS; ++i) { mbar_init(&tc.full[i], 1); mbar_init(&tc.empty[i], 1); }
      mbar_init(tc.acc, 4);  // one arrival per issuing thread (lane 0 of warps 0-3)
      fence_mbar_init();
    }
    if (warp == 2) {
      tmem_alloc(tmem_slot, DT_TMEM_COLS);
      tmem_relinquish();
    }
    // token slots n .. 15 of the B operand are never written again: zero them once (their D columns are never read, but keep them finite)
    for (int i = threadIdx.x; i < (KI / 64) * 2048 / 16; i += DC_THREADS)
      *reinterpret_cast<uint4*>(tc.bop + (size_t)i * 16) = make_uint4(0u, 0u, 0u, 0u);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    tc.tmem = *tmem_slot;
    dt_init(tc, p);
    if (threadIdx.x == 0)
      for (int i = 0; i < tc.depth; ++i) dt_issue(tc, p, rank);
  } else {
    if (threadIdx.x == 0) {
      for (int i = 0; i < DC_SLOTS; ++i) mbar_init(&full[i], 1);
      fence_mbar_init();
    }
    __syncthreads();
    dc_stream_init(st, p);
    for (int i = 0; i < DC_DEPTH; ++i) dc_stream_issue(st, p, rank, slots, full);
  }

  // ---- wide state (tower features, bf16): projection Ws[H][sd] . s, column slice per CTA, one warp per output --------------------
  int wide_tok = -1;
  if (!p.sd_small)
    for (int t = 0; t < n; ++t)
      if ((p0 + t) % 3 == 1) wide_tok = t;  // at most one state token among <= 3 consecutive tokens
  if (wide_tok >= 0) {
    const int k = (p0 + wide_tok) / 3, per = H / DC_CTAS;
    const bf16* srow = p.t_dev ? p.state_new : p.states_bf16 + (size_t)k * p.sd;
    // 8 consecutive outputs per warp and pass; per 256-wide k chunk the state row and the 8 weight rows are loaded before the first FMA
    const int end = (rank + 1) * per;
    for (int cb = rank * per + warp * 8; cb < end; cb += (DC_THREADS / 32) * 8) {
      float a[8];
#pragma unroll
      for (int o = 0; o < 8; ++o) a[o] = 0.f;
#pragma unroll 2
      for (int j = lane * 8; j < p.sd; j += 256) {  // sd % 8 == 0 (checked at dtb_dt_create)
        const uint4 sv = __ldg(reinterpret_cast<const uint4*>(srow + j));
        uint4 wv[8];
#pragma unroll
        for (int o = 0; o < 8; ++o)
          wv[o] = cb + o < end ? __ldg(reinterpret_cast<const uint4*>(Pb + p.e_off[E_WS] + (size_t)(cb + o) * p.sd + j)) : make_uint4(0u, 0u, 0u, 0u);
        const float2 s0 = unpack_bf16x2(sv.x), s1 = unpack_bf16x2(sv.y), s2 = unpack_bf16x2(sv.z), s3 = unpack_bf16x2(sv.w);
#pragma unroll
        for (int o = 0; o < 8; ++o) {
          const float2 w0 = unpack_bf16x2(wv[o].x), w1 = unpack_bf16x2(wv[o].y), w2 = unpack_bf16x2(wv[o].z), w3 = unpack_bf16x2(wv[o].w);
          float v = a[o];
          v = fmaf(w0.x, s0.x, v); v = fmaf(w0.y, s0.y, v); v = fmaf(w1.x, s1.x, v); v = fmaf(w1.y, s1.y, v);
          v = fmaf(w2.x, s2.x, v); v = fmaf(w2.y, s2.y, v); v = fmaf(w3.x, s3.x, v); v = fmaf(w3.y, s3.y, v);
          a[o] = v;
        }
      }
#pragma unroll
      for (int o = 0; o < 8; ++o) {
        const float r = warp_sum(a[o]);
        if (lane == 0 && cb + o < end) s_sp[cb + o] = r;
      }
    }
    if (TC) __syncthreads(); else dc_cluster_sync();
  }

  // ---- embedding + embed_ln of the new tokens (HF:611-628), replicated in every CTA ------------------------------------------------
  for (int idx = threadIdx.x; idx < n * H; idx += DC_THREADS) {  // all tokens at once: their dependent loads (timestep -> table row) overlap
    const int t = idx / H, c = idx - t * H;
    const int tok = p0 + t, k = tok / 3, type = tok - 3 * k;
    long long ts = p.timesteps[k - hbase];
    ts = ts < 0 ? 0 : (ts >= p.E ? p.E - 1 : ts);
    const float* te = P + p.e_off[E_TABLE] + (size_t)ts * H;
    float v = __ldg(te + c);
    if (type == 0) {
      v += p.rtg[k - hbase] * __ldg(P + p.e_off[E_WR] + c) + __ldg(P + p.e_off[E_BR] + c);
    } else if (type == 1) {
      v += __ldg(P + p.e_off[E_BS] + c);
      if (p.sd_small) {
        const float* w = P + p.e_off[E_WS] + (size_t)c * p.sd;
        const float* srow = p.states + (size_t)(k - hbase) * p.sd;
        for (int j = 0; j < p.sd; ++j) v = fmaf(__ldg(w + j), __ldg(srow + j), v);
      } else {
        v += (TC ? s_sp[c] : __ldcg(s_sp + c));
      }
    } else {
      v += __ldg(P + p.e_off[E_BA] + c);
      const float* w = P + p.e_off[E_WA] + (size_t)c * p.ad;
      const float* arow = p.actions + (size_t)(k - hbase) * p.ad;
      for (int j = 0; j < p.ad; ++j) v = fmaf(__ldg(w + j), __ldg(arow + j), v);
    }
    hb[t * H + c] = v;
  }
  __syncthreads();
  if (warp < n) {  // embed_ln -> x (fp32, not rounded: the residual stream)
    const float* xr = hb + warp * H;
    float s = 0.f;
    for (int c = lane; c < H; c += 32) s += xr[c];
    const float mean = warp_sum(s) / H;
    float q = 0.f;
    for (int c = lane; c < H; c += 32) { const float dl = xr[c] - mean; q += dl * dl; }
    const float rstd = rsqrtf(warp_sum(q) / H + p.eps);
    for (int c = lane; c < H; c += 32)
      x[warp * H + c] = (xr[c] - mean) * rstd * __ldg(P + p.e_off[E_LNW] + c) + __ldg(P + p.e_off[E_LNB] + c);
  }
  __syncthreads();

  const float scale_log2 = 1.4426950408889634f * rsqrtf((float)d);
  dc_stamp(p, sk);  // 1: embedded
  for (int l = 0; l < p.L; ++l) {
    const long long lo = (long long)l * p.layer_stride;
    bf16* kc = p.kc + (size_t)l * p.Tmax * H;
    bf16* vc = p.vc + (size_t)l * p.Tmax * H;
    // ---- A: ln_1 + this CTA's columns of q | k | v (HF:272, :179-185) ---------------------------------------------------------------
    if (TC) {
      // this layer's LayerNorm vectors and biases -> shared memory, one float4 per thread and pass: a single L2 round trip per layer instead
      // of one in front of every phase
      const int seg_len[8] = {H, H, 3 * H, H, H, H, I, H};
      const int seg_id[8] = {L_LN1W, L_LN1B, L_BQKV, L_BO, L_LN2W, L_LN2B, L_B1, L_B2};
      for (int i = threadIdx.x * 4; i < 9 * H + I; i += DC_THREADS * 4) {
        int sgm = 0, off = i;
        while (off >= seg_len[sgm]) { off -= seg_len[sgm]; ++sgm; }
        *reinterpret_cast<float4*>(pl + i) = __ldg(reinterpret_cast<const float4*>(P + p.l_off[seg_id[sgm]] + lo + off));
      }
      __syncthreads();
      dc_layernorm<true>(x, hb, n, H, pl, pl + H, p.eps);
    } else {
      dc_layernorm(x, hb, n, H, P + p.l_off[L_LN1W] + lo, P + p.l_off[L_LN1B] + lo, p.eps);
    }
    {
      const int ncols = 3 * H / DC_CTAS, c0 = rank * ncols;
      if (TC) {
        dt_gemv(tc, p, rank, 0, hb, n, [&](int c, const float (&o)[DC_NT]) {
          const float b = pl[2 * H + c];
          for (int t = 0; t < n; ++t) {
            const float v = o[t] + b;
            if (c < H) s_q[t * H + c] = dc_round_bf16(v);
            else if (c < 2 * H) kc[(size_t)(p0 + t) * H + (c - H)] = __float2bfloat16(v);
            else vc[(size_t)(p0 + t) * H + (c - 2 * H)] = __float2bfloat16(v);
          }
        });
      } else {
        dc_gemv(st, p, rank, slots, full, consumed, 0, hb, n, red);
      }
      for (int i = threadIdx.x; i < n * ncols && !TC; i += DC_THREADS) {
        const int t = i / ncols, c = c0 + (i - t * ncols);
        const float v = red[i] + __ldg(P + p.l_off[L_BQKV] + lo + c);
        if (c < H) s_q[t * H + c] = dc_round_bf16(v);
        else if (c < 2 * H) kc[(size_t)(p0 + t) * H + (c - H)] = __float2bfloat16(v);
        else vc[(size_t)(p0 + t) * H + (c - 2 * H)] = __float2bfloat16(v);
      }
    }
    if (!TC) dc_cluster_sync();  // (tensor-core path: dt_gemv ends with a block barrier and there is no peer CTA)
    dc_stamp(p, sk);  // +0: q | k | v done
    // ---- B: causal attention of the new tokens over the cache (HF:47-65); heads rank, rank + 8, ... ---------------------------------
    // (Tried and dropped: one warp per (token, head) with lanes owning whole keys — 16.6 us instead of 4.7 us per layer at c1: with three
    // busy warps the scattered K rows and the in-order P.V walk are pure load latency.  Keys spread over all 16 warps it stays.)
    for (int hh = rank; hh < p.nh; hh += DC_CTAS) {
      // warp w takes keys w, w + 8, ...; lane owns d / 32 dims; per (warp, token) online softmax, merged in warp order below
      const int per = d >> 5;  // 1, 2 or 4
      float qv[DC_NT][4], acc[DC_NT][4], m[DC_NT], lsum[DC_NT];
#pragma unroll
      for (int t = 0; t < DC_NT; ++t) {
        m[t] = -INFINITY; lsum[t] = 0.f;
#pragma unroll
        for (int e = 0; e < 4; ++e) { acc[t][e] = 0.f; qv[t][e] = (t < n && e < per) ? (TC ? s_q[t * H + hh * d + lane * per + e] : __ldcg(s_q + t * H + hh * d + lane * per + e)) : 0.f; }
      }
      const int last = p0 + n - 1;
      // four keys per pass (j0, j0 + 8, j0 + 16, j0 + 24): their K / V rows (one vector load per lane and row) are all in flight before
      // the first dot product.  ld.global.cg: the rows of this launch's tokens were written by other CTAs of the cluster a barrier ago.
      auto load_row = [&](const bf16* base, int j, float (&out)[4]) {
        const bf16* ptr = base + (size_t)j * H + hh * d + lane * per;
        if (per == 4) {
          const uint2 u = __ldcg(reinterpret_cast<const uint2*>(ptr));
          const float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y);
          out[0] = a.x; out[1] = a.y; out[2] = b.x; out[3] = b.y;
        } else if (per == 2) {
          const float2 a = unpack_bf16x2(__ldcg(reinterpret_cast<const uint32_t*>(ptr)));
          out[0] = a.x; out[1] = a.y; out[2] = 0.f; out[3] = 0.f;
        } else {
          out[0] = __bfloat162float(__ushort_as_bfloat16(__ldcg(reinterpret_cast<const unsigned short*>(ptr))));
          out[1] = out[2] = out[3] = 0.f;
        }
      };
      for (int j0 = warp; j0 <= last; j0 += 4 * (DC_THREADS / 32)) {
        float kv[4][4], vv[4][4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int j = j0 + u * (DC_THREADS / 32);
          if (j <= last) { load_row(kc, j, kv[u]); load_row(vc, j, vv[u]); }
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int j = j0 + u * (DC_THREADS / 32);
#pragma unroll
          for (int t = 0; t < DC_NT; ++t) {
            if (t < n && j <= p0 + t) {  // warp-uniform (implies j <= last)
              float sc = 0.f;
#pragma unroll
              for (int e = 0; e < 4; ++e) sc = fmaf(qv[t][e], kv[u][e], sc);
              sc = warp_sum(sc) * scale_log2;
              const float mn = fmaxf(m[t], sc), alpha = fast_exp2(m[t] - mn), pe = fast_exp2(sc - mn);
              lsum[t] = lsum[t] * alpha + pe;
#pragma unroll
              for (int e = 0; e < 4; ++e) acc[t][e] = acc[t][e] * alpha + pe * vv[u][e];
              m[t] = mn;
            }
          }
        }
      }
      // merge the 8 warps: red[(w * NT + t) * (d + 2)] = m, l, acc[d]
      __syncthreads();
#pragma unroll
      for (int t = 0; t < DC_NT; ++t) {
        float* r = red + ((size_t)warp * DC_NT + t) * (d + 2);
        if (lane == 0) { r[0] = m[t]; r[1] = lsum[t]; }
#pragma unroll
        for (int e = 0; e < 4; ++e)
          if (e < per) r[2 + lane * per + e] = acc[t][e];
      }
      __syncthreads();
      for (int i = threadIdx.x; i < n * d; i += DC_THREADS) {
        const int t = i / d, e = i - t * d;
        float M = -INFINITY;
        for (int w = 0; w < DC_THREADS / 32; ++w) M = fmaxf(M, red[((size_t)w * DC_NT + t) * (d + 2)]);
        float L = 0.f, o = 0.f;
        for (int w = 0; w < DC_THREADS / 32; ++w) {
          const float* r = red + ((size_t)w * DC_NT + t) * (d + 2);
          const float f = r[0] == -INFINITY ? 0.f : fast_exp2(r[0] - M);
          L += r[1] * f;
          o += r[2 + e] * f;
        }
        s_o[t * H + hh * d + e] = dc_round_bf16(L > 0.f ? o / L : 0.f);
      }
      __syncthreads();
    }
    if (!TC) dc_cluster_sync();
    dc_stamp(p, sk);  // +1: attention done
    // ---- C: attn.c_proj + residual (HF:217-218, :281): this CTA's columns of x ---------------------------------------------------------
    if (TC) {
      dt_gemv(tc, p, rank, 1, hb, n, [&](int c, const float (&o)[DC_NT]) {
        const float b = pl[5 * H + c];
        for (int t = 0; t < n; ++t) x[t * H + c] += o[t] + b;  // residual stream updated in place by the owner of column c
      });
    } else {
      for (int i = threadIdx.x; i < n * H; i += DC_THREADS) hb[i] = __ldcg(s_o + i);
      __syncthreads();
      const int ncols = H / DC_CTAS, c0 = rank * ncols;
      dc_gemv(st, p, rank, slots, full, consumed, 1, hb, n, red);
      for (int i = threadIdx.x; i < n * ncols; i += DC_THREADS) {
        const int t = i / ncols, c = c0 + (i - t * ncols);
        s_x1[t * H + c] = x[t * H + c] + red[i] + __ldg(P + p.l_off[L_BO] + lo + c);
      }
      dc_cluster_sync();
      for (int i = threadIdx.x; i < n * H; i += DC_THREADS) x[i] = __ldcg(s_x1 + i);
      __syncthreads();
    }
    dc_stamp(p, sk);  // +2: attn.c_proj done
    // ---- D: ln_2 + mlp.c_fc + activation (HF:303, :234-235): this CTA's columns of the hidden layer ------------------------------------
    // ---- F: mlp.c_proj + residual (HF:236-237, :306) -----------------------------------------------------------------------------------
    if (TC) {
      dc_layernorm<true>(x, hb, n, H, pl + 6 * H, pl + 7 * H, p.eps);
      dt_gemv(tc, p, rank, 2, hb, n, [&](int c, const float (&o)[DC_NT]) {  // hb (the LayerNorm output) is free again: dt_gemv works on its operand image
        const float b = pl[8 * H + c];
        for (int t = 0; t < n; ++t) hb[t * I + c] = dc_round_bf16(dc_act(o[t] + b, p.act));
      });
      dc_stamp(p, sk);  // +3: mlp.c_fc done
      dt_gemv(tc, p, rank, 3, hb, n, [&](int c, const float (&o)[DC_NT]) {
        const float b = pl[8 * H + I + c];
        for (int t = 0; t < n; ++t) x[t * H + c] += o[t] + b;
      });
    } else {
      dc_layernorm(x, hb, n, H, P + p.l_off[L_LN2W] + lo, P + p.l_off[L_LN2B] + lo, p.eps);
      {
        const int ncols = I / DC_CTAS, c0 = rank * ncols;
        dc_gemv(st, p, rank, slots, full, consumed, 2, hb, n, red);
        for (int i = threadIdx.x; i < n * ncols; i += DC_THREADS) {
          const int t = i / ncols, c = c0 + (i - t * ncols);
          s_fc[t * I + c] = dc_round_bf16(dc_act(red[i] + __ldg(P + p.l_off[L_B1] + lo + c), p.act));
        }
      }
      dc_cluster_sync();
      dc_stamp(p, sk);  // +3: mlp.c_fc done
      for (int i = threadIdx.x; i < n * I; i += DC_THREADS) hb[i] = __ldcg(s_fc + i);
      __syncthreads();
      {
        const int ncols = H / DC_CTAS, c0 = rank * ncols;
        dc_gemv(st, p, rank, slots, full, consumed, 3, hb, n, red);
        for (int i = threadIdx.x; i < n * ncols; i += DC_THREADS) {
          const int t = i / ncols, c = c0 + (i - t * ncols);
          s_x2[t * H + c] = x[t * H + c] + red[i] + __ldg(P + p.l_off[L_B2] + lo + c);
        }
      }
      dc_cluster_sync();
      for (int i = threadIdx.x; i < n * H; i += DC_THREADS) x[i] = __ldcg(s_x2 + i);
      __syncthreads();
    }
    dc_stamp(p, sk);  // +4: mlp.c_proj done
  }

  dc_stamp(p, sk);  // after every layer's +4 below: all layers done
  // ---- ln_f + predict_action on the last token when it is a state token (HF:450, :652) ------------------------------------------------
  const int last_tok = p0 + n - 1;
  if (rank == 0 && last_tok % 3 == 1 && p.action_out) {
    dc_layernorm(x + (n - 1) * H, hb, 1, H, P + p.t_off[T_LNFW], P + p.t_off[T_LNFB], p.eps);
    for (int a = warp; a < p.ad; a += DC_THREADS / 32) {
      const float* w = P + p.t_off[T_WPA] + (size_t)a * H;
      float s = 0.f;
      for (int c = lane; c < H; c += 32) s = fmaf(__ldg(w + c), hb[c], s);
      s = warp_sum(s) + __ldg(P + p.t_off[T_BPA] + a);
      if (lane == 0) p.action_out[a] = p.tanh_act ? tanhf(s) : s;
    }
  }
  if (p.t_dev && rank == 0 && threadIdx.x == 0) *p.t_dev = hbase + 3;  // t + 1: the next step's count
  dc_stamp(p, sk);  // last: head done
  dc_cluster_sync();  // no CTA leaves while a peer may still be reading what it wrote (and: one completion point for the next launch)
  if (TC) {
    tc_fence_before();
    __syncthreads();
    if (warp == 2) {
      tc_fence_after();
      tmem_dealloc(tc.tmem, DT_TMEM_COLS);
    }
  }
}

}  // namespace dtb

using namespace dtb;

extern "C" {

// Tensor-core path (dt_gemv): hidden a multiple of 128, one CTA per 128 hidden columns — every CTA then owns 384 / 128 / (I / C) / 128 columns
// of the four matrices, whole 128-feature MMA tiles.  DTB_DECODE_FFMA=1 forces the CUDA-core kernel (A/B measurements, tests).
static bool dc_use_tc(const dtb_dt_config& c) {
  static int force_ffma = -1;
  if (force_ffma < 0) {
    const char* e = getenv("DTB_DECODE_FFMA");
    force_ffma = (e && e[0] == '1') ? 1 : 0;
  }
  if (force_ffma || c.hidden != 128) return false;  // one CTA (C = 1): the exchange between the phases stays in shared memory
  const int C = c.hidden / 128;
  return C <= DC_MAX_CTAS && c.n_inner % (128 * C) == 0 && c.n_inner / C <= 128 * (DT_TMEM_COLS / DT_NTOK) && c.hidden / c.n_head <= 128;
}
// Cluster size.  Tensor-core path: hidden / 128.  CUDA-core path: the largest power of two <= 16 such that a CTA streams at most ~0.5 MB
// of weights per layer, and every column slice (H / C, 3H / C, I / C) is a whole number of 8-column groups.
static int dc_cluster_size(const dtb_dt_config& c) {
  if (dc_use_tc(c)) return c.hidden / 128;
  const double layer_bytes = 2.0 * (4.0 * c.hidden * c.hidden + 2.0 * c.hidden * c.n_inner);
  int C = 1;
  while (C < DC_MAX_CTAS && layer_bytes / C > 512.0 * 1024.0 && c.hidden % (16 * C) == 0 && c.n_inner % (16 * C) == 0) C *= 2;
  return C;
}
// tensor-core path: dynamic shared memory = x | hb | red (tc_red_floats) | pad to 1 KB | B-operand image | ring | barriers
static int dc_tc_red_floats(const dtb_dt_config& c) {
  const int C = c.hidden / 128, d = c.hidden / c.n_head;
  int widest = 3 * c.hidden / C;
  if (c.n_inner / C > widest) widest = c.n_inner / C;
  int r = DC_NT * widest, a = (DC_THREADS / 32) * DC_NT * (d + 2);
  return ((r > a ? r : a) + 63) & ~63;
}
static size_t dc_tc_smem(const dtb_dt_config& c, int* slots_out) {
  const int KI = c.hidden > c.n_inner ? c.hidden : c.n_inner;
  const size_t fixed = ((size_t)DC_NT * c.hidden + (size_t)DC_NT * KI + dc_tc_red_floats(c) + 13 * c.hidden + DC_NT * c.n_inner + 9 * c.hidden + c.n_inner) * sizeof(float) +
                       1024 + (size_t)(KI / 64) * 2048 +
                       (2 * DT_MAX_SLOTS + 1) * sizeof(uint64_t) + 64;
  const size_t limit = 227 * 1024;
  int slots = fixed + DT_BLOCK_BYTES <= limit ? (int)((limit - fixed) / DT_BLOCK_BYTES) : 0;
  if (slots > DT_MAX_SLOTS) slots = DT_MAX_SLOTS;
  while (slots & (slots - 1)) --slots;  // a power of two (the kernel masks instead of dividing)
  if (slots_out) *slots_out = slots;
  return fixed + (size_t)slots * DT_BLOCK_BYTES;
}
static size_t dc_layer_elems(const dtb_dt_config& c) { return (size_t)4 * c.hidden * c.hidden + (size_t)2 * c.hidden * c.n_inner; }
// byte offset of the decode-layout weights inside the cache buffer: after the K/V cache and the exchange scratch
static size_t dc_wd_offset(const dtb_dt_config& c, int K) {
  const size_t kv = ((size_t)2 * c.n_layer * 3 * K * c.hidden * sizeof(bf16) + 255) & ~size_t(255);
  return kv + ((dc_scratch_floats(c.hidden, c.n_inner) * sizeof(float) + 255) & ~size_t(255));
}
int dtb_dt_decode_supported(const dtb_dt* h) {
  if (!h) return 0;
  const dtb_dt_config& c = h->c;
  const int C = dc_cluster_size(c);
  const int widest = (3 * c.hidden > c.n_inner ? 3 * c.hidden : c.n_inner) / C;
  if (dc_use_tc(c)) {
    int slots = 0;
    dc_tc_smem(c, &slots);
    return slots >= 2 && (h->sd_small || c.state_dim % 8 == 0) ? 2 : 0;  // 2 = the tensor-core kernel (the host's automatic choice)
  }
  return c.hidden % 8 == 0 && c.hidden <= DC_MAXH && c.n_inner % 8 == 0 && widest <= 1536 && (h->sd_small || (c.state_dim % 8 == 0 && c.hidden % C == 0)) ? 1 : 0;
}

// bytes of the decode state: key + value cache [L][3 K][H] bf16 each, then the exchange scratch
size_t dtb_dt_decode_cache_bytes(const dtb_dt* h, int K) {
  if (!h || K < 1) return 0;
  const dtb_dt_config& c = h->c;
  return dc_wd_offset(c, K) + dc_layer_elems(c) * c.n_layer * sizeof(bf16) + 256;
}

static void dc_fill(const dtb_dt* h, DecodeParams& p, const float* params, const void* params_bf16, void* cache, int K) {
  const dtb_dt_config& c = h->c;
  memset(&p, 0, sizeof(p));
  p.H = c.hidden; p.I = c.n_inner; p.L = c.n_layer; p.nh = c.n_head; p.d = c.hidden / c.n_head; p.sd = c.state_dim; p.ad = c.act_dim;
  p.E = c.max_ep_len; p.act = c.activation; p.tanh_act = c.action_tanh; p.sd_small = h->sd_small; p.eps = c.ln_eps;
  p.P = params; p.Pb = reinterpret_cast<const bf16*>(params_bf16);
  for (int i = 0; i < E_COUNT; ++i) p.e_off[i] = h->e(i);
  for (int i = 0; i < L_COUNT; ++i) p.l_off[i] = h->l(0, i);
  for (int i = 0; i < T_COUNT; ++i) p.t_off[i] = h->t(i);
  p.layer_stride = c.n_layer > 1 ? h->l(1, 0) - h->l(0, 0) : 0;
  uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(cache) + 255) & ~uintptr_t(255));
  p.Tmax = 3 * K;
  p.kc = reinterpret_cast<bf16*>(base);
  p.vc = p.kc + (size_t)c.n_layer * p.Tmax * c.hidden;
  const size_t kv = ((size_t)2 * c.n_layer * p.Tmax * c.hidden * sizeof(bf16) + 255) & ~size_t(255);
  p.scratch = reinterpret_cast<float*>(base + kv);
  p.Wd = p.Wd_out = reinterpret_cast<bf16*>(base + dc_wd_offset(c, K));
  p.C = dc_cluster_size(c);
}

// (Re)build the decode layout of the weights inside `cache` from the bf16 shadow: call once, and again whenever the parameters changed.
int dtb_dt_decode_stage(const dtb_dt* h, void* stream, const void* params_bf16, int K, void* cache, size_t cache_bytes) {
  DTB_REQUIRE(h && params_bf16 && cache, DTB_ERR_BAD_ARG, "dtb_dt_decode_stage: null argument");
  DTB_REQUIRE(dtb_dt_decode_supported(h), DTB_ERR_UNSUPPORTED, "dtb_dt_decode_stage: hidden %d / n_inner %d unsupported", h->c.hidden, h->c.n_inner);
  DTB_REQUIRE(cache_bytes >= dtb_dt_decode_cache_bytes(h, K), DTB_ERR_WORKSPACE, "dtb_dt_decode_stage: cache too small");
  DecodeParams p;
  dc_fill(h, p, nullptr, params_bf16, cache, K);
  const size_t chunks = dc_layer_elems(h->c) * h->c.n_layer / 8;
  size_t blocks = (chunks + 255) / 256;
  if (blocks > (size_t)num_sms() * 8) blocks = (size_t)num_sms() * 8;
  if (dc_use_tc(h->c)) DTB_LAUNCH(dt_decode_stage_tc_kernel, (unsigned)blocks, 256, 0, (cudaStream_t)stream, p, dc_layer_elems(h->c));
  else DTB_LAUNCH(dt_decode_stage_kernel, (unsigned)blocks, 256, 0, (cudaStream_t)stream, p, dc_layer_elems(h->c));
  DTB_CHECK_LAUNCH();
  return DTB_OK;
}

static int dc_launch(const dtb_dt* h, cudaStream_t stream, DecodeParams& p, int t, int t_cached) {
  const dtb_dt_config& c = h->c;
  {
    static int stamp = -1;
    if (stamp < 0) { const char* e = getenv("DTB_DECODE_STAMP"); stamp = (e && e[0] == '1') ? 1 : 0; }
    if (stamp && dc_use_tc(c)) p.stamps = reinterpret_cast<unsigned long long*>(p.scratch);  // the L2 exchange scratch is unused on this path
  }
  const int KI = c.hidden > c.n_inner ? c.hidden : c.n_inner;
  const bool tc = dc_use_tc(c);
  size_t smem = ((size_t)DC_NT * c.hidden + (size_t)DC_NT * KI + DC_RED_FLOATS) * sizeof(float) + (size_t)DC_SLOTS * DC_SLOT_BYTES + 64;
  if (tc) {
    smem = dc_tc_smem(c, &p.tc_slots);
    p.tc_red_floats = dc_tc_red_floats(c);
  }
  static size_t attr[2] = {0, 0};
  if (smem > attr[tc ? 1 : 0]) {
    if (tc) {
      DTB_CUDA(cudaFuncSetAttribute(dt_decode_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      DTB_CUDA(cudaFuncSetAttribute(dt_decode_kernel<true>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    } else {
      DTB_CUDA(cudaFuncSetAttribute(dt_decode_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      DTB_CUDA(cudaFuncSetAttribute(dt_decode_kernel<false>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    }
    attr[tc ? 1 : 0] = smem;
  }
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(p.C); cfg.blockDim = dim3(DC_THREADS); cfg.dynamicSmemBytes = smem; cfg.stream = stream;
  cudaLaunchAttribute at[2];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = p.C; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = pdl_enabled() ? 2 : 1;
  if (p.t_dev) {  // step mode: one launch, the token range comes from the device-side count
    p.p0 = 0; p.n = 0;
    if (tc) DTB_CUDA(cudaLaunchKernelEx(&cfg, dt_decode_kernel<true>, p));
    else DTB_CUDA(cudaLaunchKernelEx(&cfg, dt_decode_kernel<false>, p));
    DTB_CHECK_LAUNCH();
    return DTB_OK;
  }
  // tokens [p0, p1): the action token of the last cached timestep (its action was a placeholder when it was cached), then R, s of every
  // new timestep; the newest timestep's own action token comes after its state token and is not needed for the prediction
  const int p1 = 3 * t - 1;
  for (int p0 = t_cached > 0 ? 3 * t_cached - 1 : 0; p0 < p1; p0 += DC_NT) {
    p.p0 = p0;
    p.n = p1 - p0 < DC_NT ? p1 - p0 : DC_NT;
    if (tc) DTB_CUDA(cudaLaunchKernelEx(&cfg, dt_decode_kernel<true>, p));
    else DTB_CUDA(cudaLaunchKernelEx(&cfg, dt_decode_kernel<false>, p));
    DTB_CHECK_LAUNCH();
  }
  return DTB_OK;
}

int dtb_dt_decode(const dtb_dt* h, void* stream, const float* params, const void* params_bf16, const float* states, const void* states_bf16,
                  const float* actions, const float* returns_to_go, const long long* timesteps, int t, int t_cached, int K, void* cache,
                  size_t cache_bytes, float* action_out) {
  DTB_REQUIRE(h && params && params_bf16 && actions && returns_to_go && timesteps && cache && action_out, DTB_ERR_BAD_ARG, "dtb_dt_decode: null argument");
  DTB_REQUIRE(dtb_dt_decode_supported(h), DTB_ERR_UNSUPPORTED, "dtb_dt_decode: hidden %d / n_inner %d unsupported", h->c.hidden, h->c.n_inner);
  DTB_REQUIRE(t >= 1 && t <= K && t_cached >= 0 && t_cached < t, DTB_ERR_BAD_ARG, "dtb_dt_decode: need 0 <= t_cached (%d) < t (%d) <= K (%d)", t_cached, t, K);
  DTB_REQUIRE(h->sd_small ? states != nullptr : states_bf16 != nullptr, DTB_ERR_BAD_ARG, "dtb_dt_decode: states (fp32, state_dim <= 64) or states_bf16 (wide) required");
  DTB_REQUIRE(cache_bytes >= dtb_dt_decode_cache_bytes(h, K), DTB_ERR_WORKSPACE, "dtb_dt_decode: cache too small");
  DTB_TRY(require_sm100());
  DecodeParams p;
  dc_fill(h, p, params, params_bf16, cache, K);
  p.states = states; p.states_bf16 = reinterpret_cast<const bf16*>(states_bf16); p.actions = actions; p.rtg = returns_to_go; p.timesteps = timesteps;
  p.action_out = action_out;
  return dc_launch(h, (cudaStream_t)stream, p, t, t_cached);
}

int dtb_dt_decode_step(const dtb_dt* h, void* stream, const float* params, const void* params_bf16, const float* states2, const void* state_new_bf16,
                       const float* actions2, const float* returns_to_go2, const long long* timesteps2, int* t_dev, int K, void* cache,
                       size_t cache_bytes, float* action_out) {
  DTB_REQUIRE(h && params && params_bf16 && actions2 && returns_to_go2 && timesteps2 && t_dev && cache && action_out, DTB_ERR_BAD_ARG,
              "dtb_dt_decode_step: null argument");
  DTB_REQUIRE(dtb_dt_decode_supported(h), DTB_ERR_UNSUPPORTED, "dtb_dt_decode_step: hidden %d / n_inner %d unsupported", h->c.hidden, h->c.n_inner);
  DTB_REQUIRE(h->sd_small ? states2 != nullptr : state_new_bf16 != nullptr, DTB_ERR_BAD_ARG,
              "dtb_dt_decode_step: states2 (fp32 [2, state_dim], state_dim <= 64) or state_new_bf16 (wide, [state_dim]) required");
  DTB_REQUIRE(cache_bytes >= dtb_dt_decode_cache_bytes(h, K), DTB_ERR_WORKSPACE, "dtb_dt_decode_step: cache too small");
  DTB_TRY(require_sm100());
  DecodeParams p;
  dc_fill(h, p, params, params_bf16, cache, K);
  p.states = states2; p.state_new = reinterpret_cast<const bf16*>(state_new_bf16); p.actions = actions2; p.rtg = returns_to_go2; p.timesteps = timesteps2;
  p.t_dev = t_dev;
  p.action_out = action_out;
  return dc_launch(h, (cudaStream_t)stream, p, 0, 0);
}

}  // extern "C"
