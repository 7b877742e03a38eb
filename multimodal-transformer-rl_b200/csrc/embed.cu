// Fused token embedding for the interleaved (return-to-go, state, action) stream — HF:611-628 (A2).
//   te  = embed_timestep[timesteps[b,k]]                       (integer gather, bit-exact index rule)
//   e_R = rtg * wR + bR + te ; e_s = Ws s + bs + te ; e_a = Wa a + ba + te
//   x[b, 3k+0] = LN(e_R), x[b, 3k+1] = LN(e_s), x[b, 3k+2] = LN(e_a)     (embed_ln, HF:628)
// One warp per timestep, lanes own float4 column groups (128-bit loads/stores), transposed weights in
// shared memory; LayerNorm in registers (two-pass, fp32).  Training mode also saves the pre-LN rows, the
// LN statistics and a packed bf16 copy U = [rtg | actions | states] that the weight-gradient GEMM reads.
// When state_dim is large (multimodal towers: 512/1024) the state projection comes from the tcgen05
// GEMM (`state_proj`, fp32 [B*K, H], no bias) and only bias + te + LN are applied here.
#include "kernels.h"

namespace dtb {

constexpr int EMB_MAX_VEC = 8;  // H <= 1024

struct EmbedFwdArgs {
  const float* states;      // [BK, sd] or null when state_proj is given
  const float* state_proj;  // [BK, H] or null
  const float* actions;     // [BK, ad]
  const float* rtg;         // [BK]
  const long long* timesteps;
  const float* table;       // [E, H]
  const float* wR; const float* bR;
  const float* Ws; const float* bs;   // Ws [H, sd]
  const float* Wa; const float* ba;   // Wa [H, ad]
  const float* gamma; const float* beta;
  int BK, sd, ad, H, E;
  float eps;
  float* x;      // [BK*3, H]
  float* pre;    // [BK*3, H] or null
  float* mean;   // [BK*3] or null
  float* rstd;   // [BK*3] or null
  bf16* U;       // [BK, P8] or null
  int P8;
  int u_states;  // 1: U also carries the states (small-state path)
  DropSite drop; // embd dropout on the LayerNorm output (GPT2Model.drop; idx = row * H + col)
  // optional: ln_1 of the first block fused behind embed_ln + dropout (the warp still holds the row): h1 = LN(x) in bf16 + its statistics
  const float* ln1_gamma; const float* ln1_beta;
  bf16* h1;
  float* mean1; float* rstd1;
};

__global__ void __launch_bounds__(256) embed_fwd_kernel(const EmbedFwdArgs a) {
  pdl_trigger();
  pdl_wait();
  extern __shared__ __align__(16) float sm_emb[];
  const uint32_t drop_key = a.drop.state ? drop_site_key(a.drop) : 0u;
  const int H = a.H, sd = a.states ? a.sd : 0, ad = a.ad;
  float* sWs = sm_emb;                 // [sd][H]  (transposed)
  float* sWa = sWs + (size_t)sd * H;   // [ad][H]
  float* sVec = sWa + (size_t)ad * H;  // wR, bR, bs, ba, gamma, beta : [6][H]
  for (int i = threadIdx.x; i < sd * H; i += blockDim.x) {
    const int j = i / H, c = i - j * H;
    sWs[i] = a.Ws[(size_t)c * a.sd + j];
  }
  for (int i = threadIdx.x; i < ad * H; i += blockDim.x) {
    const int j = i / H, c = i - j * H;
    sWa[i] = a.Wa[(size_t)c * ad + j];
  }
  for (int i = threadIdx.x; i < H; i += blockDim.x) {
    sVec[i] = a.wR[i];
    sVec[H + i] = a.bR[i];
    sVec[2 * H + i] = a.bs[i];
    sVec[3 * H + i] = a.ba[i];
    sVec[4 * H + i] = a.gamma[i];
    sVec[5 * H + i] = a.beta[i];
  }
  __syncthreads();

  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  for (int t = blockIdx.x * nwarp + warp; t < a.BK; t += gridDim.x * nwarp) {
    long long ts = a.timesteps[t];
    ts = ts < 0 ? 0 : (ts >= a.E ? a.E - 1 : ts);  // memory safety only; in-range indices are gathered exactly
    const float* te = a.table + (size_t)ts * H;
    const float r = a.rtg[t];
    // inputs: lane j holds element j and j+32
    float s0 = 0.f, s1 = 0.f, a0 = 0.f, a1 = 0.f;
    if (sd) {
      if (lane < sd) s0 = a.states[(size_t)t * a.sd + lane];
      if (lane + 32 < sd) s1 = a.states[(size_t)t * a.sd + lane + 32];
    }
    if (lane < ad) a0 = a.actions[(size_t)t * ad + lane];
    if (lane + 32 < ad) a1 = a.actions[(size_t)t * ad + lane + 32];
    if (a.U) {
      bf16* u = a.U + (size_t)t * a.P8;
      for (int j = lane; j < a.P8; j += 32) {
        float v = 0.f;
        if (j == 0) v = r;
        else if (j <= ad) v = a.actions[(size_t)t * ad + (j - 1)];
        else if (a.u_states && j < 1 + ad + a.sd) v = a.states[(size_t)t * a.sd + (j - 1 - ad)];
        u[j] = __float2bfloat16(v);
      }
    }
#pragma unroll
    for (int tok = 0; tok < 3; ++tok) {
      float4 v[EMB_MAX_VEC];
      // bias + timestep embedding (+ rtg term / GEMM-projected state)
#pragma unroll
      for (int i = 0; i < EMB_MAX_VEC; ++i) {
        const int c = i * 128 + lane * 4;
        v[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (c < H) {
          float4 acc = ldg_f4(te + c);
          const float4 b = *reinterpret_cast<const float4*>(sVec + (tok == 0 ? 1 : tok + 1) * H + c);
          acc.x += b.x; acc.y += b.y; acc.z += b.z; acc.w += b.w;
          if (tok == 0) {
            const float4 w = *reinterpret_cast<const float4*>(sVec + c);
            acc.x += r * w.x; acc.y += r * w.y; acc.z += r * w.z; acc.w += r * w.w;
          } else if (tok == 1 && a.state_proj) {
            const float4 p = ldg_f4(a.state_proj + (size_t)t * H + c);
            acc.x += p.x; acc.y += p.y; acc.z += p.z; acc.w += p.w;
          }
          v[i] = acc;
        }
      }
      // small projections: input element j broadcast by shuffle (warp-uniform), weights from shared memory
      if (tok != 0) {
        const int nin = tok == 1 ? sd : ad;
        const float* sW = tok == 1 ? sWs : sWa;
        const float lo = tok == 1 ? s0 : a0, hi = tok == 1 ? s1 : a1;
        for (int j = 0; j < nin; ++j) {
          const float xj = __shfl_sync(0xffffffffu, j < 32 ? lo : hi, j & 31);
#pragma unroll
          for (int i = 0; i < EMB_MAX_VEC; ++i) {
            const int c = i * 128 + lane * 4;
            if (c < H) {
              const float4 w = *reinterpret_cast<const float4*>(sW + (size_t)j * H + c);
              v[i].x += xj * w.x; v[i].y += xj * w.y; v[i].z += xj * w.z; v[i].w += xj * w.w;
            }
          }
        }
      }
      float sum = 0.f;
#pragma unroll
      for (int i = 0; i < EMB_MAX_VEC; ++i) {
        const int c = i * 128 + lane * 4;
        if (c < H) sum += v[i].x + v[i].y + v[i].z + v[i].w;
      }
      const float mean = warp_sum(sum) / H;
      float q = 0.f;
#pragma unroll
      for (int i = 0; i < EMB_MAX_VEC; ++i) {
        const int c = i * 128 + lane * 4;
        if (c < H) {
          const float d0 = v[i].x - mean, d1 = v[i].y - mean, d2 = v[i].z - mean, d3 = v[i].w - mean;
          q += d0 * d0 + d1 * d1 + d2 * d2 + d3 * d3;
        }
      }
      const float rstd = rsqrtf(warp_sum(q) / H + a.eps);
      const size_t row = (size_t)t * 3 + tok;  // token order R_k, s_k, a_k (HF:623-627)
      if (lane == 0) {
        if (a.mean) a.mean[row] = mean;
        if (a.rstd) a.rstd[row] = rstd;
      }
#pragma unroll
      for (int i = 0; i < EMB_MAX_VEC; ++i) {
        const int c = i * 128 + lane * 4;
        if (c < H) {
          if (a.pre) *reinterpret_cast<float4*>(a.pre + row * H + c) = v[i];
          const float4 g = *reinterpret_cast<const float4*>(sVec + 4 * H + c), b = *reinterpret_cast<const float4*>(sVec + 5 * H + c);
          float4 o;
          o.x = (v[i].x - mean) * rstd * g.x + b.x;
          o.y = (v[i].y - mean) * rstd * g.y + b.y;
          o.z = (v[i].z - mean) * rstd * g.z + b.z;
          o.w = (v[i].w - mean) * rstd * g.w + b.w;
          if (a.drop.state) {
            const uint32_t k = drop_key + (uint32_t)row * (uint32_t)H + (uint32_t)c;
            o.x = drop_apply(o.x, k, 0u, a.drop.thr, a.drop.scale); o.y = drop_apply(o.y, k, 1u, a.drop.thr, a.drop.scale);
            o.z = drop_apply(o.z, k, 2u, a.drop.thr, a.drop.scale); o.w = drop_apply(o.w, k, 3u, a.drop.thr, a.drop.scale);
          }
          *reinterpret_cast<float4*>(a.x + row * H + c) = o;
          v[i] = o;
        }
      }
      if (a.h1) {  // ln_1 of block 0 (HF:272) on the row just produced: same two-pass statistics as ln_fwd_kernel
        float s2 = 0.f;
#pragma unroll
        for (int i = 0; i < EMB_MAX_VEC; ++i) {
          const int c = i * 128 + lane * 4;
          if (c < H) s2 += v[i].x + v[i].y + v[i].z + v[i].w;
        }
        const float m2 = warp_sum(s2) / H;
        float q2 = 0.f;
#pragma unroll
        for (int i = 0; i < EMB_MAX_VEC; ++i) {
          const int c = i * 128 + lane * 4;
          if (c < H) {
            const float d0 = v[i].x - m2, d1 = v[i].y - m2, d2 = v[i].z - m2, d3 = v[i].w - m2;
            q2 += d0 * d0 + d1 * d1 + d2 * d2 + d3 * d3;
          }
        }
        const float r2 = rsqrtf(warp_sum(q2) / H + a.eps);
        if (lane == 0) {
          if (a.mean1) a.mean1[row] = m2;
          if (a.rstd1) a.rstd1[row] = r2;
        }
#pragma unroll
        for (int i = 0; i < EMB_MAX_VEC; ++i) {
          const int c = i * 128 + lane * 4;
          if (c < H) {
            const float4 g = ldg_f4(a.ln1_gamma + c), b = ldg_f4(a.ln1_beta + c);
            *reinterpret_cast<uint2*>(a.h1 + row * H + c) =
                make_uint2(pack_bf16x2((v[i].x - m2) * r2 * g.x + b.x, (v[i].y - m2) * r2 * g.y + b.y),
                           pack_bf16x2((v[i].z - m2) * r2 * g.z + b.z, (v[i].w - m2) * r2 * g.w + b.w));
          }
        }
      }
    }
  }
}

int embed_fwd(cudaStream_t stream, const float* states, const float* state_proj, const float* actions, const float* rtg,
              const long long* timesteps, const float* table, const float* wR, const float* bR, const float* Ws, const float* bs,
              const float* Wa, const float* ba, const float* gamma, const float* beta, int BK, int sd, int ad, int H, int E, float eps,
              float* x, float* pre, float* mean, float* rstd, bf16* U, int P8, int u_states, const DropSite& drop, const float* ln1_gamma,
              const float* ln1_beta, bf16* h1, float* mean1, float* rstd1) {
  DTB_REQUIRE(H % 4 == 0 && H <= EMB_MAX_VEC * 128, DTB_ERR_UNSUPPORTED, "embed: hidden size %d unsupported (multiple of 4, <= 1024)", H);
  DTB_REQUIRE(ad >= 1 && ad <= 64, DTB_ERR_UNSUPPORTED, "embed: act_dim %d unsupported (1..64)", ad);
  DTB_REQUIRE((states != nullptr) != (state_proj != nullptr), DTB_ERR_BAD_ARG, "embed: exactly one of states / state_proj must be given");
  DTB_REQUIRE(states == nullptr || sd <= 64, DTB_ERR_UNSUPPORTED, "embed: fused state projection needs state_dim <= 64 (got %d)", sd);
  if (BK == 0) return DTB_OK;
  EmbedFwdArgs a{states, state_proj, actions, rtg, timesteps, table, wR, bR, Ws, bs, Wa, ba, gamma, beta,
                 BK, sd, ad, H, E, eps, x, pre, mean, rstd, U, P8, u_states, drop, ln1_gamma, ln1_beta, h1, mean1, rstd1};
  const size_t smem = ((size_t)(states ? sd : 0) + ad + 6) * H * sizeof(float);
  DTB_REQUIRE(smem <= 200 * 1024, DTB_ERR_UNSUPPORTED, "embed: (state_dim + act_dim + 6) * hidden too large for shared memory");
  static bool attr_set = false;
  if (smem > 48 * 1024 && !attr_set) {
    DTB_CUDA(cudaFuncSetAttribute(embed_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    attr_set = true;
  }
  int blocks = (BK + 7) / 8;
  const int cap = num_sms() * 2;
  if (blocks > cap) blocks = cap;
  ProfScope prof(stream, PF_EMBED_FWD, (pre ? 8.0 : 4.0) * BK * 3 * H + 4.0 * BK * (sd + ad + 3));
  DTB_LAUNCH(embed_fwd_kernel, blocks, 256, smem, stream, a);
  DTB_CHECK_LAUNCH();
  return DTB_OK;
}

// ---- backward pieces -----------------------------------------------------------------------------------------------
// d_pre: bf16 [BK, 3H] (row t = timestep, columns R | s | a) = gradient w.r.t. the pre-LN embeddings (from ln_bwd).
// d_table[e] = sum over the timesteps t with timesteps[t] == e, in ascending t, of d_pre_R + d_pre_s + d_pre_a.
// One warp per table row, eight rows per CTA: the CTA stages the (clamped) timestep vector through shared memory in chunks, every warp
// scans the chunk for its row (32 indices per shared-memory load, matches found by ballot) and adds the matching rows in index order —
// no atomics, so the embedding gradient is bit-reproducible run to run (SURVEY.md §8b "Determinism"), and every table row is written
// exactly once (rows nobody gathered get zeros: no memset pass).  (A first version scanned global memory directly: one dependent
// L2 round trip per 32 timesteps made it 170 us at c2 — on the tail of the backward.)
constexpr int TE_MAX_PAIRS = 16;  // H <= 1024: 16 bf16 pairs per lane
constexpr int TE_CHUNK = 4096;    // timesteps staged per pass (16 KB of shared memory)
__global__ void __launch_bounds__(256) embed_te_scatter_kernel(const bf16* __restrict__ d_pre, const long long* __restrict__ timesteps,
                                                               const float* __restrict__ mask, int BK, int H, int E,
                                                               float* __restrict__ d_table) {
  pdl_trigger();
  pdl_wait();
  __shared__ int s_ts[TE_CHUNK];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  const int e = blockIdx.x * nwarp + warp;  // grid covers all E rows (ceil(E / 8) blocks)
  float2 acc[TE_MAX_PAIRS];
#pragma unroll
  for (int i = 0; i < TE_MAX_PAIRS; ++i) acc[i] = make_float2(0.f, 0.f);
  for (int c0 = 0; c0 < BK; c0 += TE_CHUNK) {
    const int n = min(TE_CHUNK, BK - c0);
    __syncthreads();
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      long long ts = timesteps[c0 + i];
      ts = ts < 0 ? 0 : (ts >= E ? E - 1 : ts);  // same clamp as the forward gather
      // padded timesteps (mask 0) all carry timestep 0 and an exactly-zero gradient (no loss row, never a key): skipping them is
      // bit-exact and keeps table row 0 from becoming a serial hot spot
      s_ts[i] = (mask && mask[c0 + i] <= 0.f) ? -1 : (int)ts;
    }
    __syncthreads();
    if (e < E) {
      for (int t0 = 0; t0 < n; t0 += 32) {
        const int ts = t0 + lane < n ? s_ts[t0 + lane] : -1;
        unsigned int m = __ballot_sync(0xffffffffu, ts == e);
        while (m) {  // up to four matching rows per pass: all their loads are in flight together, the adds stay in timestep order
          const bf16* p[4];
          int nm = 0;
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            p[q] = d_pre;
            if (m) {
              p[q] = d_pre + (size_t)(c0 + t0 + __ffs(m) - 1) * 3 * H;
              m &= m - 1;
              nm = q + 1;
            }
          }
#pragma unroll
          for (int i = 0; i < TE_MAX_PAIRS; ++i) {
            const int c = i * 64 + lane * 2;
            if (c < H) {
              uint32_t v[4][3];
#pragma unroll
              for (int q = 0; q < 4; ++q)
                if (q < nm) {
                  v[q][0] = __ldg(reinterpret_cast<const uint32_t*>(p[q] + c));
                  v[q][1] = __ldg(reinterpret_cast<const uint32_t*>(p[q] + H + c));
                  v[q][2] = __ldg(reinterpret_cast<const uint32_t*>(p[q] + 2 * H + c));
                }
#pragma unroll
              for (int q = 0; q < 4; ++q)
                if (q < nm) {
                  const float2 r = unpack_bf16x2(v[q][0]), s = unpack_bf16x2(v[q][1]), a = unpack_bf16x2(v[q][2]);
                  acc[i].x += r.x + s.x + a.x;
                  acc[i].y += r.y + s.y + a.y;
                }
            }
          }
        }
      }
    }
  }
  if (e < E) {
#pragma unroll
    for (int i = 0; i < TE_MAX_PAIRS; ++i) {
      const int c = i * 64 + lane * 2;
      if (c < H) *reinterpret_cast<float2*>(d_table + (size_t)e * H + c) = acc[i];
    }
  }
}

int embed_te_scatter(cudaStream_t stream, const bf16* d_pre, const long long* timesteps, const float* mask, int BK, int H, int E,
                     float* d_table) {
  DTB_REQUIRE(H % 2 == 0 && H <= 64 * TE_MAX_PAIRS, DTB_ERR_UNSUPPORTED, "embed_te_scatter: hidden size %d unsupported", H);
  if (BK == 0) {
    DTB_CUDA(cudaMemsetAsync(d_table, 0, (size_t)E * H * sizeof(float), stream));
    return DTB_OK;
  }
  DTB_LAUNCH(embed_te_scatter_kernel, (E + 7) / 8, 256, 0, stream, d_pre, timesteps, mask, BK, H, E, d_table);
  DTB_CHECK_LAUNCH();
  return DTB_OK;
}

// d_states[t, j] = sum_c d_pre_s[t, c] Ws[c, j]   (gradient w.r.t. fp32 vector states, state_dim <= 64; Ws fp32 [H, sd] as nn.Linear stores it)
// One warp per timestep, lanes over j (coalesced Ws rows), the d_pre element broadcast to the warp.
__global__ void __launch_bounds__(256) embed_dstates_kernel(const bf16* __restrict__ d_pre, const float* __restrict__ Ws, int BK, int H, int sd,
                                                            float* __restrict__ d_states) {
  pdl_trigger();
  pdl_wait();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  for (int t = blockIdx.x * nwarp + warp; t < BK; t += gridDim.x * nwarp) {
    const bf16* g = d_pre + (size_t)t * 3 * H + H;  // the state-token row
    float a0 = 0.f, a1 = 0.f;
    for (int c = 0; c < H; ++c) {
      const float gc = __bfloat162float(g[c]);
      if (lane < sd) a0 = fmaf(gc, __ldg(Ws + (size_t)c * sd + lane), a0);
      if (lane + 32 < sd) a1 = fmaf(gc, __ldg(Ws + (size_t)c * sd + lane + 32), a1);
    }
    if (lane < sd) d_states[(size_t)t * sd + lane] = a0;
    if (lane + 32 < sd) d_states[(size_t)t * sd + lane + 32] = a1;
  }
}
int embed_dstates(cudaStream_t stream, const bf16* d_pre, const float* Ws, int BK, int H, int sd, float* d_states) {
  DTB_REQUIRE(sd <= 64, DTB_ERR_UNSUPPORTED, "embed_dstates: state_dim %d > 64 goes through the tensor-core path", sd);
  if (BK == 0) return DTB_OK;
  int blocks = (BK + 7) / 8;
  if (blocks > num_sms() * 4) blocks = num_sms() * 4;
  DTB_LAUNCH(embed_dstates_kernel, blocks, 256, 0, stream, d_pre, Ws, BK, H, sd, d_states);
  DTB_CHECK_LAUNCH();
  return DTB_OK;
}

// G = d_pre^T U : fp32 [3H, P8] (tcgen05 wgrad GEMM); bsum = column sums of d_pre : fp32 [3H].
// Only the diagonal blocks are gradients:  d wR[c] = G[c][0];  d Wa[c][j] = G[2H+c][1+j];  d Ws[c][j] = G[H+c][1+ad+j].
__global__ void embed_unpack_kernel(const float* __restrict__ G, const float* __restrict__ bsum, int H, int sd, int ad, int P8,
                                    int u_states, float* __restrict__ dwR, float* __restrict__ dbR, float* __restrict__ dWs,
                                    float* __restrict__ dbs, float* __restrict__ dWa, float* __restrict__ dba) {
  pdl_trigger();
  pdl_wait();
  const int per_c = 3 + ad + (u_states ? sd : 0);
  const int total = H * per_c;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int c = i / per_c, e = i - c * per_c;
    if (e == 0) dwR[c] = G[(size_t)c * P8];
    else if (e == 1) dbR[c] = bsum[c];
    else if (e == 2) { dbs[c] = bsum[H + c]; dba[c] = bsum[2 * H + c]; }
    else if (e < 3 + ad) dWa[(size_t)c * ad + (e - 3)] = G[(size_t)(2 * H + c) * P8 + 1 + (e - 3)];
    else dWs[(size_t)c * sd + (e - 3 - ad)] = G[(size_t)(H + c) * P8 + 1 + ad + (e - 3 - ad)];
  }
}

int embed_unpack(cudaStream_t stream, const float* G, const float* bsum, int H, int sd, int ad, int P8, int u_states, float* dwR,
                 float* dbR, float* dWs, float* dbs, float* dWa, float* dba) {
  const int total = H * (3 + ad + (u_states ? sd : 0));
  int blocks = (total + 255) / 256;
  if (blocks > num_sms() * 4) blocks = num_sms() * 4;
  DTB_LAUNCH(embed_unpack_kernel, blocks, 256, 0, stream, G, bsum, H, sd, ad, P8, u_states, dwR, dbR, dWs, dbs, dWa, dba);
  DTB_CHECK_LAUNCH();
  return DTB_OK;
}

}  // namespace dtb
