// Prediction heads + masked action-MSE loss (A7, A8, A10) on the ln_f output of the interleaved stream.
//   y = ln_f(x) bf16 [B*K*3, H] (rows 3t+0 = R token, 3t+1 = state token, 3t+2 = action token; HF:650)
//   action_preds[t] = [tanh](Wpa y[3t+1] + b)      (HF:655)      state_preds[t] = Wps y[3t+2] + b   (HF:654)
//   return_preds[t] = wpr . y[3t+2] + b            (HF:653)
//   loss = mean over rows with attention_mask > 0 of (action_preds - target)^2      (upstream train_step)
// Row selection (3t+1 / 3t+2, mask > 0) is integer-exact; arithmetic is fp32.  One warp per timestep, warp-shuffle dots.
#include "kernels.h"

namespace dtb {

constexpr int HD_MAX_VEC = 4;  // lane owns 8 bf16 columns per 256-wide chunk: H <= 1024

struct HeadsFwdArgs {
  const bf16* y;  // [BK*3, H]
  const float* Wa; const float* ba;  // [ad, H], [ad]
  const float* Ws; const float* bs;  // [sd, H], [sd]
  const float* Wr; const float* br;  // [H], [1]
  int BK, H, sd, ad, tanh_act;
  float* action_preds;  // [BK, ad]
  float* state_preds;   // [BK, sd] or null
  float* return_preds;  // [BK] or null
};

__device__ __forceinline__ void load_row8(const bf16* row, int H, int lane, float (&v)[HD_MAX_VEC][8]) {
#pragma unroll
  for (int i = 0; i < HD_MAX_VEC; ++i) {
    const int c = i * 256 + lane * 8;
    if (c < H) {
      const uint4 u = ldg_u4(row + c);
      const float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), cc = unpack_bf16x2(u.z), d = unpack_bf16x2(u.w);
      v[i][0] = a.x; v[i][1] = a.y; v[i][2] = b.x; v[i][3] = b.y; v[i][4] = cc.x; v[i][5] = cc.y; v[i][6] = d.x; v[i][7] = d.y;
    } else {
#pragma unroll
      for (int e = 0; e < 8; ++e) v[i][e] = 0.f;
    }
  }
}
__device__ __forceinline__ float dot_row8(const float* __restrict__ w, int H, int lane, const float (&v)[HD_MAX_VEC][8]) {
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < HD_MAX_VEC; ++i) {
    const int c = i * 256 + lane * 8;
    if (c < H) {
      const float4 w0 = ldg_f4(w + c), w1 = ldg_f4(w + c + 4);
      s += v[i][0] * w0.x + v[i][1] * w0.y + v[i][2] * w0.z + v[i][3] * w0.w + v[i][4] * w1.x + v[i][5] * w1.y + v[i][6] * w1.z +
           v[i][7] * w1.w;
    }
  }
  return warp_sum(s);
}

__global__ void __launch_bounds__(256) heads_fwd_kernel(const HeadsFwdArgs a) {
  pdl_trigger();
  pdl_wait();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  for (int t = blockIdx.x * nwarp + warp; t < a.BK; t += gridDim.x * nwarp) {
    float v[HD_MAX_VEC][8];
    load_row8(a.y + ((size_t)t * 3 + 1) * a.H, a.H, lane, v);  // state token -> action prediction
    for (int j = 0; j < a.ad; ++j) {
      float z = dot_row8(a.Wa + (size_t)j * a.H, a.H, lane, v) + a.ba[j];
      if (a.tanh_act) z = tanhf(z);
      if (lane == 0) a.action_preds[(size_t)t * a.ad + j] = z;
    }
    if (a.state_preds || a.return_preds) {
      load_row8(a.y + ((size_t)t * 3 + 2) * a.H, a.H, lane, v);  // action token -> state / return predictions
      if (a.return_preds) {
        const float z = dot_row8(a.Wr, a.H, lane, v) + a.br[0];
        if (lane == 0) a.return_preds[t] = z;
      }
      if (a.state_preds)
        for (int j = 0; j < a.sd; ++j) {
          const float z = dot_row8(a.Ws + (size_t)j * a.H, a.H, lane, v) + a.bs[j];
          if (lane == 0) a.state_preds[(size_t)t * a.sd + j] = z;
        }
    }
  }
}

int heads_fwd(cudaStream_t stream, const bf16* y, const float* Wa, const float* ba, const float* Ws, const float* bs, const float* Wr,
              const float* br, int BK, int H, int sd, int ad, int tanh_act, float* action_preds, float* state_preds,
              float* return_preds) {
  DTB_REQUIRE(H % 8 == 0 && H <= HD_MAX_VEC * 256, DTB_ERR_UNSUPPORTED, "heads: hidden size %d unsupported (multiple of 8, <= 1024)", H);
  if (BK == 0) return DTB_OK;
  HeadsFwdArgs a{y, Wa, ba, Ws, bs, Wr, br, BK, H, sd, ad, tanh_act, action_preds, state_preds, return_preds};
  int blocks = (BK + 7) / 8;
  if (blocks > num_sms() * 8) blocks = num_sms() * 8;
  DTB_LAUNCH(heads_fwd_kernel, blocks, 256, 0, stream, a);
  DTB_CHECK_LAUNCH();
  return DTB_OK;
}

// ---- masked MSE ----------------------------------------------------------------------------------------------------
constexpr int MSE_BLOCKS = 128;

__global__ void __launch_bounds__(256) mse_partial_kernel(const float* __restrict__ pred, const float* __restrict__ target,
                                                          const float* __restrict__ mask, int BK, int ad, float* __restrict__ part) {
  pdl_trigger();
  pdl_wait();
  __shared__ float ssum[8], scnt[8];
  float s = 0.f, n = 0.f;
  for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < BK; t += gridDim.x * blockDim.x) {
    if (mask == nullptr || mask[t] > 0.f) {
      n += 1.f;
      for (int j = 0; j < ad; ++j) {
        const float d = pred[(size_t)t * ad + j] - target[(size_t)t * ad + j];
        s += d * d;
      }
    }
  }
  s = warp_sum(s);
  n = warp_sum(n);
  if ((threadIdx.x & 31) == 0) { ssum[threadIdx.x >> 5] = s; scnt[threadIdx.x >> 5] = n; }
  __syncthreads();
  if (threadIdx.x == 0) {
    float S = 0.f, N = 0.f;
    for (int i = 0; i < 8; ++i) { S += ssum[i]; N += scnt[i]; }
    part[2 * blockIdx.x] = S;
    part[2 * blockIdx.x + 1] = N;
  }
}

// loss_out[0] = loss, loss_out[1] = number of valid rows; d_pred = grad_scale * 2 (pred - target) / (N_valid * ad) on valid rows, 0 elsewhere.
__global__ void __launch_bounds__(256) mse_finalize_kernel(const float* __restrict__ part, int nparts, const float* __restrict__ pred,
                                                           const float* __restrict__ target, const float* __restrict__ mask, int BK,
                                                           int ad, float grad_scale, float* __restrict__ loss_out,
                                                           float* __restrict__ d_pred) {
  pdl_trigger();
  pdl_wait();
  float S = 0.f, N = 0.f;
  for (int i = 0; i < nparts; ++i) { S += part[2 * i]; N += part[2 * i + 1]; }  // every thread: same fixed order
  const float denom = N * ad;
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    loss_out[0] = denom > 0.f ? S / denom : 0.f;
    loss_out[1] = N;
  }
  if (d_pred == nullptr) return;
  const float g = denom > 0.f ? grad_scale * 2.f / denom : 0.f;
  const int total = BK * ad;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int t = i / ad;
    const bool valid = mask == nullptr || mask[t] > 0.f;
    d_pred[i] = valid ? g * (pred[i] - target[i]) : 0.f;
  }
}

size_t mse_scratch_floats() { return 2 * MSE_BLOCKS; }

int mse_loss(cudaStream_t stream, const float* pred, const float* target, const float* mask, int BK, int ad, float grad_scale,
             float* loss_out, float* d_pred, float* scratch) {
  int blocks = (BK + 255) / 256;
  if (blocks > MSE_BLOCKS) blocks = MSE_BLOCKS;
  if (blocks < 1) blocks = 1;
  DTB_LAUNCH(mse_partial_kernel, blocks, 256, 0, stream, pred, target, mask, BK, ad, scratch);
  DTB_CHECK_LAUNCH();
  int fblocks = (BK * ad + 255) / 256;
  if (fblocks > num_sms() * 4) fblocks = num_sms() * 4;
  if (fblocks < 1) fblocks = 1;
  DTB_LAUNCH(mse_finalize_kernel, fblocks, 256, 0, stream, scratch, blocks, pred, target, mask, BK, ad, grad_scale, loss_out, d_pred);
  DTB_CHECK_LAUNCH();
  return DTB_OK;
}

// ---- heads backward ------------------------------------------------------------------------------------------------
// dz_a = d_action * tanh'(action_preds) -> dzA bf16 [BK, ZA8];  dy[3t+1] = dz_a @ Wpa
// dz_s = d_state, dz_r = d_return       -> dzS bf16 [BK, ZS8] (cols 0..sd-1 state, col sd return);  dy[3t+2] = dz_s @ Wps + dz_r * wpr
// dy[3t+0] = 0.  dy is bf16 [BK*3, H] and feeds ln_f's backward; the dz copies feed the head weight-gradient GEMMs.
struct HeadsBwdArgs {
  const float* d_action; const float* d_state; const float* d_return;  // any may be null
  const float* action_preds;
  const float* Wa; const float* Ws; const float* Wr;
  int BK, H, sd, ad, tanh_act, ZA8, ZS8;
  bf16* dy; bf16* dzA; bf16* dzS;
};

__global__ void __launch_bounds__(256) heads_bwd_kernel(const HeadsBwdArgs a) {
  pdl_trigger();
  pdl_wait();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  const int H = a.H;
  for (int t = blockIdx.x * nwarp + warp; t < a.BK; t += gridDim.x * nwarp) {
    bf16* dyr = a.dy + (size_t)t * 3 * H;
    // R-token row: no head reads it
    for (int c = lane * 8; c < H; c += 256) *reinterpret_cast<uint4*>(dyr + c) = make_uint4(0, 0, 0, 0);
    // state-token row <- action head
    {
      float acc[HD_MAX_VEC][8];
#pragma unroll
      for (int i = 0; i < HD_MAX_VEC; ++i)
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[i][e] = 0.f;
      for (int j = 0; j < a.ZA8; ++j) {
        float dz = 0.f;
        if (j < a.ad && a.d_action) {
          dz = a.d_action[(size_t)t * a.ad + j];
          if (a.tanh_act) {
            const float p = a.action_preds[(size_t)t * a.ad + j];
            dz *= 1.f - p * p;
          }
        }
        if (lane == 0 && a.dzA) a.dzA[(size_t)t * a.ZA8 + j] = __float2bfloat16(dz);
        if (j < a.ad && a.d_action) {
          const float* w = a.Wa + (size_t)j * H;
#pragma unroll
          for (int i = 0; i < HD_MAX_VEC; ++i) {
            const int c = i * 256 + lane * 8;
            if (c < H) {
              const float4 w0 = ldg_f4(w + c), w1 = ldg_f4(w + c + 4);
              acc[i][0] += dz * w0.x; acc[i][1] += dz * w0.y; acc[i][2] += dz * w0.z; acc[i][3] += dz * w0.w;
              acc[i][4] += dz * w1.x; acc[i][5] += dz * w1.y; acc[i][6] += dz * w1.z; acc[i][7] += dz * w1.w;
            }
          }
        }
      }
#pragma unroll
      for (int i = 0; i < HD_MAX_VEC; ++i) {
        const int c = i * 256 + lane * 8;
        if (c < H)
          *reinterpret_cast<uint4*>(dyr + H + c) = make_uint4(pack_bf16x2(acc[i][0], acc[i][1]), pack_bf16x2(acc[i][2], acc[i][3]),
                                                              pack_bf16x2(acc[i][4], acc[i][5]), pack_bf16x2(acc[i][6], acc[i][7]));
      }
    }
    // action-token row <- state / return heads
    {
      float acc[HD_MAX_VEC][8];
#pragma unroll
      for (int i = 0; i < HD_MAX_VEC; ++i)
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[i][e] = 0.f;
      if (a.d_state || a.d_return) {
        for (int j = 0; j < a.ZS8; ++j) {
          float dz = 0.f;
          const float* w = nullptr;
          if (j < a.sd && a.d_state) { dz = a.d_state[(size_t)t * a.sd + j]; w = a.Ws + (size_t)j * H; }
          else if (j == a.sd && a.d_return) { dz = a.d_return[t]; w = a.Wr; }
          if (lane == 0 && a.dzS) a.dzS[(size_t)t * a.ZS8 + j] = __float2bfloat16(dz);
          if (w) {
#pragma unroll
            for (int i = 0; i < HD_MAX_VEC; ++i) {
              const int c = i * 256 + lane * 8;
              if (c < H) {
                const float4 w0 = ldg_f4(w + c), w1 = ldg_f4(w + c + 4);
                acc[i][0] += dz * w0.x; acc[i][1] += dz * w0.y; acc[i][2] += dz * w0.z; acc[i][3] += dz * w0.w;
                acc[i][4] += dz * w1.x; acc[i][5] += dz * w1.y; acc[i][6] += dz * w1.z; acc[i][7] += dz * w1.w;
              }
            }
          }
        }
      }
#pragma unroll
      for (int i = 0; i < HD_MAX_VEC; ++i) {
        const int c = i * 256 + lane * 8;
        if (c < H)
          *reinterpret_cast<uint4*>(dyr + 2 * H + c) = make_uint4(pack_bf16x2(acc[i][0], acc[i][1]), pack_bf16x2(acc[i][2], acc[i][3]),
                                                                  pack_bf16x2(acc[i][4], acc[i][5]), pack_bf16x2(acc[i][6], acc[i][7]));
      }
    }
  }
}

int heads_bwd(cudaStream_t stream, const float* d_action, const float* d_state, const float* d_return, const float* action_preds,
              const float* Wa, const float* Ws, const float* Wr, int BK, int H, int sd, int ad, int tanh_act, int ZA8, int ZS8, bf16* dy,
              bf16* dzA, bf16* dzS) {
  DTB_REQUIRE(H % 8 == 0 && H <= HD_MAX_VEC * 256, DTB_ERR_UNSUPPORTED, "heads bwd: hidden size %d unsupported", H);
  if (BK == 0) return DTB_OK;
  HeadsBwdArgs a{d_action, d_state, d_return, action_preds, Wa, Ws, Wr, BK, H, sd, ad, tanh_act, ZA8, ZS8, dy, dzA, dzS};
  int blocks = (BK + 7) / 8;
  if (blocks > num_sms() * 8) blocks = num_sms() * 8;
  DTB_LAUNCH(heads_bwd_kernel, blocks, 256, 0, stream, a);
  DTB_CHECK_LAUNCH();
  return DTB_OK;
}

// Copy rows/cols out of a padded fp32 matrix: dst[r][c] = src[r * lds + c], r < R, c < C (head weight grads from the padded GEMM result).
__global__ void copy2d_kernel(const float* __restrict__ src, int lds, int R, int C, float* __restrict__ dst, int ldd) {
  pdl_trigger();
  pdl_wait();
  const int total = R * C;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int r = i / C, c = i - r * C;
    dst[(size_t)r * ldd + c] = src[(size_t)r * lds + c];
  }
}
int copy2d(cudaStream_t stream, const float* src, int lds, int R, int C, float* dst, int ldd) {
  if (R * C == 0) return DTB_OK;
  int blocks = (R * C + 255) / 256;
  if (blocks > num_sms() * 4) blocks = num_sms() * 4;
  DTB_LAUNCH(copy2d_kernel, blocks, 256, 0, stream, src, lds, R, C, dst, ldd);
  DTB_CHECK_LAUNCH();
  return DTB_OK;
}

}  // namespace dtb
