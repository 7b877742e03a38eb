// Fused optimizer step of the reference trainer: upstream's train_step() does
//     torch.nn.utils.clip_grad_norm_(model.parameters(), .25);  optimizer.step()      with torch.optim.AdamW(lr, weight_decay)
// after loss.backward().  Here the parameters / gradients of a model are single flat fp32 buffers, so the whole tail of the step is
// three launches: per-block sum of squares -> (1 block) global norm, clip coefficient, step count and bias corrections on the
// device -> one pass that clips the gradient, applies decoupled weight decay + Adam and re-stages the bf16 shadow the GEMMs read
// (the separate fp32 -> bf16 staging pass of the next forward disappears).  No host synchronisation: the step count and an optional
// learning rate live in device memory, so the step can sit inside the CUDA graph of the training step.
// Arithmetic follows torch (clip_grad_norm_: coef = min(1, max_norm / (norm + 1e-6));  AdamW single-tensor path:
//   p *= 1 - lr * wd;  m = b1 m + (1 - b1) g;  v = b2 v + (1 - b2) g^2;  p -= (lr / (1 - b1^t)) * m / (sqrt(v) / sqrt(1 - b2^t) + eps)).
#include "kernels.h"

namespace dtb {

constexpr int OPT_MAX_SEG = 8;
constexpr int OPT_BLOCKS_PER_SEG = 2 * DTB_NUM_SMS_FALLBACK;  // fixed (not the live SM count) so the scratch layout is static
constexpr int OPT_HDR = 8;                                     // scratch[0..4]: clip, lr / (1 - b1^t), sqrt(1 - b2^t), norm, 1 - lr wd

struct OptSegs {
  float* p[OPT_MAX_SEG];
  float* g[OPT_MAX_SEG];
  float* m[OPT_MAX_SEG];
  float* v[OPT_MAX_SEG];
  bf16* pb[OPT_MAX_SEG];
  long long n[OPT_MAX_SEG];
  int nseg;
};

__global__ void __launch_bounds__(256) opt_sumsq_kernel(const OptSegs s, float* __restrict__ scratch) {
  pdl_trigger();
  pdl_wait();
  const int seg = blockIdx.y;
  const float* g = s.g[seg];
  const long long n = s.n[seg], n4 = n >> 2;
  float acc = 0.f;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
    const float4 v = ldg_f4(g + 4 * i);
    acc += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
  }
  if (blockIdx.x == 0 && threadIdx.x < (int)(n & 3)) {
    const float v = g[(n4 << 2) + threadIdx.x];
    acc += v * v;
  }
  __shared__ float red[8];
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int w = 0; w < 8; ++w) t += red[w];  // fixed order: deterministic
    scratch[OPT_HDR + seg * OPT_BLOCKS_PER_SEG + blockIdx.x] = t;
  }
}

// one block: global norm (partials combined in index order, in double), clip coefficient, step count and the per-step scalars.
// torch evaluates those scalars in Python doubles before they meet fp32 tensors; the same is done here (hyper-parameters are doubles).
__global__ void opt_prepare_kernel(float* __restrict__ scratch, int nparts, double max_norm, double beta1, double beta2, double lr,
                                   double weight_decay, int warmup_steps, const float* __restrict__ lr_dev, int* __restrict__ step_dev) {
  pdl_trigger();
  pdl_wait();
  if (threadIdx.x != 0) return;
  double total = 0.0;
  for (int i = 0; i < nparts; ++i) total += (double)scratch[OPT_HDR + i];
  const float norm = (float)sqrt(total);
  float clip = 1.f;
  if (max_norm > 0.0) {
    clip = (float)max_norm / (norm + 1e-6f);
    clip = clip < 1.f ? clip : 1.f;
  }
  const int t = *step_dev + 1;
  *step_dev = t;
  if (lr_dev) lr = (double)*lr_dev;
  // upstream scheduler: LambdaLR(optimizer, lambda steps: min((steps + 1) / warmup_steps, 1)) -> at the t-th optimizer step lr * min(t / warmup, 1)
  if (warmup_steps > 0 && t < warmup_steps) lr *= (double)t / (double)warmup_steps;
  scratch[0] = clip;
  scratch[1] = (float)(lr / (1.0 - pow(beta1, (double)t)));   // step_size = lr / bias_correction1
  scratch[2] = (float)sqrt(1.0 - pow(beta2, (double)t));      // bias_correction2_sqrt
  scratch[3] = norm;
  scratch[4] = (float)(1.0 - lr * weight_decay);              // decoupled decay factor
}

struct OptScalars { float b2, omb1, omb2, eps; };  // beta2, 1 - beta1, 1 - beta2 (rounded from doubles, as torch's scalar arguments are)

__device__ __forceinline__ void adamw_one(float& p, float& g, float& m, float& v, float clip, float decay, const OptScalars& c, float step_size,
                                          float bc2_sqrt) {
  g *= clip;
  p *= decay;
  m = m + (g - m) * c.omb1;            // exp_avg.lerp_(grad, 1 - beta1)
  v = v * c.b2 + c.omb2 * g * g;       // exp_avg_sq.mul_(beta2).addcmul_(grad, grad, value = 1 - beta2)
  const float denom = sqrtf(v) / bc2_sqrt + c.eps;
  p -= step_size * (m / denom);        // param.addcdiv_(exp_avg, denom, value = -step_size)
}

__global__ void __launch_bounds__(256) opt_adamw_kernel(const OptSegs s, const float* scratch, const OptScalars c, int write_grads) {
  pdl_trigger();
  pdl_wait();
  const int seg = blockIdx.y;
  float* __restrict__ P = s.p[seg];
  float* __restrict__ G = s.g[seg];
  float* __restrict__ M = s.m[seg];
  float* __restrict__ V = s.v[seg];
  bf16* __restrict__ PB = s.pb[seg];
  const long long n = s.n[seg], n4 = n >> 2;
  // volatile: the compiler hoists plain `const __restrict__` loads above griddepcontrol.wait (seen in SASS: LDG.E.CONSTANT before
  // ACQBULK), i.e. before opt_prepare_kernel of THIS step has written them (tests/test_build_cpu.py scans every kernel for this)
  const volatile float* hdr = scratch;
  const float clip = hdr[0], step_size = hdr[1], bc2s = hdr[2], decay = hdr[4];
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
    float4 p = *reinterpret_cast<const float4*>(P + 4 * i), g = *reinterpret_cast<const float4*>(G + 4 * i);
    float4 m = *reinterpret_cast<const float4*>(M + 4 * i), v = *reinterpret_cast<const float4*>(V + 4 * i);
    adamw_one(p.x, g.x, m.x, v.x, clip, decay, c, step_size, bc2s);
    adamw_one(p.y, g.y, m.y, v.y, clip, decay, c, step_size, bc2s);
    adamw_one(p.z, g.z, m.z, v.z, clip, decay, c, step_size, bc2s);
    adamw_one(p.w, g.w, m.w, v.w, clip, decay, c, step_size, bc2s);
    *reinterpret_cast<float4*>(P + 4 * i) = p;
    *reinterpret_cast<float4*>(M + 4 * i) = m;
    *reinterpret_cast<float4*>(V + 4 * i) = v;
    if (write_grads) *reinterpret_cast<float4*>(G + 4 * i) = g;  // clip_grad_norm_ scales .grad in place
    if (PB) *reinterpret_cast<uint2*>(PB + 4 * i) = make_uint2(pack_bf16x2(p.x, p.y), pack_bf16x2(p.z, p.w));
  }
  if (blockIdx.x == 0 && threadIdx.x < (int)(n & 3)) {
    const long long i = (n4 << 2) + threadIdx.x;
    float p = P[i], g = G[i], m = M[i], v = V[i];
    adamw_one(p, g, m, v, clip, decay, c, step_size, bc2s);
    P[i] = p; M[i] = m; V[i] = v;
    if (write_grads) G[i] = g;
    if (PB) PB[i] = __float2bfloat16(p);
  }
}

size_t adamw_scratch_floats() { return OPT_HDR + (size_t)OPT_MAX_SEG * OPT_BLOCKS_PER_SEG; }

int clip_adamw_step(cudaStream_t stream, int nseg, const dtb_adamw_segment* segs, const dtb_adamw_hparams* hp, int* step_dev,
                    float* scratch) {
  DTB_REQUIRE(nseg >= 1 && nseg <= OPT_MAX_SEG && segs && hp && step_dev && scratch, DTB_ERR_BAD_ARG, "clip_adamw_step: bad arguments");
  DTB_REQUIRE(hp->beta1 >= 0.0 && hp->beta1 < 1.0 && hp->beta2 >= 0.0 && hp->beta2 < 1.0 && hp->eps > 0.0, DTB_ERR_BAD_ARG,
              "clip_adamw_step: bad hyper-parameters");
  OptSegs s;
  memset(&s, 0, sizeof(s));
  s.nseg = nseg;
  long long nmax = 0;
  for (int i = 0; i < nseg; ++i) {
    const dtb_adamw_segment& g = segs[i];
    DTB_REQUIRE(g.params && g.grads && g.exp_avg && g.exp_avg_sq && g.n >= 0, DTB_ERR_BAD_ARG, "clip_adamw_step: segment %d has a null buffer", i);
    const uintptr_t al = reinterpret_cast<uintptr_t>(g.params) | reinterpret_cast<uintptr_t>(g.grads) | reinterpret_cast<uintptr_t>(g.exp_avg) |
                         reinterpret_cast<uintptr_t>(g.exp_avg_sq);
    DTB_REQUIRE((al & 15) == 0 && (reinterpret_cast<uintptr_t>(g.params_bf16) & 7) == 0, DTB_ERR_BAD_ARG,
                "clip_adamw_step: segment %d buffers must be 16-byte aligned", i);
    s.p[i] = g.params; s.g[i] = g.grads; s.m[i] = g.exp_avg; s.v[i] = g.exp_avg_sq; s.pb[i] = reinterpret_cast<bf16*>(g.params_bf16);
    s.n[i] = g.n;
    if (g.n > nmax) nmax = g.n;
  }
  if (nmax == 0) return DTB_OK;
  long long want = (nmax / 4 + 255) / 256;
  const int blocks = (int)(want < 1 ? 1 : (want > OPT_BLOCKS_PER_SEG ? OPT_BLOCKS_PER_SEG : want));
  // every [seg][OPT_BLOCKS_PER_SEG] partial slot is written on every call (blocks without work store 0), then summed in index order
  DTB_LAUNCH(opt_sumsq_kernel, dim3(OPT_BLOCKS_PER_SEG, nseg), 256, 0, stream, s, scratch);
  DTB_CHECK_LAUNCH();
  DTB_LAUNCH(opt_prepare_kernel, 1, 32, 0, stream, scratch, nseg * OPT_BLOCKS_PER_SEG, hp->max_norm, hp->beta1, hp->beta2, hp->lr,
             hp->weight_decay, hp->warmup_steps, hp->lr_dev, step_dev);
  DTB_CHECK_LAUNCH();
  OptScalars c;
  c.b2 = (float)hp->beta2; c.omb1 = (float)(1.0 - hp->beta1); c.omb2 = (float)(1.0 - hp->beta2); c.eps = (float)hp->eps;
  DTB_LAUNCH(opt_adamw_kernel, dim3(blocks, nseg), 256, 0, stream, s, scratch, c, hp->write_clipped_grads);
  DTB_CHECK_LAUNCH();
  return DTB_OK;
}

}  // namespace dtb
