// Audio front-end on the GPU: wav samples -> 13 MFCCs per 25 ms frame -> bicubic-resized, min-max normalised image.
// Replaces REF environments/Minecraft/Minecraft.py:226-245 (`_get_mfcc` / `_wav_to_mfcc`; same code in Doorpuzzle.py:264-282):
//   python_speech_features.mfcc(signal, samplerate)  [defaults of release 0.6: winlen 25 ms, winstep 10 ms, rectangular window,
//   nfft = next power of two >= window, 26 mel filters over 0..sr/2, log, DCT-II (ortho) -> 13, lifter 22, c0 := log(frame energy),
//   pre-emphasis 0.97]  ->  swapaxes  ->  PIL Image.resize((W, H), BICUBIC) on a float32 image  ->  (x - min) / (max - min) * 2 - 1.
// The reference computes the cepstra in float64 (numpy) and the image in float32 (Pillow mode "F"); so does this file: one CTA per
// frame runs a shared-memory radix-2 FFT in double (f64 is the arithmetic type of this path), the image kernel follows Pillow's
// ImagingResample (Resample.c: precompute_coeffs + horizontal pass then vertical pass, double accumulators, float32 storage) and the
// reference's float32 normalisation.  HBM traffic is the signal once (2 B / sample) — the path is latency / FP64-bound, not HBM-bound.
#include "../../include/dtb200.h"
#include "kernels.h"

#include <math.h>
#include <string.h>

namespace dtb {
int require_sm100();

namespace {

constexpr int NFILT = 26, NCEP = 13, MAX_NFFT = 4096, FFT_THREADS = 256;
constexpr double CEPLIFTER = 22.0, PREEMPH = 0.97;
constexpr double F64_EPS = 2.220446049250313e-16;  // numpy.finfo(float).eps

struct MfccGeom {
  int frame_len, frame_step, nfft, log2n;
  long long numframes;
  int bin[NFILT + 2];
};

// decimal ROUND_HALF_UP of a positive double (sigproc.round_half_up)
long long round_half_up(double x) { return (long long)floor(x + 0.5); }

int make_geom(long long n_samples, int samplerate, MfccGeom& g) {
  DTB_REQUIRE(n_samples >= 1 && samplerate >= 1000, DTB_ERR_BAD_ARG, "mfcc: need n_samples >= 1 and samplerate >= 1000");
  const double wl = 0.025 * (double)samplerate, ws = 0.01 * (double)samplerate;
  g.frame_len = (int)round_half_up(wl);
  g.frame_step = (int)round_half_up(ws);
  int nfft = 1, l2 = 0;
  while ((double)nfft < wl) { nfft *= 2; ++l2; }  // base.calculate_nfft
  DTB_REQUIRE(nfft <= MAX_NFFT && nfft >= 64, DTB_ERR_UNSUPPORTED, "mfcc: samplerate %d needs an FFT of %d points (supported: 64..%d)",
              samplerate, nfft, MAX_NFFT);
  DTB_REQUIRE(g.frame_len <= nfft, DTB_ERR_UNSUPPORTED, "mfcc: frame longer than the FFT");
  g.nfft = nfft; g.log2n = l2;
  if (n_samples <= g.frame_len) g.numframes = 1;
  else g.numframes = 1 + (long long)ceil((1.0 * (double)n_samples - g.frame_len) / g.frame_step);
  // base.get_filterbanks: bin = floor((nfft + 1) * mel2hz(linspace(hz2mel(0), hz2mel(sr / 2), nfilt + 2)) / samplerate)
  const double lowmel = 2595.0 * log10(1.0 + 0.0 / 700.0), highmel = 2595.0 * log10(1.0 + ((double)samplerate / 2) / 700.0);
  const double step = (highmel - lowmel) / (double)(NFILT + 1);
  for (int i = 0; i < NFILT + 2; ++i) {
    const double mel = i == NFILT + 1 ? highmel : (double)i * step + lowmel;  // numpy.linspace
    const double hz = 700.0 * (pow(10.0, mel / 2595.0) - 1.0);
    g.bin[i] = (int)floor((double)(nfft + 1) * hz / (double)samplerate);
  }
  return DTB_OK;
}

__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// One CTA per (clip, frame).  smem: re[nfft], im[nfft], twiddle cos/sin [nfft / 2] (doubles), then pspec reuses re[].
template <bool I16>
__global__ void __launch_bounds__(FFT_THREADS) mfcc_frame_kernel(const void* __restrict__ signal_, long long n_samples, long long clip_stride,
                                                                 int lag, MfccGeom g, double* __restrict__ cep) {
  pdl_trigger();
  pdl_wait();
  extern __shared__ double sm[];
  const int n = g.nfft, half = n >> 1;
  double* re = sm;
  double* im = sm + n;
  double* tc = sm + 2 * n;
  double* ts = tc + half;
  __shared__ double red[FFT_THREADS / 32];
  __shared__ double fb[NFILT];
  __shared__ double energy_s;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const long long clip = blockIdx.x / g.numframes, frame = blockIdx.x % g.numframes;
  const long long start = frame * g.frame_step;

  // twiddles e^{-2 pi i k / n}
  for (int k = tid; k < half; k += FFT_THREADS) {
    double s, c;
    sincospi(2.0 * (double)k / (double)n, &s, &c);
    tc[k] = c;
    ts[k] = -s;
  }
  // pre-emphasised frame, bit-reversed order.  sigproc.preemphasis: append(x[0], x[1:] - 0.97 x[:-1]) acts per channel ROW of an
  // [n, channels] array and is then flattened: in interleaved samples y[p] = x[p] - 0.97 x[p - lag] (p >= lag), y[p] = x[p] (p < lag),
  // lag = channels.  Zero padding is appended AFTER the pre-emphasis.
  for (int i = tid; i < n; i += FFT_THREADS) {
    double v = 0.0;
    const long long p = start + i;
    if (i < g.frame_len && p < n_samples) {
      double x0, x1 = 0.0;
      if (I16) {
        const short* s = reinterpret_cast<const short*>(signal_) + clip * clip_stride;
        x0 = (double)s[p];
        if (p >= lag) x1 = (double)s[p - lag];
      } else {
        const float* s = reinterpret_cast<const float*>(signal_) + clip * clip_stride;
        x0 = (double)s[p];
        if (p >= lag) x1 = (double)s[p - lag];
      }
      v = p >= lag ? x0 - PREEMPH * x1 : x0;
    }
    const int r = (int)(__brev((unsigned)i) >> (32 - g.log2n));
    re[r] = v;
    im[r] = 0.0;
  }
  __syncthreads();
  // radix-2 decimation-in-time butterflies
  for (int s = 0; s < g.log2n; ++s) {
    const int h = 1 << s, tstride = half >> s;
    for (int t = tid; t < half; t += FFT_THREADS) {
      const int j = t & (h - 1), i0 = ((t >> s) << (s + 1)) + j, i1 = i0 + h;
      const double wr = tc[j * tstride], wi = ts[j * tstride];
      const double xr = re[i1], xi = im[i1];
      const double pr = xr * wr - xi * wi, pi = xr * wi + xi * wr;
      const double ur = re[i0], ui = im[i0];
      re[i0] = ur + pr; im[i0] = ui + pi;
      re[i1] = ur - pr; im[i1] = ui - pi;
    }
    __syncthreads();
  }
  // power spectrum (sigproc.powspec: |rfft|^2 / nfft) over bins 0..n/2, kept in tc[0..half] (half + 1 <= 2 * half doubles of tc + ts)
  double* ps = tc;
  double e = 0.0;
  __syncthreads();
  for (int k = tid; k <= half; k += FFT_THREADS) {
    const double a = re[k], b = im[k];
    const double p = (a * a + b * b) / (double)n;
    ps[k] = p;
    e += p;
  }
  e = warp_sum_d(e);
  if (lane == 0) red[wid] = e;
  __syncthreads();
  if (tid == 0) {
    double s = 0.0;
    for (int i = 0; i < FFT_THREADS / 32; ++i) s += red[i];
    energy_s = s == 0.0 ? F64_EPS : s;
  }
  // mel filterbank: filter j rises over [bin[j], bin[j+1]) and falls over [bin[j+1], bin[j+2])
  for (int j = wid; j < NFILT; j += FFT_THREADS / 32) {
    const int b0 = g.bin[j], b1 = g.bin[j + 1], b2 = g.bin[j + 2];
    double acc = 0.0;
    for (int i = b0 + lane; i < b2; i += 32) {
      const double wgt = i < b1 ? (double)(i - b0) / (double)(b1 - b0) : (double)(b2 - i) / (double)(b2 - b1);
      acc += ps[i] * wgt;
    }
    acc = warp_sum_d(acc);
    if (lane == 0) fb[j] = log(acc == 0.0 ? F64_EPS : acc);
  }
  __syncthreads();
  // DCT-II, norm='ortho' (scipy.fftpack.dct), first 13 coefficients; lifter 1 + 11 sin(pi k / 22); c0 := log(energy)
  if (tid < NCEP) {
    const int k = tid;
    double s = 0.0;
    for (int m = 0; m < NFILT; ++m) s += fb[m] * cospi((double)k * (double)(2 * m + 1) / (double)(2 * NFILT));
    s *= 2.0 * (k == 0 ? sqrt(1.0 / (4.0 * NFILT)) : sqrt(1.0 / (2.0 * NFILT)));
    s *= 1.0 + (CEPLIFTER / 2.0) * sinpi((double)k / CEPLIFTER);
    if (k == 0) s = log(energy_s);
    cep[(clip * g.numframes + frame) * NCEP + k] = s;
  }
}

// ---- image: Pillow's ImagingResample (bicubic, a = -0.5, antialiased support) + the reference's min-max normalisation ------------
// Every double operation is written with the non-contracting intrinsics (__dmul_rn / __dadd_rn ...): Pillow's C code is compiled
// without FMA contraction, and the image is compared with Pillow bit for bit.
__device__ __forceinline__ double bicubic_filter(double x) {  // Resample.c bicubic_filter, a = -0.5
  const double a = -0.5;
  if (x < 0.0) x = -x;
  if (x < 1.0) return __dadd_rn(__dmul_rn(__dmul_rn(__dsub_rn(__dmul_rn(a + 2.0, x), a + 3.0), x), x), 1.0);
  if (x < 2.0) return __dmul_rn(__dsub_rn(__dmul_rn(__dadd_rn(__dmul_rn(__dsub_rn(x, 5.0), x), 8.0), x), 4.0), a);
  return 0.0;
}

struct ResampleAxis { double scale, support, ss; int in_size; };
__device__ __forceinline__ ResampleAxis make_axis(int in_size, int out_size) {  // precompute_coeffs, box = the whole image
  ResampleAxis ax;
  ax.in_size = in_size;
  ax.scale = __ddiv_rn((double)in_size, (double)out_size);
  const double fs = ax.scale < 1.0 ? 1.0 : ax.scale;
  ax.support = __dmul_rn(2.0, fs);  // bicubic support 2.0
  ax.ss = __ddiv_rn(1.0, fs);
  return ax;
}
// bounds and normaliser of output position xx
__device__ __forceinline__ void axis_bounds(const ResampleAxis& ax, int xx, double& center, int& xmin, int& cnt, double& ww) {
  center = __dmul_rn(__dadd_rn((double)xx, 0.5), ax.scale);  // in0 = 0
  xmin = (int)__dadd_rn(__dsub_rn(center, ax.support), 0.5);
  if (xmin < 0) xmin = 0;
  int xmax = (int)__dadd_rn(__dadd_rn(center, ax.support), 0.5);
  if (xmax > ax.in_size) xmax = ax.in_size;
  cnt = xmax - xmin;
  ww = 0.0;
  for (int x = 0; x < cnt; ++x) ww = __dadd_rn(ww, bicubic_filter(__dmul_rn(__dadd_rn(__dsub_rn((double)(x + xmin), center), 0.5), ax.ss)));
}
__device__ __forceinline__ double axis_coeff(const ResampleAxis& ax, double center, int xmin, int x, double ww) {
  const double w = bicubic_filter(__dmul_rn(__dadd_rn(__dsub_rn((double)(x + xmin), center), 0.5), ax.ss));
  return ww != 0.0 ? __ddiv_rn(w, ww) : w;
}

// One CTA per clip.  cep [frames, 13] (double) is the transposed source image: pixel (row y = cepstral index, column x = frame) as float32.
__global__ void __launch_bounds__(512) mfcc_image_kernel(const double* __restrict__ cep, long long numframes, int W, int H,
                                                         float* __restrict__ tmp_all, float* __restrict__ image_all) {
  pdl_trigger();
  pdl_wait();
  __shared__ float smin[16], smax[16];
  const double* src = cep + (size_t)blockIdx.x * numframes * NCEP;
  float* tmp = tmp_all + (size_t)blockIdx.x * NCEP * W;   // [13, W] after the horizontal pass
  float* img = image_all + (size_t)blockIdx.x * H * W;    // [H, W]
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const bool need_h = numframes != (long long)W, need_v = H != NCEP;
  // horizontal pass (ImagingResampleHorizontal_32bpc)
  {
    const ResampleAxis ax = make_axis((int)numframes, W);
    for (int o = tid; o < NCEP * W; o += blockDim.x) {
      const int yy = o / W, xx = o % W;
      float r;
      if (need_h) {
        double center, ww; int xmin, cnt;
        axis_bounds(ax, xx, center, xmin, cnt, ww);
        double ss = 0.0;
        for (int x = 0; x < cnt; ++x)
          ss = __dadd_rn(ss, __dmul_rn((double)(float)src[(size_t)(x + xmin) * NCEP + yy], axis_coeff(ax, center, xmin, x, ww)));
        r = (float)ss;
      } else {
        r = (float)src[(size_t)xx * NCEP + yy];
      }
      tmp[o] = r;
    }
  }
  __syncthreads();
  // vertical pass (ImagingResampleVertical_32bpc) + running min / max
  float mn = INFINITY, mx = -INFINITY;
  {
    const ResampleAxis ax = make_axis(NCEP, H);
    for (int o = tid; o < H * W; o += blockDim.x) {
      const int yy = o / W, xx = o % W;
      float r;
      if (need_v) {
        double center, ww; int ymin, cnt;
        axis_bounds(ax, yy, center, ymin, cnt, ww);
        double ss = 0.0;
        for (int y = 0; y < cnt; ++y) ss = __dadd_rn(ss, __dmul_rn((double)tmp[(y + ymin) * W + xx], axis_coeff(ax, center, ymin, y, ww)));
        r = (float)ss;
      } else {
        r = tmp[o];
      }
      img[o] = r;
      mn = fminf(mn, r);
      mx = fmaxf(mx, r);
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  }
  if (lane == 0) { smin[wid] = mn; smax[wid] = mx; }
  __syncthreads();
  mn = smin[0]; mx = smax[0];
  for (int i = 1; i < (int)(blockDim.x >> 5); ++i) { mn = fminf(mn, smin[i]); mx = fmaxf(mx, smax[i]); }
  // Minecraft.py:240-243 in float32: (x - min) / (max - min), then * 2 - 1 (separate roundings, no FMA)
  const float range = __fsub_rn(mx, mn);
  for (int o = tid; o < H * W; o += blockDim.x)
    img[o] = __fsub_rn(__fmul_rn(__fdiv_rn(__fsub_rn(img[o], mn), range), 2.0f), 1.0f);
}

int launch_cepstra(cudaStream_t st, const void* signal, int is_i16, int channels, int n_clips, long long n_samples, long long clip_stride,
                   const MfccGeom& g, double* cep) {
  DTB_REQUIRE(channels >= 1 && channels <= 8, DTB_ERR_BAD_ARG, "mfcc: channels must be 1..8");
  const long long blocks = (long long)n_clips * g.numframes;
  DTB_REQUIRE(blocks <= 0x7fffffffLL, DTB_ERR_UNSUPPORTED, "mfcc: too many frames in one launch");
  const size_t smem = (size_t)3 * g.nfft * sizeof(double);
  static bool attr_set = false;
  if (!attr_set) {
    DTB_CUDA(cudaFuncSetAttribute(mfcc_frame_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 3 * MAX_NFFT * (int)sizeof(double)));
    DTB_CUDA(cudaFuncSetAttribute(mfcc_frame_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 3 * MAX_NFFT * (int)sizeof(double)));
    attr_set = true;
  }
  if (is_i16) DTB_LAUNCH(mfcc_frame_kernel<true>, (unsigned)blocks, FFT_THREADS, smem, st, signal, n_samples, clip_stride, channels, g, cep);
  else DTB_LAUNCH(mfcc_frame_kernel<false>, (unsigned)blocks, FFT_THREADS, smem, st, signal, n_samples, clip_stride, channels, g, cep);
  DTB_CHECK_LAUNCH();
  return DTB_OK;
}

inline size_t al256(size_t n) { return (n + 255) & ~size_t(255); }

}  // namespace
}  // namespace dtb

using namespace dtb;

extern "C" {

long long dtb_mfcc_num_frames(long long n_samples, int samplerate) {
  MfccGeom g;
  DTB_TRY(make_geom(n_samples, samplerate, g));
  return g.numframes;
}
int dtb_mfcc_filterbank_bins(int samplerate, int* nfft, int* bins) {
  MfccGeom g;
  DTB_TRY(make_geom(1, samplerate, g));
  if (nfft) *nfft = g.nfft;
  if (bins) for (int i = 0; i < NFILT + 2; ++i) bins[i] = g.bin[i];
  return DTB_OK;
}
size_t dtb_mfcc_workspace_bytes(int n_clips, long long n_samples, int samplerate, int out_w) {
  MfccGeom g;
  if (n_clips < 1 || out_w < 1 || make_geom(n_samples, samplerate, g) != DTB_OK) return 0;
  return al256((size_t)n_clips * g.numframes * NCEP * sizeof(double)) + al256((size_t)n_clips * NCEP * out_w * sizeof(float)) + 256;
}
int dtb_mfcc_cepstra(void* stream, const void* signal, int is_i16, int channels, int n_clips, long long n_samples, long long clip_stride, int samplerate,
                     double* cepstra) {
  DTB_REQUIRE(signal && cepstra && n_clips >= 1, DTB_ERR_BAD_ARG, "dtb_mfcc_cepstra: null argument");
  DTB_TRY(require_sm100());
  MfccGeom g;
  DTB_TRY(make_geom(n_samples, samplerate, g));
  return launch_cepstra((cudaStream_t)stream, signal, is_i16, channels, n_clips, n_samples, clip_stride, g, cepstra);
}
int dtb_mfcc_resize_normalize(void* stream, const double* cepstra, int n_clips, long long numframes, int out_w, int out_h, float* scratch,
                              float* image) {
  DTB_REQUIRE(cepstra && scratch && image && n_clips >= 1 && numframes >= 1 && out_w >= 1 && out_h >= 1, DTB_ERR_BAD_ARG,
              "dtb_mfcc_resize_normalize: bad argument");
  DTB_REQUIRE(numframes <= 0x7fffffffLL / NCEP, DTB_ERR_UNSUPPORTED, "mfcc: clip too long");
  DTB_TRY(require_sm100());
  DTB_LAUNCH(mfcc_image_kernel, n_clips, 512, 0, (cudaStream_t)stream, cepstra, numframes, out_w, out_h, scratch, image);
  DTB_CHECK_LAUNCH();
  return DTB_OK;
}
int dtb_wav_to_mfcc_image(void* stream, const void* signal, int is_i16, int channels, int n_clips, long long n_samples, long long clip_stride,
                          int samplerate, int out_w, int out_h, void* workspace, size_t workspace_bytes, float* image) {
  DTB_REQUIRE(signal && workspace && image && n_clips >= 1 && out_w >= 1 && out_h >= 1, DTB_ERR_BAD_ARG, "dtb_wav_to_mfcc_image: bad argument");
  DTB_TRY(require_sm100());
  MfccGeom g;
  DTB_TRY(make_geom(n_samples, samplerate, g));
  uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(workspace) + 255) & ~uintptr_t(255));
  const size_t cep_bytes = al256((size_t)n_clips * g.numframes * NCEP * sizeof(double));
  const size_t need = (size_t)(base - reinterpret_cast<uint8_t*>(workspace)) + cep_bytes + al256((size_t)n_clips * NCEP * out_w * sizeof(float));
  DTB_REQUIRE(need <= workspace_bytes, DTB_ERR_WORKSPACE, "mfcc: workspace too small (%zu bytes given, %zu needed)", workspace_bytes, need);
  double* cep = reinterpret_cast<double*>(base);
  float* scratch = reinterpret_cast<float*>(base + cep_bytes);
  DTB_TRY(launch_cepstra((cudaStream_t)stream, signal, is_i16, channels, n_clips, n_samples, clip_stride, g, cep));
  return dtb_mfcc_resize_normalize(stream, cep, n_clips, g.numframes, out_w, out_h, scratch, image);
}

}  // extern "C"
