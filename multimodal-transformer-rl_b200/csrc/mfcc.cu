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

// One CTA per (clip, frame).  The real frame of n = nfft samples is packed into m = n / 2 complex points z[i] = y[2i] + i y[2i+1];
// an m-point FFT (radix-4 passes in registers: two butterfly stages per shared-memory round trip, one radix-2 pass first when log2 m
// is odd) and the split X[k] = E[k] + e^{-2 pi i k / n} O[k] give the one-sided spectrum.  smem (doubles): zr[m + 1], zi[m + 1],
// twiddle table e^{-2 pi i k / n} for k < m (the m-point twiddles are its even entries); the power spectrum overwrites zr in place.
// Bank conflicts (ncu v1: 64 % of the shared wavefronts): zr / zi are padded by one double per 16 (PD) and, while the half size of a
// pass is < 16, lanes run over butterfly groups instead of over j, which makes every pass conflict-free; the frame is first staged in
// natural order (padded by one double per 128, PS) in the twiddle area and then gathered in bit-reversed order, so neither the
// global loads nor the shared stores of the load phase are scattered.
__device__ __forceinline__ int PD(int i) { return i + (i >> 4); }
__device__ __forceinline__ int PS(int i) { return i + (i >> 7); }
// Tables shared by every frame of a launch (one small kernel per call, kept in the caller's workspace): the twiddles e^{-2 pi i k / n},
// k < n / 2 (cos, -sin) and the 13 x 26 DCT-II matrix with scipy's 'ortho' scaling and the lifter folded in.
constexpr int DCT_TABLE = NCEP * NFILT;
// Twiddle table length (complex entries) for an n-point real frame, m = n / 2: [0, m / 2] the split factors e^{-2 pi i k / n}, then per
// radix-4 pass (half size h) 3h entries laid out contiguously so that lanes running over j read consecutive doubles (ncu v2: the strided
// reads tc[j * stride] of a single shared table were 4- to 32-way bank conflicts): W_{2h}^j (j < h), then W_{4h}^j (j < 2h).
__host__ __device__ inline int tw_len(int n) {
  const int m = n >> 1;
  int lm = 0;
  while ((1 << lm) < m) ++lm;
  int len = m / 2 + 1;
  for (int s = lm & 1; s < lm; s += 2) len += 3 << s;
  return len;
}
// Layout: [0, m) cos, [m, 2m) -sin, [2m, 2m + 338) DCT, then the mel weights of every bin: up[i] = rising weight of the filter whose
// rising edge [bin[j], bin[j+1]) holds bin i, dn[i] = the falling weight of filter j - 1 over the same segment (base.get_filterbanks).
__global__ void __launch_bounds__(256) mfcc_tables_kernel(int n, MfccGeom g, double* __restrict__ tab) {
  pdl_trigger();
  pdl_wait();
  const int m = n >> 1, lt = tw_len(n), lm = g.log2n - 1;
  const int tid = blockIdx.x * blockDim.x + threadIdx.x, nth = gridDim.x * blockDim.x;
  for (int k = tid; k <= m / 2; k += nth) {
    double sn, cs;
    sincospi(2.0 * (double)k / (double)n, &sn, &cs);
    tab[k] = cs;
    tab[lt + k] = -sn;
  }
  int off = m / 2 + 1;
  for (int s = lm & 1; s < lm; s += 2) {
    const int h = 1 << s;
    for (int t = tid; t < 3 * h; t += nth) {
      double sn, cs;
      if (t < h) sincospi((double)t / (double)h, &sn, &cs);             // W_{2h}^t = e^{-2 pi i t / (2h)}
      else sincospi((double)(t - h) / (double)(2 * h), &sn, &cs);       // W_{4h}^{t-h}
      tab[off + t] = cs;
      tab[lt + off + t] = -sn;
    }
    off += 3 * h;
  }
  for (int i = tid; i < DCT_TABLE; i += nth) {
    const int k = i / NFILT, f = i % NFILT;
    double v = cospi((double)k * (double)(2 * f + 1) / (double)(2 * NFILT));
    v *= 2.0 * (k == 0 ? sqrt(1.0 / (4.0 * NFILT)) : sqrt(1.0 / (2.0 * NFILT)));  // scipy.fftpack.dct(type=2, norm='ortho')
    v *= 1.0 + (CEPLIFTER / 2.0) * sinpi((double)k / CEPLIFTER);                  // base.lifter, L = 22
    tab[2 * lt + i] = v;
  }
  double* up = tab + 2 * lt + DCT_TABLE;
  double* dn = up + (m + 1);
  for (int i = tid; i <= m; i += nth) {
    double u = 0.0, d = 0.0;
    for (int j = 0; j <= NFILT; ++j) {
      const int b0 = g.bin[j], b1 = g.bin[j + 1];
      if (i >= b0 && i < b1) {
        u = (double)(i - b0) / (double)(b1 - b0);
        d = (double)(b1 - i) / (double)(b1 - b0);
      }
    }
    up[i] = u;
    dn[i] = d;
  }
}

template <bool I16>
__global__ void __launch_bounds__(FFT_THREADS) mfcc_frame_kernel(const void* __restrict__ signal_, long long n_samples, long long clip_stride,
                                                                 int lag, MfccGeom g, const double* __restrict__ tab,
                                                                 double* __restrict__ cep) {
  pdl_trigger();
  pdl_wait();
  extern __shared__ double sm[];
  const int n = g.nfft, m = n >> 1, lm = g.log2n - 1;
  const int zlen = m + (m >> 4) + 1, lt = tw_len(n);
  double* zr = sm;
  double* zi = sm + zlen;
  double* tc = sm + 2 * zlen;  // 3m + 2 doubles: staging area of the frame first (n + n / 128 + 1), then the twiddles (cos | -sin)
  double* ts = tc + lt;
  double* stage = tc;
  __shared__ double red[FFT_THREADS / 32];
  __shared__ double fb[NFILT];
  __shared__ double energy_s;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const long long clip = blockIdx.x / g.numframes, frame = blockIdx.x % g.numframes;
  const long long start = frame * g.frame_step;

  // pre-emphasised frame.  sigproc.preemphasis: append(x[0], x[1:] - 0.97 x[:-1]) acts per channel ROW of an [n, channels] array and
  // is then flattened: in interleaved samples y[p] = x[p] - 0.97 x[p - lag] (p >= lag), y[p] = x[p] (p < lag), lag = channels.
  // Zero padding is appended AFTER the pre-emphasis.  Sample 2i goes to Re z[i], 2i+1 to Im z[i], z stored in bit-reversed order.
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
    stage[PS(i)] = v;
  }
  __syncthreads();
  for (int r = tid; r < m; r += FFT_THREADS) {
    const int i = (int)(__brev((unsigned)r) >> (32 - lm));
    zr[PD(r)] = stage[PS(2 * i)];
    zi[PD(r)] = stage[PS(2 * i + 1)];
  }
  __syncthreads();
  for (int k = tid; k < 2 * lt; k += FFT_THREADS) tc[k] = tab[k];  // cos in tc[0..lt), -sin in ts[0..lt)
  __syncthreads();
  int s0 = 0;
  if (lm & 1) {  // radix-2 pass (stage 0: twiddle 1)
    for (int t = tid; t < (m >> 1); t += FFT_THREADS) {
      const int i0 = PD(2 * t), i1 = PD(2 * t + 1);
      const double ar = zr[i0], ai = zi[i0], br = zr[i1], bi = zi[i1];
      zr[i0] = ar + br; zi[i0] = ai + bi;
      zr[i1] = ar - br; zi[i1] = ai - bi;
    }
    __syncthreads();
    s0 = 1;
  }
  int poff = m / 2 + 1;
  for (int s = s0; s < lm; poff += 3 << s, s += 2) {  // radix-4 pass = stages s and s + 1 (half sizes h and 2h)
    const int h = 1 << s;
    const int groups = m >> (s + 2);
    for (int q = tid; q < (m >> 2); q += FFT_THREADS) {
      // h >= 16: lanes over j (16 consecutive doubles); h < 16: lanes over groups (stride 4h doubles, spread over the banks by PD)
      const int j = h >= 16 ? (q & (h - 1)) : (q >> (lm - s - 2)), blk = h >= 16 ? (q >> s) : (q & (groups - 1));
      const int b0 = (blk << (s + 2)) + j;
      const int i0 = PD(b0), i1 = PD(b0 + h), i2 = PD(b0 + 2 * h), i3 = PD(b0 + 3 * h);
      const double w1r = tc[poff + j], w1i = ts[poff + j];                  // W_{2h}^j
      const double w2r = tc[poff + h + j], w2i = ts[poff + h + j];          // W_{4h}^j
      const double w3r = tc[poff + 2 * h + j], w3i = ts[poff + 2 * h + j];  // W_{4h}^{j+h}
      const double x0r = zr[i0], x0i = zi[i0], x1r = zr[i1], x1i = zi[i1], x2r = zr[i2], x2i = zi[i2], x3r = zr[i3], x3i = zi[i3];
      double tr = x1r * w1r - x1i * w1i, ti = x1r * w1i + x1i * w1r;
      const double a0r = x0r + tr, a0i = x0i + ti, a1r = x0r - tr, a1i = x0i - ti;
      tr = x3r * w1r - x3i * w1i; ti = x3r * w1i + x3i * w1r;
      const double a2r = x2r + tr, a2i = x2i + ti, a3r = x2r - tr, a3i = x2i - ti;
      tr = a2r * w2r - a2i * w2i; ti = a2r * w2i + a2i * w2r;
      zr[i0] = a0r + tr; zi[i0] = a0i + ti;
      zr[i2] = a0r - tr; zi[i2] = a0i - ti;
      tr = a3r * w3r - a3i * w3i; ti = a3r * w3i + a3i * w3r;
      zr[i1] = a1r + tr; zi[i1] = a1i + ti;
      zr[i3] = a1r - tr; zi[i3] = a1i - ti;
    }
    __syncthreads();
  }
  // split + power spectrum (sigproc.powspec: |rfft|^2 / nfft) for bins 0..m, pair (k, m - k) per thread, written over zr (ps[m] in zr[m])
  double* ps = zr;
  double e = 0.0;
  const double inv_n = 1.0 / (double)n;
  for (int k = tid; k <= (m >> 1); k += FFT_THREADS) {
    const int k2 = (m - k) & (m - 1);
    const double ar = zr[PD(k)], ai = zi[PD(k)], br = zr[PD(k2)], bi = -zi[PD(k2)];  // A = Z[k], B = conj(Z[m - k])
    const double er = 0.5 * (ar + br), ei = 0.5 * (ai + bi);
    const double dr = 0.5 * (ar - br), di = 0.5 * (ai - bi);  // O = -i (A - B) / 2 = (di, -dr)
    const double orr = di, oi = -dr;
    const double c = tc[k], sn = ts[k];  // (cos, -sin) of 2 pi k / n
    // X[k] = E + W O
    const double xr = er + (orr * c - oi * sn), xi = ei + (orr * sn + oi * c);
    // X[m - k]: E' = conj(E), O' = conj(O), W' = (-c, sn)  [e^{-2 pi i (m - k) / n} = -conj(W)]
    const double yr = er + (orr * (-c) - (-oi) * sn), yi = -ei + (orr * sn + (-oi) * (-c));
    const double p0 = (xr * xr + xi * xi) * inv_n, p1 = (yr * yr + yi * yi) * inv_n;  // n is a power of two: same as dividing
    if (k == 0) {  // X[0] and X[m] both come from Z[0]
      ps[0] = p0; ps[PD(m)] = p1; e += p0 + p1;
    } else if (k == k2) {
      ps[PD(k)] = p0; e += p0;
    } else {
      ps[PD(k)] = p0; ps[PD(m - k)] = p1; e += p0 + p1;
    }
  }
  e = warp_sum_d(e);
  if (lane == 0) red[wid] = e;
  __syncthreads();
  if (tid == 0) {
    double s = 0.0;
    for (int i = 0; i < FFT_THREADS / 32; ++i) s += red[i];
    energy_s = s == 0.0 ? F64_EPS : s;
  }
  // mel filterbank: filter j rises over [bin[j], bin[j+1]) and falls over [bin[j+1], bin[j+2]) (weights from the table)
  const double* mel_up = tab + 2 * lt + DCT_TABLE;
  const double* mel_dn = mel_up + (m + 1);
  for (int j = wid; j < NFILT; j += FFT_THREADS / 32) {
    const int b0 = g.bin[j], b1 = g.bin[j + 1], b2 = g.bin[j + 2];
    double acc = 0.0;
    for (int i = b0 + lane; i < b2; i += 32) acc += ps[PD(i)] * (i < b1 ? mel_up[i] : mel_dn[i]);
    acc = warp_sum_d(acc);
    if (lane == 0) fb[j] = log(acc == 0.0 ? F64_EPS : acc);
  }
  __syncthreads();
  // DCT-II, norm='ortho' (scipy.fftpack.dct), first 13 coefficients, lifter 1 + 11 sin(pi k / 22) (both in the table); c0 := log(energy)
  const double* dct = tab + 2 * lt;
  for (int k = wid; k < NCEP; k += FFT_THREADS / 32) {
    double s = lane < NFILT ? fb[lane] * dct[k * NFILT + lane] : 0.0;
    s = warp_sum_d(s);
    if (k == 0) s = log(energy_s);
    if (lane == 0) cep[(clip * g.numframes + frame) * NCEP + k] = s;
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

// Horizontal pass (ImagingResampleHorizontal_32bpc), one warp per output column.  cep [frames, 13] (double) is the transposed
// source image: pixel (row y = cepstral index, column x = frame) as float32.  Pillow sums the weights and then the 13 rows' products
// sequentially in tap order; double addition is not associative, so those chains stay sequential here (every lane carries the same
// running sum), while what feeds them — the bicubic weights, their division by the normaliser, the pixel loads — is spread over the
// 32 lanes: lanes evaluate 32 taps at a time, lanes 0..12 each own one image row.
__global__ void __launch_bounds__(256) mfcc_resample_h_kernel(const double* __restrict__ cep, long long numframes, int W,
                                                              float* __restrict__ tmp_all) {
  pdl_trigger();
  pdl_wait();
  __shared__ float pix[8][32 * NCEP];
  const int lane = threadIdx.x & 31;
  const int xx = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (xx >= W) return;
  const double* src = cep + (size_t)blockIdx.y * numframes * NCEP;
  float* tmp = tmp_all + (size_t)blockIdx.y * NCEP * W;  // [13, W]
  if (numframes == (long long)W) {  // Pillow skips the pass when the width is unchanged
    if (lane < NCEP) tmp[lane * W + xx] = (float)src[(size_t)xx * NCEP + lane];
    return;
  }
  const ResampleAxis ax = make_axis((int)numframes, W);
  const double center = __dmul_rn(__dadd_rn((double)xx, 0.5), ax.scale);
  int xmin = (int)__dadd_rn(__dsub_rn(center, ax.support), 0.5);
  if (xmin < 0) xmin = 0;
  int xmax = (int)__dadd_rn(__dadd_rn(center, ax.support), 0.5);
  if (xmax > ax.in_size) xmax = ax.in_size;
  const int cnt = xmax - xmin;
  double ww = 0.0;
  for (int base = 0; base < cnt; base += 32) {
    const int x = base + lane;
    const double w = x < cnt ? bicubic_filter(__dmul_rn(__dadd_rn(__dsub_rn((double)(x + xmin), center), 0.5), ax.ss)) : 0.0;
    const int m = min(32, cnt - base);
    for (int l = 0; l < m; ++l) ww = __dadd_rn(ww, __shfl_sync(0xffffffffu, w, l));
  }
  double ss = 0.0;
  float* px = pix[threadIdx.x >> 5];
  for (int base = 0; base < cnt; base += 32) {
    const int x = base + lane;
    double k = x < cnt ? bicubic_filter(__dmul_rn(__dadd_rn(__dsub_rn((double)(x + xmin), center), 0.5), ax.ss)) : 0.0;
    if (ww != 0.0) k = __ddiv_rn(k, ww);
    const int m = min(32, cnt - base);
    // the chunk's m x 13 source pixels are contiguous: one coalesced sweep into shared memory (as the float32 image Pillow sees)
    for (int i = lane; i < m * NCEP; i += 32) px[i] = (float)src[(size_t)(xmin + base) * NCEP + i];
    __syncwarp();
    for (int l = 0; l < m; ++l) {
      const double kk = __shfl_sync(0xffffffffu, k, l);
      if (lane < NCEP) ss = __dadd_rn(ss, __dmul_rn((double)px[l * NCEP + lane], kk));
    }
    __syncwarp();
  }
  if (lane < NCEP) tmp[lane * W + xx] = (float)ss;
}

// Vertical pass + min-max normalisation, one CTA per clip.
__global__ void __launch_bounds__(512) mfcc_image_kernel(int W, int H, const float* __restrict__ tmp_all, float* __restrict__ image_all) {
  pdl_trigger();
  pdl_wait();
  __shared__ float smin[16], smax[16];
  extern __shared__ double vk[];  // [H][13] normalised weights of each output row, then ymin / count as ints
  int* vb = reinterpret_cast<int*>(vk + (size_t)H * NCEP);
  const float* tmp = tmp_all + (size_t)blockIdx.x * NCEP * W;   // [13, W] after the horizontal pass
  float* img = image_all + (size_t)blockIdx.x * H * W;          // [H, W]
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const bool need_v = H != NCEP;
  // vertical pass (ImagingResampleVertical_32bpc) + running min / max
  float mn = INFINITY, mx = -INFINITY;
  {
    const ResampleAxis ax = make_axis(NCEP, H);
    if (need_v)
      for (int yy = tid; yy < H; yy += blockDim.x) {  // precompute_coeffs of the vertical axis: at most 13 taps (the source has 13 rows)
        double center, ww; int ymin, cnt;
        axis_bounds(ax, yy, center, ymin, cnt, ww);
        for (int y = 0; y < cnt; ++y) vk[yy * NCEP + y] = axis_coeff(ax, center, ymin, y, ww);
        vb[2 * yy] = ymin; vb[2 * yy + 1] = cnt;
      }
    __syncthreads();
    for (int o = tid; o < H * W; o += blockDim.x) {
      const int yy = o / W, xx = o % W;
      float r;
      if (need_v) {
        const int ymin = vb[2 * yy], cnt = vb[2 * yy + 1];
        double ss = 0.0;
        for (int y = 0; y < cnt; ++y) ss = __dadd_rn(ss, __dmul_rn((double)tmp[(y + ymin) * W + xx], vk[yy * NCEP + y]));
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

inline size_t frame_smem_doubles(int nfft) {  // zr, zi (m + m / 16 + 1 each) + staging / twiddle area (n + n / 128 + 1)
  const size_t m = (size_t)nfft / 2;
  const size_t stage = (size_t)nfft + (size_t)nfft / 128 + 1, tw = 2 * (size_t)tw_len(nfft);
  return 2 * (m + m / 16 + 1) + (stage > tw ? stage : tw);
}
inline size_t table_doubles(int nfft) { return 2 * (size_t)tw_len(nfft) + DCT_TABLE + 2 * ((size_t)nfft / 2 + 1); }

int launch_cepstra(cudaStream_t st, const void* signal, int is_i16, int channels, int n_clips, long long n_samples, long long clip_stride,
                   const MfccGeom& g, const double* tab, double* cep) {
  DTB_REQUIRE(channels >= 1 && channels <= 8, DTB_ERR_BAD_ARG, "mfcc: channels must be 1..8");
  const long long blocks = (long long)n_clips * g.numframes;
  DTB_REQUIRE(blocks <= 0x7fffffffLL, DTB_ERR_UNSUPPORTED, "mfcc: too many frames in one launch");
  const size_t smem = frame_smem_doubles(g.nfft) * sizeof(double);
  static bool attr_set = false;
  if (!attr_set) {
    DTB_CUDA(cudaFuncSetAttribute(mfcc_frame_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(frame_smem_doubles(MAX_NFFT) * sizeof(double))));
    DTB_CUDA(cudaFuncSetAttribute(mfcc_frame_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(frame_smem_doubles(MAX_NFFT) * sizeof(double))));
    attr_set = true;
  }
  if (is_i16)
    DTB_LAUNCH(mfcc_frame_kernel<true>, (unsigned)blocks, FFT_THREADS, smem, st, signal, n_samples, clip_stride, channels, g, tab, cep);
  else
    DTB_LAUNCH(mfcc_frame_kernel<false>, (unsigned)blocks, FFT_THREADS, smem, st, signal, n_samples, clip_stride, channels, g, tab, cep);
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
size_t dtb_mfcc_table_doubles(int samplerate) {
  MfccGeom g;
  if (make_geom(1, samplerate, g) != DTB_OK) return 0;
  return table_doubles(g.nfft);
}
size_t dtb_mfcc_workspace_bytes(int n_clips, long long n_samples, int samplerate, int out_w) {
  MfccGeom g;
  if (n_clips < 1 || out_w < 1 || make_geom(n_samples, samplerate, g) != DTB_OK) return 0;
  return al256(table_doubles(g.nfft) * sizeof(double)) + al256((size_t)n_clips * g.numframes * NCEP * sizeof(double)) +
         al256((size_t)n_clips * NCEP * out_w * sizeof(float)) + 256;
}
int dtb_mfcc_build_tables(void* stream, int samplerate, double* tables) {
  DTB_REQUIRE(tables != nullptr, DTB_ERR_BAD_ARG, "dtb_mfcc_build_tables: null argument");
  DTB_TRY(require_sm100());
  MfccGeom g;
  DTB_TRY(make_geom(1, samplerate, g));
  DTB_LAUNCH(mfcc_tables_kernel, 8, 256, 0, (cudaStream_t)stream, g.nfft, g, tables);
  DTB_CHECK_LAUNCH();
  return DTB_OK;
}
int dtb_mfcc_cepstra(void* stream, const void* signal, int is_i16, int channels, int n_clips, long long n_samples, long long clip_stride, int samplerate,
                     const double* tables, double* cepstra) {
  DTB_REQUIRE(signal && cepstra && tables && n_clips >= 1, DTB_ERR_BAD_ARG, "dtb_mfcc_cepstra: null argument");
  DTB_TRY(require_sm100());
  MfccGeom g;
  DTB_TRY(make_geom(n_samples, samplerate, g));
  return launch_cepstra((cudaStream_t)stream, signal, is_i16, channels, n_clips, n_samples, clip_stride, g, tables, cepstra);
}
int dtb_mfcc_resize_normalize(void* stream, const double* cepstra, int n_clips, long long numframes, int out_w, int out_h, float* scratch,
                              float* image) {
  DTB_REQUIRE(cepstra && scratch && image && n_clips >= 1 && numframes >= 1 && out_w >= 1 && out_h >= 1, DTB_ERR_BAD_ARG,
              "dtb_mfcc_resize_normalize: bad argument");
  DTB_REQUIRE(numframes <= 0x7fffffffLL / NCEP, DTB_ERR_UNSUPPORTED, "mfcc: clip too long");
  DTB_TRY(require_sm100());
  DTB_LAUNCH(mfcc_resample_h_kernel, dim3((out_w + 7) / 8, n_clips), 256, 0, (cudaStream_t)stream, cepstra, numframes, out_w, scratch);
  DTB_CHECK_LAUNCH();
  const size_t vsmem = (size_t)out_h * NCEP * sizeof(double) + (size_t)out_h * 2 * sizeof(int);
  DTB_REQUIRE(vsmem <= 200 * 1024, DTB_ERR_UNSUPPORTED, "mfcc: out_h %d too large", out_h);
  static size_t vsmem_set = 48 * 1024;
  if (vsmem > vsmem_set) {
    DTB_CUDA(cudaFuncSetAttribute(mfcc_image_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    vsmem_set = 200 * 1024;
  }
  DTB_LAUNCH(mfcc_image_kernel, n_clips, 512, vsmem, (cudaStream_t)stream, out_w, out_h, (const float*)scratch, image);
  DTB_CHECK_LAUNCH();
  return DTB_OK;
}
int dtb_wav_to_mfcc_image(void* stream, const void* signal, int is_i16, int channels, int n_clips, long long n_samples, long long clip_stride,
                          int samplerate, int out_w, int out_h, void* workspace, size_t workspace_bytes, const double* tables, float* image) {
  DTB_REQUIRE(signal && workspace && image && n_clips >= 1 && out_w >= 1 && out_h >= 1, DTB_ERR_BAD_ARG, "dtb_wav_to_mfcc_image: bad argument");
  DTB_TRY(require_sm100());
  MfccGeom g;
  DTB_TRY(make_geom(n_samples, samplerate, g));
  uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(workspace) + 255) & ~uintptr_t(255));
  const size_t tab_bytes = al256(table_doubles(g.nfft) * sizeof(double));
  const size_t cep_bytes = al256((size_t)n_clips * g.numframes * NCEP * sizeof(double));
  const size_t need = (size_t)(base - reinterpret_cast<uint8_t*>(workspace)) + tab_bytes + cep_bytes +
                      al256((size_t)n_clips * NCEP * out_w * sizeof(float));
  DTB_REQUIRE(need <= workspace_bytes, DTB_ERR_WORKSPACE, "mfcc: workspace too small (%zu bytes given, %zu needed)", workspace_bytes, need);
  const double* tab = tables;
  if (!tab) {  // no cached tables: build them into the workspace first
    DTB_TRY(dtb_mfcc_build_tables(stream, samplerate, reinterpret_cast<double*>(base)));
    tab = reinterpret_cast<const double*>(base);
  }
  double* cep = reinterpret_cast<double*>(base + tab_bytes);
  float* scratch = reinterpret_cast<float*>(base + tab_bytes + cep_bytes);
  DTB_TRY(launch_cepstra((cudaStream_t)stream, signal, is_i16, channels, n_clips, n_samples, clip_stride, g, tab, cep));
  return dtb_mfcc_resize_normalize(stream, cep, n_clips, g.numframes, out_w, out_h, scratch, image);
}

}  // extern "C"
