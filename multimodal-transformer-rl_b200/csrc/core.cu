// Error plumbing + device queries shared by every translation unit of libdtb200.so.
#include "common.cuh"

#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

namespace dtb {

static thread_local char g_err[1024] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

const char* last_error() { return g_err; }

int cuda_fail(cudaError_t e, const char* what, const char* file, int line) {
  set_error("CUDA error %d (%s) at %s:%d: %s", (int)e, cudaGetErrorString(e), file, line, what);
  return DTB_ERR_CUDA;
}

bool pdl_enabled() {
  static int on = -1;
  if (on < 0) {
    const char* e = getenv("DTB_NO_PDL");
    on = (e && e[0] == '1') ? 0 : 1;
  }
  return on == 1;
}

// ---- runtime options (dtb_set_option): A/B switches between kernel forms inside one process; defaults from the environment ---------
static int env_flag(const char* name, int dflt) {
  const char* e = getenv(name);
  return e ? (e[0] == '1' ? 1 : 0) : dflt;
}
struct Options {
  int splitk_inkernel = env_flag("DTB_SPLITK_INKERNEL", 0);  // split-K slices reduced by the last-arriving CTA inside the GEMM
  int ln_fused = !env_flag("DTB_LN_SEPARATE", 0);            // LayerNorm in the c_proj GEMM epilogue (H <= 256)
  int conv_implicit = !env_flag("DTB_CONV_EXPLICIT", 0);     // implicit-GEMM conv2 / conv3 (read when an encoder is created)
  double balance = 212.0;                                    // machine balance (flop / byte) that splits the GEMM profiling families
};
static Options& opts() { static Options o; return o; }
int opt_splitk_inkernel() { return opts().splitk_inkernel; }
int opt_ln_fused() { return opts().ln_fused; }
int opt_conv_implicit() { return opts().conv_implicit; }
double opt_balance() { return opts().balance; }
void opt_set_balance(double b) { if (b > 0) opts().balance = b; }
int opt_set(const char* name, int value) {
  if (!strcmp(name, "splitk_inkernel")) opts().splitk_inkernel = value != 0;
  else if (!strcmp(name, "ln_fused")) opts().ln_fused = value != 0;
  else if (!strcmp(name, "conv_implicit")) opts().conv_implicit = value != 0;
  else { set_error("dtb_set_option: unknown option '%s'", name); return DTB_ERR_BAD_ARG; }
  return DTB_OK;
}

static long long g_launches = 0;
void count_launch() { __atomic_add_fetch(&g_launches, 1, __ATOMIC_RELAXED); }
long long launch_count() { return __atomic_load_n(&g_launches, __ATOMIC_RELAXED); }

// ---- live profiling ------------------------------------------------------------------------------------------------
static bool g_prof = false;
static const int PROF_SLOTS = 8192;
static cudaEvent_t g_pe0[PROF_SLOTS], g_pe1[PROF_SLOTS];
static int g_pfam[PROF_SLOTS];
static double g_pwork[PROF_SLOTS];
static int g_pn = 0, g_pcreated = 0;
bool prof_on() { return g_prof; }
void prof_enable(bool on) { g_prof = on; }
void prof_reset() { g_pn = 0; }
int prof_begin(cudaStream_t stream, int family, double work) {
  if (g_pn >= PROF_SLOTS) return -1;
  cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
  if (cudaStreamIsCapturing(stream, &cs) != cudaSuccess) return -1;
  const int i = g_pn++;
  if (i >= g_pcreated) {
    cudaEventCreate(&g_pe0[i]);
    cudaEventCreate(&g_pe1[i]);
    g_pcreated = i + 1;
  }
  g_pfam[i] = family;
  g_pwork[i] = work;
  // inside a stream capture the record becomes an external event-record node: every graph replay re-times the kernel in place
  cudaEventRecordWithFlags(g_pe0[i], stream, cs == cudaStreamCaptureStatusNone ? cudaEventRecordDefault : cudaEventRecordExternal);
  return i;
}
void prof_end(cudaStream_t stream, int slot) {
  cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
  cudaStreamIsCapturing(stream, &cs);
  cudaEventRecordWithFlags(g_pe1[slot], stream, cs == cudaStreamCaptureStatusNone ? cudaEventRecordDefault : cudaEventRecordExternal);
}
int prof_report(int family, double* total_ms, double* work, long long* launches) {
  double t = 0, w = 0;
  long long n = 0;
  for (int i = 0; i < g_pn; ++i) {
    if (g_pfam[i] != family) continue;
    if (cudaEventSynchronize(g_pe1[i]) != cudaSuccess) return DTB_ERR_CUDA;
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, g_pe0[i], g_pe1[i]) != cudaSuccess) return DTB_ERR_CUDA;
    t += ms; w += g_pwork[i]; ++n;
  }
  if (total_ms) *total_ms = t;
  if (work) *work = w;
  if (launches) *launches = n;
  return DTB_OK;
}

// span i of the last profiled step: family and [start, end] in ms relative to the first recorded launch (timeline reconstruction)
int prof_span(int i, int* family, double* start_ms, double* end_ms) {
  if (i < 0 || i >= g_pn) return DTB_ERR_BAD_ARG;
  if (cudaEventSynchronize(g_pe1[i]) != cudaSuccess) return DTB_ERR_CUDA;
  float a = 0.f, b = 0.f;
  if (cudaEventElapsedTime(&a, g_pe0[0], g_pe0[i]) != cudaSuccess) return DTB_ERR_CUDA;
  if (cudaEventElapsedTime(&b, g_pe0[0], g_pe1[i]) != cudaSuccess) return DTB_ERR_CUDA;
  if (family) *family = g_pfam[i];
  if (start_ms) *start_ms = a;
  if (end_ms) *end_ms = b;
  return DTB_OK;
}
int prof_count() { return g_pn; }

static int g_num_sms = 0;
int num_sms() {
  if (g_num_sms > 0) return g_num_sms;
  int dev = 0, n = 0;
  if (cudaGetDevice(&dev) == cudaSuccess && cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && n > 0)
    g_num_sms = n;
  else
    g_num_sms = DTB_NUM_SMS_FALLBACK;
  return g_num_sms;
}

// 0 when the current device is compute capability 10.x, error otherwise (no fallback path exists).
int require_sm100() {
  int dev = 0, major = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return cuda_fail(e, "cudaGetDevice", __FILE__, __LINE__);
  e = cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
  if (e != cudaSuccess) return cuda_fail(e, "cudaDeviceGetAttribute", __FILE__, __LINE__);
  if (major != 10) {
    set_error("dtb200 requires an sm_100a (Blackwell B200) device; found compute capability major %d", major);
    return DTB_ERR_ARCH;
  }
  return DTB_OK;
}

}  // namespace dtb
