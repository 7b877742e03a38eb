// Error plumbing + device queries shared by every translation unit of libdtb200.so.
#include "common.cuh"

#include <stdarg.h>
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

static long long g_launches = 0;
void count_launch() { __atomic_add_fetch(&g_launches, 1, __ATOMIC_RELAXED); }
long long launch_count() { return __atomic_load_n(&g_launches, __ATOMIC_RELAXED); }

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
