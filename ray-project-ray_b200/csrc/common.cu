// Error bookkeeping + small device queries shared by every b200x entry point.
#include "common.cuh"
#include "../../include/b200x.h"
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

namespace b200x {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int cuda_fail(cudaError_t e, const char* what) {
  set_error("CUDA error %d (%s) at %s", (int)e, cudaGetErrorString(e), what);
  // Clear the sticky-less error state so that the next call starts clean.
  cudaGetLastError();
  return -(int)e;
}

int sm_count() {
  static thread_local int cached_dev = -1;
  static thread_local int cached_sms = 148;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;
  if (dev != cached_dev) {
    int sms = 0;
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && sms > 0) {
      cached_sms = sms;
      cached_dev = dev;
    }
  }
  return cached_sms;
}

}  // namespace b200x

extern "C" {

int b200x_version(void) { return B200X_VERSION; }

const char* b200x_last_error_string(void) { return b200x::g_err; }

int b200x_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}

}  // extern "C"
