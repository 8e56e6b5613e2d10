// Error bookkeeping + small device queries shared by every b200x entry point.
#include "common.cuh"
#include "../../include/b200x.h"
#include <stdarg.h>
#include <stdio.h>
#include <string.h>
#include <atomic>
#include <utility>
#include <vector>

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

DeviceGuard::DeviceGuard(const void* ptr) {
  if (ptr == nullptr) return;
  cudaPointerAttributes attr;
  if (cudaPointerGetAttributes(&attr, ptr) != cudaSuccess) {
    cudaGetLastError();   // not a CUDA pointer: nothing to do, and no sticky error for the caller
    return;
  }
  if (attr.type != cudaMemoryTypeDevice && attr.type != cudaMemoryTypeManaged) return;
  if (cudaGetDevice(&prev) != cudaSuccess) return;
  if (attr.device != prev && cudaSetDevice(attr.device) == cudaSuccess) switched = true;
}

DeviceGuard::~DeviceGuard() {
  if (switched) cudaSetDevice(prev);
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

static std::atomic<unsigned long long> g_launches{0};
void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }

static thread_local bool t_prof = false;
static thread_local std::vector<std::pair<cudaEvent_t, cudaEvent_t>> t_events;
static thread_local std::vector<int> t_kernel;
static thread_local size_t t_used = 0;
static thread_local bool t_open = false;

void prof_begin(cudaStream_t s, int kernel) {
  if (!t_prof) return;
  if (t_used == t_events.size()) {
    cudaEvent_t a, b;
    if (cudaEventCreate(&a) != cudaSuccess || cudaEventCreate(&b) != cudaSuccess) {
      cudaGetLastError();
      return;
    }
    t_events.emplace_back(a, b);
    t_kernel.push_back(0);
  }
  t_kernel[t_used] = kernel;
  t_open = cudaEventRecord(t_events[t_used].first, s) == cudaSuccess;
}

void prof_end(cudaStream_t s) {
  if (!t_prof || !t_open) return;
  if (cudaEventRecord(t_events[t_used].second, s) == cudaSuccess) ++t_used;
  t_open = false;
}

}  // namespace b200x

extern "C" {

unsigned long long b200x_launch_count(void) { return b200x::g_launches.load(std::memory_order_relaxed); }

int b200x_profile_enable(int on) {
  b200x::t_prof = on != 0;
  b200x::t_used = 0;
  b200x::t_open = false;
  return B200X_OK;
}

int b200x_profile_read(int kernel, double* total_ms, int* launches) {
  double tot = 0.0;
  int cnt = 0;
  for (size_t i = 0; i < b200x::t_used; ++i) {
    if (b200x::t_kernel[i] != kernel) continue;
    float ms = 0.f;
    B200X_CUDA(cudaEventSynchronize(b200x::t_events[i].second));
    B200X_CUDA(cudaEventElapsedTime(&ms, b200x::t_events[i].first, b200x::t_events[i].second));
    // a kernel that returns at once because the device chose the other path (a few microseconds) is not counted
    if (ms > 0.1f) {
      tot += ms;
      ++cnt;
    }
  }
  if (total_ms) *total_ms = tot;
  if (launches) *launches = cnt;
  return B200X_OK;
}

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
