// Shared host-side plumbing for libtfc_b200.so: error text, launch accounting, CUDA checks.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <string>

#include "tfc_b200.h"

namespace tfc {

std::string& last_error_ref();
extern std::atomic<int64_t> g_launches;

inline int fail(int code, const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  last_error_ref() = buf;
  return code;
}

inline int cuda_fail(cudaError_t e, const char* what) {
  return fail(TFC_ERR_CUDA, "CUDA failure in %s: %s", what, cudaGetErrorString(e));
}

#define TFC_CUDA(expr)                                   \
  do {                                                   \
    cudaError_t e__ = (expr);                            \
    if (e__ != cudaSuccess) return ::tfc::cuda_fail(e__, #expr); \
  } while (0)

// call after every kernel launch
#define TFC_LAUNCHED(name)                                        \
  do {                                                            \
    ::tfc::g_launches.fetch_add(1, std::memory_order_relaxed);    \
    cudaError_t e__ = cudaGetLastError();                         \
    if (e__ != cudaSuccess) return ::tfc::cuda_fail(e__, name);   \
  } while (0)

inline int num_sms() {
  static int sms = 0;
  if (sms == 0) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess ||
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess)
      sms = 148;
  }
  return sms;
}

}  // namespace tfc
