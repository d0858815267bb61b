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
extern std::atomic<int> g_sm_budget;   // SMs the persistent transform kernels may occupy (0: all)
extern std::atomic<int> g_tiles_per_cta;  // transform kernels: tiles per CTA (0: persistent grid)
extern std::atomic<int> g_coder_lanes; // unbounded coder, one stream per lane: 0 auto, 1 off, 2 on
extern std::atomic<int> g_coder_pack;  // range-coder streams per CTA (0: one CTA per stream when
                                       // the SMs allow it — lowest latency)

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

inline int num_sms();
// grid size of a persistent kernel: every SM, or the configured budget (tfc_configure)
inline int persistent_grid() {
  const int b = g_sm_budget.load(std::memory_order_relaxed);
  const int n = num_sms();
  return (b > 0 && b < n) ? b : n;
}

extern std::atomic<int> g_debug;
extern std::atomic<int> g_conv_cluster;
inline int conv_cluster() { return g_conv_cluster.load(std::memory_order_relaxed); }
inline int debug_bits() { return g_debug.load(std::memory_order_relaxed); }
inline int tiles_per_cta() { return g_tiles_per_cta.load(std::memory_order_relaxed); }

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
