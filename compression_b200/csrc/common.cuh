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
extern std::atomic<int> g_conv_pair;       // implicit GEMM with cta_group::2 MMAs: 0 auto, 1 off, 2 on
extern std::atomic<int> g_conv_stages;     // implicit-GEMM operand ring: 0 auto (4), 2..4
extern std::atomic<int> g_coder_smem_kb;   // dynamic shared memory a warps-per-stream coder CTA may take (0: 226 KB)
extern std::atomic<int> g_coder_table_kb;  // shared-memory budget of the coder's staged CDF tables (0: 72 KB)
inline int conv_cluster() { return g_conv_cluster.load(std::memory_order_relaxed); }
inline int debug_bits() { return g_debug.load(std::memory_order_relaxed); }
inline int tiles_per_cta() { return g_tiles_per_cta.load(std::memory_order_relaxed); }

// SM count of the CURRENT device (cached per device: a process may drive several GPUs)
inline int num_sms() {
  static std::atomic<int> cache[64];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;
  std::atomic<int>& slot = cache[dev & 63];
  int sms = slot.load(std::memory_order_relaxed);
  if (sms == 0) {
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms <= 0)
      sms = 148;
    slot.store(sms, std::memory_order_relaxed);
  }
  return sms;
}

// "done once" flag kept PER DEVICE: function attributes (cudaFuncSetAttribute) belong to the
// device that was current when they were set, so a process-wide static bool would leave every
// other GPU of the process without them.  Thread-safe.
struct PerDeviceOnce {
  std::atomic<uint64_t> bits{0};
  // true the first time it is called with this device current
  bool first() {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return true;
    const uint64_t m = 1ull << (dev & 63);
    return (bits.fetch_or(m, std::memory_order_relaxed) & m) == 0;
  }
};

// Length of an encoded string as a decoder may read it: the caller's nbytes clamped to the slab
// row (an encode that overflowed its slab reports the size it NEEDED, which is larger than cap;
// reading that many bytes would run into the next stream's row or past the allocation).
__host__ __device__ inline uint32_t clamp_stream_len(int32_t nbytes, int64_t cap) {
  const int64_t n = nbytes < 0 ? 0 : (int64_t)nbytes;
  return (uint32_t)(n < cap ? n : cap);
}

}  // namespace tfc
