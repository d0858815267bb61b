// PmfToQuantizedCdf on the device: one thread per row, bit-exact with
// cc/kernels/range_coding_helper_kernels.cc:97-201 as built with this image's libstdc++/glibc.
//
// What makes the reference's result implementation-defined, and how each piece is pinned:
//   std::rint(float)        -> rintf (same IEEE round-half-even)
//   std::log2(int) (double) -> lookup table filled ON THE HOST with glibc's log2 at first use
//   penalty/gain            -> IEEE double subtract + multiply (exact on both sides)
//   std::sort tie order     -> StdSortReplay (stdsort_replay.cuh)
//   find_if + rotate        -> the same linear scan + shift
#include <algorithm>
#include <cmath>
#include <mutex>
#include <vector>

#include "common.cuh"
#include "stdsort_replay.cuh"

namespace {

constexpr int kLutSize = (1 << 17) + 2;
__device__ double g_log2_lut[kLutSize];

__device__ __forceinline__ double lut_log2(int32_t v) {
  return (v >= 0 && v < kLutSize) ? g_log2_lut[v] : log2((double)v);
}

// scratch layout: item j of row r lives at ws[j * stride + r] (neighbouring threads = rows)
struct RowAcc {
  tfc::SortItem* base;
  int64_t stride;
  __device__ __forceinline__ tfc::SortItem get(int i) const { return base[(int64_t)i * stride]; }
  __device__ __forceinline__ void set(int i, const tfc::SortItem& v) const {
    base[(int64_t)i * stride] = v;
  }
};

__device__ __forceinline__ double steal_cost(double mass, int32_t v) {
  if (v <= 1) return INFINITY;
  return mass * (lut_log2(v) - lut_log2(v - 1));
}
__device__ __forceinline__ double give_gain(double mass, int32_t v) {
  if (v < 1) return -INFINITY;
  return mass * (lut_log2(v + 1) - lut_log2(v));
}

template <bool OVER>
__device__ void rebalance(RowAcc acc, const float* pmf, int32_t* c, int n, int todo) {
  for (int i = 0; i < n; ++i) {
    tfc::SortItem it;
    it.idx = i; it.pad = 0;
    it.key = OVER ? steal_cost((double)pmf[i], c[i]) : give_gain((double)pmf[i], c[i]);
    acc.set(i, it);
  }
  tfc::StdSortReplay<RowAcc, !OVER> sorter(acc);
  sorter.sort(n);
  for (; todo > 0; --todo) {
    tfc::SortItem head = acc.get(0);
    c[head.idx] += OVER ? -1 : 1;
    head.key = OVER ? steal_cost((double)pmf[head.idx], c[head.idx])
                    : give_gain((double)pmf[head.idx], c[head.idx]);
    int j = 1;
    for (; j < n; ++j) {
      const tfc::SortItem o = acc.get(j);
      if (OVER ? head.key < o.key : head.key > o.key) break;
      acc.set(j - 1, o);
    }
    acc.set(j - 1, head);
  }
}

__global__ void pmf_to_cdf_kernel(const float* __restrict__ pmf, int64_t rows, int n,
                                  int precision, int32_t* __restrict__ cdf,
                                  tfc::SortItem* ws, int64_t stride) {
  const int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (r >= rows) return;
  const float* p = pmf + r * n;
  int32_t* c = cdf + r * (int64_t)(n + 1) + 1;
  const int32_t total = 1 << precision;
  int32_t sum = 0;
  for (int i = 0; i < n; ++i) {
    const int32_t v = max((int32_t)rintf(p[i] * (float)total), 1);
    c[i] = v;
    sum += v;
  }
  RowAcc acc{ws + r, stride};
  if (sum > total) rebalance<true>(acc, p, c, n, sum - total);
  else if (sum < total) rebalance<false>(acc, p, c, n, total - sum);
  int32_t run = 0;
  c[-1] = 0;
  for (int i = 0; i < n; ++i) { run += c[i]; c[i] = run; }
}

// ---- long rows: one WARP per row, everything in shared memory --------------------------------
// One thread per row leaves the GPU empty when there are few long rows (a Gaussian scale table
// is 64 rows of up to 1481 entries) and walks its row through global memory one dependent load
// at a time.  Here the row's items {key, index, count} live in shared memory; the keys are
// computed and the result is scanned by the whole warp; lane 0 replays std::sort (its decision
// sequence is what fixes the order of equal keys and cannot be reordered); the re-insertion of
// the adjusted head — `find_if` + `rotate` in the reference (:165-181, :182-199) — is a ballot
// over 32 candidates at a time and a cooperative shift.
struct SmemAcc {
  tfc::SortItem* base;
  __device__ __forceinline__ tfc::SortItem get(int i) const { return base[i]; }
  __device__ __forceinline__ void set(int i, const tfc::SortItem& v) const { base[i] = v; }
};

constexpr int kWarpRowBytesPerEntry = (int)sizeof(tfc::SortItem) + 8;   // item + pmf + count

template <bool OVER>
__device__ void rebalance_warp(tfc::SortItem* it, const float* sp, int32_t* cnt, int n, int todo,
                               int lane) {
  for (int i = lane; i < n; i += 32) {
    tfc::SortItem v;
    v.idx = i; v.pad = cnt[i];
    v.key = OVER ? steal_cost((double)sp[i], cnt[i]) : give_gain((double)sp[i], cnt[i]);
    it[i] = v;
  }
  __syncwarp();
  if (lane == 0) {
    tfc::StdSortReplay<SmemAcc, !OVER> sorter(SmemAcc{it});
    sorter.sort(n);
  }
  __syncwarp();
  for (; todo > 0; --todo) {
    tfc::SortItem head = it[0];
    head.pad += OVER ? -1 : 1;
    head.key = OVER ? steal_cost((double)sp[head.idx], head.pad)
                    : give_gain((double)sp[head.idx], head.pad);
    int pos = n;
    for (int base = 1; base < n; base += 32) {
      const int j = base + lane;
      tfc::SortItem o;
      bool stop = false;
      if (j < n) {
        o = it[j];
        stop = OVER ? head.key < o.key : head.key > o.key;
      }
      const unsigned m = __ballot_sync(0xFFFFFFFFu, stop);
      const int lim = m ? base + __ffs(m) - 1 : min(base + 32, n);
      __syncwarp();
      if (j < lim) it[j - 1] = o;
      __syncwarp();
      if (m) { pos = lim; break; }
    }
    if (lane == 0) it[pos - 1] = head;
    __syncwarp();
  }
  for (int i = lane; i < n; i += 32) cnt[it[i].idx] = it[i].pad;
  __syncwarp();
}

__global__ void pmf_to_cdf_warp_kernel(const float* __restrict__ pmf, int64_t rows, int n,
                                       int precision, int32_t* __restrict__ cdf) {
  extern __shared__ __align__(16) uint8_t smem_rows[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, wpc = blockDim.x >> 5;
  const int64_t r = (int64_t)blockIdx.x * wpc + warp;
  if (r >= rows) return;
  const int npad = (n + 3) & ~3;
  uint8_t* mine = smem_rows + (size_t)warp * npad * kWarpRowBytesPerEntry;
  tfc::SortItem* it = reinterpret_cast<tfc::SortItem*>(mine);
  float* sp = reinterpret_cast<float*>(it + npad);
  int32_t* cnt = reinterpret_cast<int32_t*>(sp + npad);
  const float* p = pmf + r * n;
  int32_t* c = cdf + r * (int64_t)(n + 1);
  const int32_t total = 1 << precision;
  int32_t sum = 0;
  for (int i = lane; i < n; i += 32) {
    const float pv = p[i];
    const int32_t v = max((int32_t)rintf(pv * (float)total), 1);
    sp[i] = pv;
    cnt[i] = v;
    sum += v;
  }
  sum = __reduce_add_sync(0xFFFFFFFFu, sum);
  __syncwarp();
  if (sum > total) rebalance_warp<true>(it, sp, cnt, n, sum - total, lane);
  else if (sum < total) rebalance_warp<false>(it, sp, cnt, n, total - sum, lane);
  // cdf = [0, cumsum(count)]
  int32_t run = 0;
  if (lane == 0) c[0] = 0;
  for (int base = 0; base < n; base += 32) {
    const int i = base + lane;
    int32_t v = i < n ? cnt[i] : 0;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int32_t u = __shfl_up_sync(0xFFFFFFFFu, v, o);
      if (lane >= o) v += u;
    }
    if (i < n) c[i + 1] = run + v;
    run += __shfl_sync(0xFFFFFFFFu, v, 31);
  }
}

int ensure_lut() {
  static std::mutex mu;
  static std::vector<int> done;  // per-device flag
  std::lock_guard<std::mutex> lock(mu);
  int dev = 0;
  TFC_CUDA(cudaGetDevice(&dev));
  if ((int)done.size() <= dev) done.resize(dev + 1, 0);
  if (done[dev]) return TFC_OK;
  std::vector<double> lut(kLutSize);
  lut[0] = -INFINITY;
  for (int i = 1; i < kLutSize; ++i) lut[i] = std::log2((double)i);  // glibc, as the reference
  TFC_CUDA(cudaMemcpyToSymbol(g_log2_lut, lut.data(), sizeof(double) * kLutSize));
  done[dev] = 1;
  return TFC_OK;
}

}  // namespace

extern "C" size_t tfc_pmf_to_quantized_cdf_workspace(int64_t rows, int32_t n) {
  if (rows <= 0 || n <= 0) return 0;
  const int64_t stride = (rows + 31) / 32 * 32;
  return (size_t)stride * (size_t)n * sizeof(tfc::SortItem);
}

extern "C" int tfc_pmf_to_quantized_cdf(const float* d_pmf, int64_t rows, int32_t n,
                                        int precision, int32_t* d_cdf, void* d_workspace,
                                        size_t workspace_bytes, void* stream) {
  if (precision < 1 || precision > 16)
    return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "`precision` must be in [1, 16]: %d", precision);
  if (n < 2)
    return tfc::fail(TFC_ERR_INVALID_ARGUMENT,
                     "`pmf` size should be at least 2 in the last axis.");
  if (rows <= 0) return TFC_OK;
  if (workspace_bytes < tfc_pmf_to_quantized_cdf_workspace(rows, n) || d_workspace == nullptr)
    return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "workspace too small: need %zu bytes",
                     tfc_pmf_to_quantized_cdf_workspace(rows, n));
  int rc = ensure_lut();
  if (rc) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  // long rows: a warp per row in shared memory, as many warps per CTA as fit (at most 4)
  const size_t row_bytes = (size_t)((n + 3) & ~3) * kWarpRowBytesPerEntry;
  // (also for short rows when there are few of them: a thread per row would leave the GPU empty
  // and walk its row through global memory — the EntropyBottleneck table of a training step is
  // 128 rows of 22 entries)
  if ((n >= 64 || rows <= 2048) && row_bytes <= 200 * 1024) {
    const int wpc = (int)std::max<size_t>(1, std::min<size_t>(4, (200 * 1024) / row_bytes));
    const size_t smem = wpc * row_bytes;
    static tfc::PerDeviceOnce attr;
    if (attr.first()) {
      TFC_CUDA(cudaFuncSetAttribute(pmf_to_cdf_warp_kernel,
                                    cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    }
    pmf_to_cdf_warp_kernel<<<(unsigned)((rows + wpc - 1) / wpc), wpc * 32, smem, st>>>(
        d_pmf, rows, n, precision, d_cdf);
    TFC_LAUNCHED("pmf_to_cdf_kernel");
    return TFC_OK;
  }
  const int64_t stride = (rows + 31) / 32 * 32;
  const int threads = 32;
  pmf_to_cdf_kernel<<<(unsigned)((rows + threads - 1) / threads), threads, 0, st>>>(
      d_pmf, rows, n, precision, d_cdf, static_cast<tfc::SortItem*>(d_workspace), stride);
  TFC_LAUNCHED("pmf_to_cdf_kernel");
  return TFC_OK;
}


// =========================================================================================
// tfc_pmf_to_quantized_cdf_static — the same op for callers that cannot allocate
// =========================================================================================
// No counterpart in the reference: added for callers that
// cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
// forbid it).  It runs the thread-per-row kernel above (pmf_to_cdf_kernel, the one kernel of this
// op that needs scratch memory: the sort items of every row) with its scratch in a FILE-SCOPE
// device array instead of a caller-provided workspace.
//
//   * The array is kStaticScratchBytes = 8 MiB and every GPU that loads this library holds it
//     until the process exits.  It holds round_up(rows, 32) * n sort items of 16 bytes:
//     round_up(rows, 32) * n <= 524 288 (e.g. 8192 rows x 64 entries, 4096 x 128, 352 x 1480).
//     Larger problems are rejected with TFC_ERR_INVALID_ARGUMENT; use tfc_pmf_to_quantized_cdf.
//   * In host code `g_static_scratch` is only a host-side stand-in for the symbol, NOT a device
//     address: the address the kernel gets comes from cudaGetSymbolAddress, looked up once per
//     device with that device current (static_scratch_ptr below).
//   * Nothing here calls cudaMalloc / cudaFree or an allocating library routine.  The first call
//     on a device does the one-time set-up (symbol address; the log2 table upload of ensure_lut,
//     a cudaMemcpyToSymbol into another file-scope array).
//   * Every call shares the one array and finds the previous call's contents in it.  Nothing has
//     to start at zero: the kernel writes item i of a row (rebalance: acc.set for all i < n)
//     before anything reads it, and never reads past n.  Calls that overlap on the GPU (two
//     streams, two host threads) RACE on the array: serialise them, one stream at a time.
namespace {

constexpr size_t kStaticScratchBytes = size_t(8) << 20;
constexpr int64_t kStaticScratchItems = (int64_t)(kStaticScratchBytes / sizeof(tfc::SortItem));
static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];

// device address of g_static_scratch on the CURRENT device (cached per device)
int static_scratch_ptr(tfc::SortItem** out) {
  static std::mutex mu;
  static void* ptrs[64] = {nullptr};
  int dev = 0;
  TFC_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64)
    return tfc::fail(TFC_ERR_CUDA, "static scratch: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(mu);
  if (ptrs[dev] == nullptr) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
    if (e != cudaSuccess) return tfc::cuda_fail(e, "cudaGetSymbolAddress(g_static_scratch)");
    if (p == nullptr || (reinterpret_cast<uintptr_t>(p) & 255) != 0)
      return tfc::fail(TFC_ERR_CUDA, "static scratch: bad device address %p", p);
    ptrs[dev] = p;
  }
  *out = static_cast<tfc::SortItem*>(ptrs[dev]);
  return TFC_OK;
}

}  // namespace

extern "C" size_t tfc_pmf_to_quantized_cdf_static_capacity(void) { return kStaticScratchBytes; }

extern "C" int tfc_pmf_to_quantized_cdf_static(const float* d_pmf, int64_t rows, int32_t n,
                                               int precision, int32_t* d_cdf, void* stream) {
  if (precision < 1 || precision > 16)
    return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "`precision` must be in [1, 16]: %d", precision);
  if (n < 2)
    return tfc::fail(TFC_ERR_INVALID_ARGUMENT,
                     "`pmf` size should be at least 2 in the last axis.");
  if (rows <= 0) return TFC_OK;
  if (d_pmf == nullptr || d_cdf == nullptr)
    return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "pmf_to_quantized_cdf_static: null buffer");
  const int64_t stride = (rows + 31) / 32 * 32;
  if (stride > kStaticScratchItems || stride * (int64_t)n > kStaticScratchItems)
    return tfc::fail(TFC_ERR_INVALID_ARGUMENT,
                     "pmf_to_quantized_cdf_static: %lld rows x %d entries need %lld scratch items, "
                     "the static scratch array holds %lld (%zu bytes); use "
                     "tfc_pmf_to_quantized_cdf with a caller-provided workspace",
                     (long long)rows, n, (long long)(stride * (int64_t)n),
                     (long long)kStaticScratchItems, kStaticScratchBytes);
  int rc = ensure_lut();
  if (rc) return rc;
  tfc::SortItem* scratch = nullptr;
  rc = static_scratch_ptr(&scratch);
  if (rc) return rc;
  const int threads = 32;
  pmf_to_cdf_kernel<<<(unsigned)((rows + threads - 1) / threads), threads, 0,
                      static_cast<cudaStream_t>(stream)>>>(d_pmf, rows, n, precision, d_cdf,
                                                           scratch, stride);
  TFC_LAUNCHED("pmf_to_cdf_kernel");
  return TFC_OK;
}
