// PmfToQuantizedCdf on the device: one thread per row, bit-exact with
// cc/kernels/range_coding_helper_kernels.cc:97-201 as built with this image's libstdc++/glibc.
//
// What makes the reference's result implementation-defined, and how each piece is pinned:
//   std::rint(float)        -> rintf (same IEEE round-half-even)
//   std::log2(int) (double) -> lookup table filled ON THE HOST with glibc's log2 at first use
//   penalty/gain            -> IEEE double subtract + multiply (exact on both sides)
//   std::sort tie order     -> StdSortReplay (stdsort_replay.cuh)
//   find_if + rotate        -> the same linear scan + shift
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
  const int64_t stride = (rows + 31) / 32 * 32;
  const int threads = 32;
  pmf_to_cdf_kernel<<<(unsigned)((rows + threads - 1) / threads), threads, 0, st>>>(
      d_pmf, rows, n, precision, d_cdf, static_cast<tfc::SortItem*>(d_workspace), stride);
  TFC_LAUNCHED("pmf_to_cdf_kernel");
  return TFC_OK;
}
