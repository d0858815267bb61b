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
// computed and the result is scanned by the whole warp; std::sort is replayed by the warp
// together (warp_sort below: same permutation as libstdc++'s, ties included); the re-insertion
// of the adjusted head — `find_if` + `rotate` in the reference (:165-181, :182-199) — is a
// 32-way search and a cooperative shift of the shorter side of a ring.
struct SmemAcc {
  tfc::SortItem* base;
  __device__ __forceinline__ tfc::SortItem get(int i) const { return base[i]; }
  __device__ __forceinline__ void set(int i, const tfc::SortItem& v) const { base[i] = v; }
};

constexpr int kWarpRowBytesPerEntry = (int)sizeof(tfc::SortItem) + 8;   // item + pmf + count
// (the two uint16 position lists of the sort live in the count array: the counts are in the items
// from the moment the keys are built until they are written back)
constexpr int kCoopMin = 64;    // ranges above this many items are partitioned by the whole warp

// The unguarded Hoare partition of std::sort on [first + 1, last) around the pivot at `first`,
// by the whole warp: StdSortReplay::partition_ranked (stdsort_replay.cuh; checked against the
// real std::sort on the host) with its candidate lists built by ballots.  Returns the cut.
template <bool DESC>
__device__ int partition_warp(tfc::SortItem* it, uint16_t* lpos, uint16_t* rasc, int first,
                              int last, int lane) {
  const double pk = it[first].key;
  const unsigned below = (1u << lane) - 1u;
  int nl = 0, nr = 0;
  for (int base = first; base < last; base += 32) {
    const int i = base + lane;
    bool fl = false, fr = false;
    if (i < last) {
      const double k = it[i].key;
      fl = i > first && !(DESC ? k > pk : k < pk);
      fr = !(DESC ? pk > k : pk < k);
    }
    const unsigned ml = __ballot_sync(0xFFFFFFFFu, fl), mr = __ballot_sync(0xFFFFFFFFu, fr);
    if (fl) lpos[nl + __popc(ml & below)] = (uint16_t)i;
    if (fr) rasc[nr + __popc(mr & below)] = (uint16_t)i;
    nl += __popc(ml);
    nr += __popc(mr);
  }
  __syncwarp();
  // K = number of leading pairs with L[k] < R[k] (monotone in k)
  const int pairs = min(nl, nr);
  int K = 0;
  for (int base = 0; base < pairs; base += 32) {
    const int k = base + lane;
    const bool ok = k < pairs && (int)lpos[k] < (int)rasc[nr - 1 - k];
    const unsigned m = __ballot_sync(0xFFFFFFFFu, ok);
    K += __popc(m);
    if (m != 0xFFFFFFFFu) break;
  }
  for (int k = lane; k < K; k += 32) {
    const int i = lpos[k], j = rasc[nr - 1 - k];
    const tfc::SortItem t = it[i];
    it[i] = it[j];
    it[j] = t;
  }
  int cut = 0x7FFFFFFF;
  if (K < nl) cut = lpos[K];
  if (K > 0) cut = min(cut, (int)rasc[nr - K]);
  __syncwarp();
  return cut;
}

// std::sort of it[0, n) with the tie order of libstdc++ (StdSortReplay), by one warp: the big
// partitions are done by all lanes together, ranges of at most kCoopMin items are handed out one
// per lane and finished serially (introsort loop + leaf insertion sorts) 32 at a time.  The
// decomposition is StdSortReplay::sort_decomposed, which tests/cpp/stdsort_check.cc compares with
// the real std::sort (ties, adversarial inputs that hit the heap-sort fallback).
template <bool DESC>
__device__ void warp_sort(tfc::SortItem* it, uint16_t* lpos, uint16_t* rasc, int n, int lane) {
  tfc::StdSortReplay<SmemAcc, DESC> ser(SmemAcc{it});
  if (n <= kCoopMin) {
    if (lane == 0) ser.sort(n);
    __syncwarp();
    return;
  }
  int lg = 0;
  for (int m = n; m > 1; m >>= 1) ++lg;
  int sf[48], sl[48], sd[48];
  int sp = 1;
  sf[0] = 0; sl[0] = n; sd[0] = 2 * lg;
  int myf = 0, myl = 0, myd = 0, pending = 0;   // this lane's small range; ranges handed out
  auto flush = [&]() {
    __syncwarp();
    if (lane < pending) ser.sort_range(myf, myl, myd);
    __syncwarp();
    pending = 0;
  };
  while (sp > 0) {
    --sp;
    const int first = sf[sp], last = sl[sp];
    int depth = sd[sp];
    if (last - first <= kCoopMin) {
      if (last - first >= 2) {
        if (lane == pending) { myf = first; myl = last; myd = depth; }
        if (++pending == 32) flush();
      }
      continue;
    }
    if (depth == 0) {
      __syncwarp();
      if (lane == 0) ser.heap_sort(first, last);
      __syncwarp();
      continue;
    }
    --depth;
    if (lane == 0) ser.median_to_first(first, first + 1, first + (last - first) / 2, last - 1);
    __syncwarp();
    const int cut = partition_warp<DESC>(it, lpos, rasc, first, last, lane);
    sf[sp] = cut; sl[sp] = last; sd[sp] = depth; ++sp;
    sf[sp] = first; sl[sp] = cut; sd[sp] = depth; ++sp;
  }
  flush();
}

// The queue of the reference (`std::sort`, then per adjustment `find_if` + `rotate`,
// range_coding_helper_kernels.cc:165-199) as a RING over the sorted items: logical position j is
// physical (start + j) mod n.  The adjusted head is re-inserted before the first item it
// compares less than; the items are sorted, so that position comes from a 32-way search instead
// of a walk, and the rotate moves whichever side of the ring is shorter — an item whose count
// drops to 1 gets an infinite key and goes to the back, which in the ring is its own slot.
template <bool OVER>
__device__ void rebalance_warp(tfc::SortItem* it, uint16_t* lpos, uint16_t* rasc, const float* sp,
                               int32_t* cnt, int n, int todo, int lane) {
  for (int i = lane; i < n; i += 32) {
    tfc::SortItem v;
    v.idx = i; v.pad = cnt[i];
    v.key = OVER ? steal_cost((double)sp[i], cnt[i]) : give_gain((double)sp[i], cnt[i]);
    it[i] = v;
  }
  __syncwarp();
  warp_sort<!OVER>(it, lpos, rasc, n, lane);
  int start = 0;
  auto phys = [&](int j) { const int q = start + j; return q >= n ? q - n : q; };
  for (; todo > 0; --todo) {
    tfc::SortItem head = it[start];
    head.pad += OVER ? -1 : 1;
    head.key = OVER ? steal_cost((double)sp[head.idx], head.pad)
                    : give_gain((double)sp[head.idx], head.pad);
    // pos = first logical j in [1, n) whose item stops the head, n if none
    // (the first round looks at the next 32 items one by one: an adjusted head usually moves a
    // few places; only then does the search widen to 32 evenly spaced probes)
    int lo = 1, hi = n;
    bool near = true;
    while (lo < hi) {
      const int step = near ? 1 : (hi - lo + 31) / 32;
      near = false;
      const int j = lo + lane * step;
      bool stop = true;
      if (j < hi) {
        const double k = it[phys(j)].key;
        stop = OVER ? head.key < k : head.key > k;
      }
      const unsigned m = __ballot_sync(0xFFFFFFFFu, stop);
      const int f = m ? __ffs(m) - 1 : 32;
      if (f == 0) { hi = lo; break; }
      const int nlo = lo + (f - 1) * step + 1;
      if (f < 32) hi = min(hi, lo + f * step);     // f == 32: every probe was before the stop
      lo = nlo;
    }
    const int pos = lo;
    __syncwarp();
    if (pos - 1 <= n - pos) {
      // logical [1, pos) one place towards the front, the head lands at pos - 1
      for (int base = 1; base < pos; base += 32) {
        const int j = base + lane;
        tfc::SortItem o;
        if (j < pos) o = it[phys(j)];
        __syncwarp();
        if (j < pos) it[phys(j - 1)] = o;
        __syncwarp();
      }
      if (lane == 0) it[phys(pos - 1)] = head;
    } else {
      // drop the head's slot from the front (start + 1: every logical index goes down by one)
      // and move the new logical [pos - 1, n - 1) one place towards the back, into that slot
      start = start + 1 == n ? 0 : start + 1;
      for (int top = n - 1; top > pos - 1; top -= 32) {
        const int j = top - 1 - lane;                // source, going down from n - 2
        tfc::SortItem o;
        if (j >= pos - 1) o = it[phys(j)];
        __syncwarp();
        if (j >= pos - 1) it[phys(j + 1)] = o;
        __syncwarp();
      }
      if (lane == 0) it[phys(pos - 1)] = head;
    }
    __syncwarp();
  }
  for (int i = lane; i < n; i += 32) cnt[it[i].idx] = it[i].pad;
  __syncwarp();
}

// `lens` (ragged form): row r has lens[r] entries (clamped to [2, nmax]) at pmf + r * pmf_stride
// and its cdf goes to cdf + r * cdf_stride, zero-filled behind its lens[r] + 1 entries; without
// `lens` every row has nmax entries and the strides are nmax / nmax + 1.
__global__ void pmf_to_cdf_warp_kernel(const float* __restrict__ pmf, int64_t rows, int nmax,
                                       int precision, int32_t* __restrict__ cdf,
                                       const int32_t* __restrict__ lens, int pmf_stride,
                                       int cdf_stride) {
  extern __shared__ __align__(16) uint8_t smem_rows[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, wpc = blockDim.x >> 5;
  const int64_t r = (int64_t)blockIdx.x * wpc + warp;
  if (r >= rows) return;
  const int n = lens ? min(max(__ldg(lens + r), 2), nmax) : nmax;
  const int npad = (nmax + 3) & ~3;
  uint8_t* mine = smem_rows + (size_t)warp * npad * kWarpRowBytesPerEntry;
  tfc::SortItem* it = reinterpret_cast<tfc::SortItem*>(mine);
  float* sp = reinterpret_cast<float*>(it + npad);
  int32_t* cnt = reinterpret_cast<int32_t*>(sp + npad);
  uint16_t* lpos = reinterpret_cast<uint16_t*>(cnt);
  uint16_t* rasc = lpos + npad;
  const float* p = pmf + r * (int64_t)pmf_stride;
  int32_t* c = cdf + r * (int64_t)cdf_stride;
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
  if (sum > total) rebalance_warp<true>(it, lpos, rasc, sp, cnt, n, sum - total, lane);
  else if (sum < total) rebalance_warp<false>(it, lpos, rasc, sp, cnt, n, total - sum, lane);
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
  for (int i = n + 1 + lane; i < cdf_stride; i += 32) c[i] = 0;
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
    // warps per CTA: the re-insertion loop is one dependent chain per row, so what matters is
    // how many rows an SM holds at once: the CTA size (1, 2 or 4 rows) that packs most rows
    // into the SM's 227 KB (1 KB reserved per CTA)
    int wpc = 1;
    size_t best = 0;
    for (int w : {4, 2, 1}) {
      if (w * row_bytes > 200 * 1024) continue;
      const size_t per_sm = std::min<size_t>(32 / w, (227 * 1024) / (w * row_bytes + 1024)) * w;
      if (per_sm > best) { best = per_sm; wpc = w; }
    }
    const size_t smem = wpc * row_bytes;
    static tfc::PerDeviceOnce attr;
    if (attr.first()) {
      TFC_CUDA(cudaFuncSetAttribute(pmf_to_cdf_warp_kernel,
                                    cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    }
    pmf_to_cdf_warp_kernel<<<(unsigned)((rows + wpc - 1) / wpc), wpc * 32, smem, st>>>(
        d_pmf, rows, n, precision, d_cdf, nullptr, n, n + 1);
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


// Rows of different lengths in one launch (the EntropyBottleneck table rebuild of a training
// step, entropy_models.py:131-146: row r is pmf[r, :len_r] ++ [tail_r], zero-padded to the longest
// row): the lengths stay on the device, so the caller neither reads them back nor groups rows.
extern "C" int tfc_pmf_to_quantized_cdf_ragged(const float* d_pmf, int64_t rows,
                                               int32_t pmf_stride, const int32_t* d_n,
                                               int32_t max_n, int precision, int32_t* d_cdf,
                                               int32_t cdf_stride, void* stream) {
  if (precision < 1 || precision > 16)
    return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "`precision` must be in [1, 16]: %d", precision);
  if (max_n < 2 || pmf_stride < max_n || cdf_stride < max_n + 1 || d_n == nullptr)
    return tfc::fail(TFC_ERR_INVALID_ARGUMENT,
                     "ragged pmf: need max_n >= 2, pmf_stride >= max_n, cdf_stride >= max_n + 1");
  if (rows <= 0) return TFC_OK;
  const size_t row_bytes = (size_t)((max_n + 3) & ~3) * kWarpRowBytesPerEntry;
  if (row_bytes > 200 * 1024)
    return tfc::fail(TFC_ERR_UNIMPLEMENTED, "ragged pmf rows of up to %d entries", max_n);
  int rc = ensure_lut();
  if (rc) return rc;
  int wpc = 1;
  size_t best = 0;
  for (int w : {4, 2, 1}) {
    if (w * row_bytes > 200 * 1024) continue;
    const size_t per_sm = std::min<size_t>(32 / w, (227 * 1024) / (w * row_bytes + 1024)) * w;
    if (per_sm > best) { best = per_sm; wpc = w; }
  }
  static tfc::PerDeviceOnce attr;
  if (attr.first()) {
    TFC_CUDA(cudaFuncSetAttribute(pmf_to_cdf_warp_kernel,
                                  cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  }
  pmf_to_cdf_warp_kernel<<<(unsigned)((rows + wpc - 1) / wpc), wpc * 32, wpc * row_bytes,
                           static_cast<cudaStream_t>(stream)>>>(
      d_pmf, rows, max_n, precision, d_cdf, d_n, pmf_stride, cdf_stride);
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
