// Batched range coder for sm_100a: one warp per independent stream.
//
// Replaces (citations relative to /root/reference/tensorflow_compression/cc/kernels/):
//   RangeEncoder::Encode/Finalize, RangeDecoder::Decode      range_coder.cc:38-377
//   UnboundedIndexRangeEncodeOp / ...DecodeOp                unbounded_index_range_coding_kernels.cc:141-376
//   RangeEncodeOp / RangeDecodeOp, BroadcastRange            range_coding_kernels.cc:63-380
//   MergeAxes                                                range_coding_kernels_util.cc:36-92
//
// Design.  The coder state is a serial dependency chain per stream, so
// the unit of parallelism is the stream (one per image, as EntropyModel.compress emits them).
//   encode: the 32 lanes of a warp fetch symbols+indexes coalesced and do every
//           state-INDEPENDENT step in parallel (offset shift, escape mapping, the two CDF
//           lookups), packing (lower, upper-1) into one 32-bit word in shared memory; lane 0
//           then runs the 32-bit state machine over the staged words.
//   decode: the search depends on the state, so all lanes carry the state and each probes one
//           CDF entry of a 32-wide window placed around the row's mode (tables are staged
//           ragged in shared memory); one ballot finds the symbol, two shuffles fetch its
//           bounds.  Rows longer than the window fall back to a 32-ary ballot search.
// All arithmetic is integer and bit-exact with the reference.
#include <algorithm>
#include <cstdlib>
#include <vector>

#include "common.cuh"

namespace {

constexpr int kChunk = 128;          // symbols staged per encode round (4 per lane)
constexpr int kMaxMerged = 6;        // range_coding_kernels.cc:210-230
constexpr int kTableBudget = 96 * 1024;

struct Broadcast {  // merged data shape + cdf row strides (in rows) for the bounded ops
  int32_t nd;
  int64_t dshape[kMaxMerged];
  int64_t rstride[kMaxMerged];
};

// ---------------------------------------------------------------------------------------
// encoder state machine (range_coder.cc:38-301)
// ---------------------------------------------------------------------------------------
struct Encoder {
  uint32_t lo, span1;
  uint64_t held;  // 0: interval inside [0,2^32); else low 16 bits = pending word (+1),
                  // upper bits = 0x20000 per pending pair of filler bytes
  uint8_t* out;
  uint32_t pos, cap;

  __device__ __forceinline__ void init(uint8_t* o, uint32_t c) {
    lo = 0; span1 = 0xFFFFFFFFu; held = 0; out = o; pos = 0; cap = c;
  }
  __device__ __forceinline__ void byte(uint32_t b) {
    if (pos < cap) out[pos] = (uint8_t)b;
    ++pos;
  }
  __device__ __forceinline__ void word(uint32_t w) { byte(w >> 8); byte(w & 0xFFu); }

  __device__ __forceinline__ void put(uint32_t c0, uint32_t c1, int prec) {
    // span*c == span1*c + c with span = span1+1 <= 2^32: one 32x32+64 multiply-add each
    const uint32_t a = (uint32_t)(((uint64_t)span1 * c0 + c0) >> prec);
    const uint32_t b = (uint32_t)(((uint64_t)span1 * c1 + c1) >> prec) - 1u;
    lo += a;
    const bool wrapped = lo < a;
    span1 = b - a;
    if (lo + span1 < lo) {  // interval straddles 2^32: carry still undecided
      if ((span1 >> 16) == 0) {
        lo <<= 16;
        span1 = (span1 << 16) | 0xFFFFu;
        held += 0x20000;
      }
      return;
    }
    if (held != 0) {
      if (!wrapped) --held;
      word((uint32_t)held & 0xFFFFu);
      const uint32_t fill = wrapped ? 0x00u : 0xFFu;
      for (uint64_t k = held >> 16; k != 0; --k) byte(fill);
      held = 0;
    }
    if ((span1 >> 16) == 0) {
      const uint32_t hi = lo >> 16;
      lo <<= 16;
      span1 = (span1 << 16) | 0xFFFFu;
      if (lo + span1 >= lo) word(hi); else held = hi + 1;
    }
  }

  __device__ __forceinline__ void flush() {
    uint32_t w;
    if (held != 0) w = (uint32_t)held & 0xFFFFu;
    else if (lo != 0) w = ((lo - 1) >> 16) + 1;
    else return;
    byte(w >> 8);
    if (w & 0xFFu) byte(w & 0xFFu);
  }

  // escape tail of the unbounded-index code (unbounded_index_range_coding_kernels.cc:232-247)
  __device__ __forceinline__ void escape(uint32_t x, int w) {
    const uint32_t all1 = (1u << w) - 1u;
    int nd = 0;
    while (nd * w < 32 && (x >> (nd * w)) != 0) ++nd;
    uint32_t u = (uint32_t)nd;
    for (; u >= all1; u -= all1) put(all1, all1 + 1, w);
    put(u, u + 1, w);
    for (int j = 0; j < nd; ++j) {
      const uint32_t d = (x >> (j * w)) & all1;
      put(d, d + 1, w);
    }
  }
};

struct EncodeParams {
  // unbounded-index form
  const int32_t* data32;
  const int32_t* index;
  int64_t index_stride;
  const int32_t* cdf_size;
  const int32_t* offset;
  int32_t rows, cols;
  int overflow_width;
  // bounded form
  const int16_t* data16;
  int64_t cdf_stream_stride;
  int32_t chip;
  Broadcast bc;
  // common
  const int32_t* cdf;
  int64_t S, n;
  int precision;
  uint8_t* out;
  int64_t cap;
  int32_t* nbytes;
  int32_t* status;
};

__device__ __forceinline__ int64_t broadcast_row(const Broadcast& bc, int64_t i) {
  int64_t row = 0;
#pragma unroll
  for (int d = kMaxMerged - 1; d >= 0; --d) {
    if (d < bc.nd) {
      const int64_t q = i / bc.dshape[d];
      row += (i - q * bc.dshape[d]) * bc.rstride[d];
      i = q;
    }
  }
  return row;
}

template <bool UNBOUNDED>
__global__ void __launch_bounds__(256) encode_kernel(const EncodeParams p) {
  extern __shared__ __align__(16) uint32_t smem_enc[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t s = (int64_t)blockIdx.x * (blockDim.x >> 5) + warp;
  if (s >= p.S) return;
  uint32_t* packed = smem_enc + warp * (2 * kChunk + 4);
  uint32_t* extra = packed + kChunk;
  uint32_t* escm = extra + kChunk;

  Encoder e;
  e.init(p.out + s * p.cap, (uint32_t)p.cap);
  const int32_t* idx = UNBOUNDED ? p.index + s * p.index_stride : nullptr;
  const int32_t* d32 = UNBOUNDED ? p.data32 + s * p.n : nullptr;
  const int16_t* d16 = UNBOUNDED ? nullptr : p.data16 + s * p.n;
  const int32_t* cdf = UNBOUNDED ? p.cdf : p.cdf + s * p.cdf_stream_stride;

  for (int64_t base = 0; base < p.n; base += kChunk) {
    const int cnt = (int)min((int64_t)kChunk, p.n - base);
#pragma unroll
    for (int u = 0; u < kChunk / 32; ++u) {
      const int j = u * 32 + lane;
      bool esc = false;
      if (j < cnt) {
        uint32_t c0, c1, x = 0;
        if (UNBOUNDED) {
          const int32_t r = __ldg(idx + base + j);
          const int32_t top = __ldg(p.cdf_size + r) - 2;
          int32_t v = __ldg(d32 + base + j) - __ldg(p.offset + r);
          if (v < 0) { x = (uint32_t)(-2 * v - 1); v = top; esc = true; }
          else if (v >= top) { x = (uint32_t)(2 * (v - top)); v = top; esc = true; }
          const int32_t* row = cdf + (int64_t)r * p.cols;
          c0 = (uint32_t)__ldg(row + v);
          c1 = (uint32_t)__ldg(row + v + 1);
        } else {
          const int64_t rowi = broadcast_row(p.bc, base + j);
          const int32_t* row = cdf + rowi * p.chip;
          const int32_t v = d16[base + j];
          c0 = (uint32_t)__ldg(row + v);
          c1 = (uint32_t)__ldg(row + v + 1);
        }
        packed[j] = (c0 & 0xFFFFu) | ((c1 - 1u) << 16);
        extra[j] = x;
      }
      const uint32_t m = __ballot_sync(0xFFFFFFFFu, esc);
      if (lane == 0) escm[u] = m;
    }
    __syncwarp();
    if (lane == 0) {
      for (int g = 0; g < cnt; g += 4) {
        const uint4 q = *reinterpret_cast<const uint4*>(packed + g);
        const uint32_t m = UNBOUNDED ? (escm[g >> 5] >> (g & 31)) : 0u;
        const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
        for (int t = 0; t < 4; ++t) {
          if (g + t < cnt) {
            e.put(w[t] & 0xFFFFu, (w[t] >> 16) + 1u, p.precision);
            if (UNBOUNDED && ((m >> t) & 1u)) e.escape(extra[g + t], p.overflow_width);
          }
        }
      }
    }
    __syncwarp();
  }
  if (lane == 0) {
    e.flush();
    p.nbytes[s] = (int32_t)e.pos;
    if (p.status) p.status[s] = e.pos > e.cap ? TFC_STREAM_CAPACITY : TFC_STREAM_OK;
  }
}

// ---------------------------------------------------------------------------------------
// decoder (range_coder.cc:303-377)
// ---------------------------------------------------------------------------------------
struct DecodeParams {
  const uint8_t* enc;
  int64_t cap;
  const int32_t* nbytes;
  // unbounded-index form
  const int32_t* index;
  int64_t index_stride;
  const int32_t* cdf_size;
  const int32_t* offset;
  int32_t rows, cols;
  int overflow_width;
  int32_t* out32;
  // bounded form
  int64_t cdf_stream_stride;
  int32_t chip;
  Broadcast bc;
  int16_t* out16;
  // common
  const int32_t* cdf;
  int64_t S, n;
  int precision;
  int32_t* status;
  int32_t table_budget;  // bytes of dynamic shared memory available for the staged tables
};

struct Decoder {
  uint32_t lo, span1, code, nxt;
  const uint8_t* src;
  uint32_t pos, len;
  bool corrupt;

  __device__ __forceinline__ uint32_t peek16(uint32_t at) const {
    uint32_t v = 0;
    if (at < len) v = (uint32_t)__ldg(src + at) << 8;
    if (at + 1 < len) v |= (uint32_t)__ldg(src + at + 1);
    return v;
  }
  __device__ __forceinline__ void init(const uint8_t* s, uint32_t n) {
    src = s; len = n; lo = 0; span1 = 0xFFFFFFFFu; corrupt = false;
    code = (peek16(0) << 16) | peek16(2);
    pos = 4;
    nxt = peek16(pos);
  }
  __device__ __forceinline__ uint64_t target(int prec) const {
    return (((uint64_t)(uint32_t)(code - lo) + 1u) << prec) - 1u;
  }
  // span * c > target ?
  __device__ __forceinline__ bool above(uint32_t c, uint64_t tgt) const {
    return (uint64_t)span1 * c + c > tgt;
  }
  __device__ __forceinline__ void narrow(uint32_t c0, uint32_t c1, int prec) {
    const uint32_t a = (uint32_t)(((uint64_t)span1 * c0 + c0) >> prec);
    const uint32_t b = (uint32_t)(((uint64_t)span1 * c1 + c1) >> prec) - 1u;
    lo += a;
    span1 = b - a;
    if ((span1 >> 16) == 0) {
      lo <<= 16;
      span1 = (span1 << 16) | 0xFFFFu;
      code = (code << 16) | nxt;
      pos += 2;
      nxt = peek16(pos);
    }
  }
  // one digit of the uniform 2^w alphabet (cdf = iota): smallest k with span*k > target is
  // floor(target/span)+1
  __device__ __forceinline__ uint32_t digit(int w) {
    const uint64_t span = (uint64_t)span1 + 1u;
    const uint64_t d = target(w) / span;
    if (d >> w) { corrupt = true; }
    const uint32_t dd = (uint32_t)min(d, (uint64_t)((1u << w) - 1u));
    narrow(dd, dd + 1u, w);
    return dd;
  }
};

// 32-ary ballot search for the smallest k in [lo, hi] with span*row[k] > target (row[lo-1]
// is known not to satisfy it).  Returns hi+1 if none does.
template <typename RowPtr>
__device__ __forceinline__ int wide_search(const Decoder& d, RowPtr row, int lo, int hi,
                                           uint64_t tgt, int lane) {
  while (hi - lo + 1 > 32) {
    const int step = (hi - lo + 1 + 31) >> 5;
    const int k = min(lo + (lane + 1) * step - 1, hi);
    const uint32_t b = __ballot_sync(0xFFFFFFFFu, d.above((uint32_t)row[k], tgt));
    if (b == 0) return hi + 1;
    const int f = __ffs(b) - 1;
    const int nlo = lo + f * step;
    hi = min(nlo + step - 1, hi);
    lo = nlo;
  }
  const int k = lo + lane;
  const uint32_t b =
      __ballot_sync(0xFFFFFFFFu, k <= hi && d.above((uint32_t)row[min(k, hi)], tgt));
  if (b == 0) return hi + 1;
  return lo + __ffs(b) - 1;
}

// Shared-memory image of the tables: ragged rows + one int4 {start, len, window base, offset}
// per row.  Falls back to the dense global table when it does not fit.
// bounded decode: every element has its own (broadcast) row of `chip` entries
__global__ void __launch_bounds__(256) decode_bounded_kernel(const DecodeParams p) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t s = (int64_t)blockIdx.x * (blockDim.x >> 5) + warp;
  if (s >= p.S) return;
  Decoder d;
  d.init(p.enc + s * p.cap, tfc::clamp_stream_len(p.nbytes[s], p.cap));
  d.corrupt = (int64_t)p.nbytes[s] > p.cap;   // a truncated string
  const int32_t* cdf = p.cdf + s * p.cdf_stream_stride;
  int16_t* out = p.out16 + s * p.n;
  const int len = p.chip, prec = p.precision;
  for (int64_t base = 0; base < p.n; base += 32) {
    const int cnt = (int)min((int64_t)32, p.n - base);
    const int64_t myrow = lane < cnt ? broadcast_row(p.bc, base + lane) : 0;
    int32_t mine = 0;
    for (int j = 0; j < cnt; ++j) {
      const int32_t* row = cdf + __shfl_sync(0xFFFFFFFFu, myrow, j) * len;
      const uint64_t tgt = d.target(prec);
      int k = wide_search(d, row, 1, len - 1, tgt, lane);
      if (k > len - 1) { d.corrupt = true; k = len - 1; }
      d.narrow((uint32_t)__ldg(row + k - 1), (uint32_t)__ldg(row + k), prec);
      if (lane == j) mine = k - 1;
    }
    if (lane < cnt) out[base + lane] = (int16_t)mine;
  }
  if (lane == 0 && p.status) p.status[s] = d.corrupt ? TFC_STREAM_CORRUPT : TFC_STREAM_OK;
}

}  // namespace
#include "rc_unbounded.cuh"
#include "rc_v4.cuh"
#include "rc_lanes.cuh"
namespace {

// ---------------------------------------------------------------------------------------
// debug_level=1 argument checks (unbounded_index_range_coding_kernels.cc:58-115,
// range_coding_kernels.cc:153-176, :250-255), run on the device; the first failure wins.
// ---------------------------------------------------------------------------------------
enum CheckCode { kCheckOk = 0, kBadIndex, kBadCdfSize, kBadCdfEnds, kNotMonotonic,
                 kBadCdfEndsBounded, kBadData };
struct CheckRecord { int32_t code; int32_t a; int32_t b; int32_t pad; };

__device__ __forceinline__ void report(CheckRecord* rec, int code, int a, int b) {
  if (atomicCAS(&rec->code, kCheckOk, code) == kCheckOk) { rec->a = a; rec->b = b; }
}

__global__ void check_index_kernel(const int32_t* index, int64_t count, int32_t rows,
                                   CheckRecord* rec) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < count;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int32_t v = index[i];
    if (v < 0 || v >= rows) report(rec, kBadIndex, v, 0);
  }
}

// ragged == true: rows have individual lengths cdf_size[r] (unbounded ops); else all `cols`
__global__ void check_cdf_kernel(const int32_t* cdf, int64_t rows, int32_t cols,
                                 const int32_t* cdf_size, bool ragged, int precision,
                                 CheckRecord* rec) {
  const int32_t top = 1 << precision;
  for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < rows;
       r += (int64_t)gridDim.x * blockDim.x) {
    int32_t len = cols;
    if (ragged) {
      len = cdf_size[r];
      if (len < 3 || len > cols) { report(rec, kBadCdfSize, len, 0); continue; }
    }
    const int32_t* row = cdf + r * cols;
    if (row[0] != 0 || row[len - 1] != top) {
      report(rec, ragged ? kBadCdfEnds : kBadCdfEndsBounded, row[0], row[len - 1]);
      continue;
    }
    for (int j = 0; j + 1 < len; ++j)
      if (row[j + 1] <= row[j]) { report(rec, kNotMonotonic, 0, 0); break; }
  }
}

__global__ void check_data16_kernel(const int16_t* data, int64_t count, int32_t chip,
                                    CheckRecord* rec) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < count;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int32_t v = data[i];
    if (v < 0 || chip <= v + 1) report(rec, kBadData, v, 0);
  }
}

int finish_checks(CheckRecord* d_rec, cudaStream_t st, int rows, int cols, int precision,
                  int chip) {
  CheckRecord h;
  TFC_CUDA(cudaMemcpyAsync(&h, d_rec, sizeof(h), cudaMemcpyDeviceToHost, st));
  TFC_CUDA(cudaStreamSynchronize(st));
  cudaFreeAsync(d_rec, st);
  switch (h.code) {
    case kCheckOk: return TFC_OK;
    case kBadIndex:
      return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "'index' has a value not in [0, %d): value=%d",
                       rows, h.a);
    case kBadCdfSize:
      return tfc::fail(TFC_ERR_INVALID_ARGUMENT,
                       "'cdf_size' has a value not in [3, %d]: value=%d", cols, h.a);
    case kBadCdfEnds:
      return tfc::fail(TFC_ERR_INVALID_ARGUMENT,
                       "Each cdf should start from 0 and end at %d: cdf[0]=%d, cdf[^1]=%d",
                       1 << precision, h.a, h.b);
    case kBadCdfEndsBounded:
      return tfc::fail(TFC_ERR_INVALID_ARGUMENT,
                       "CDF should start from 0 and end at %d: cdf[0]=%d, cdf[^1]=%d",
                       1 << precision, h.a, h.b);
    case kNotMonotonic:
      return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "CDF is not monotonic");
    case kBadData:
      return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "'data' value not in [0, %d): value=%d",
                       chip - 1, h.a);
  }
  return TFC_OK;
}

int alloc_record(CheckRecord** rec, cudaStream_t st) {
  TFC_CUDA(cudaMallocAsync(reinterpret_cast<void**>(rec), sizeof(CheckRecord), st));
  TFC_CUDA(cudaMemsetAsync(*rec, 0, sizeof(CheckRecord), st));
  return TFC_OK;
}

int check_attrs(int precision, int overflow_width, int debug_level, bool has_overflow) {
  if (precision < 1 || precision > 16)
    return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "`precision` must be in [1, 16]: %d", precision);
  if (has_overflow && (overflow_width < 1 || overflow_width > 16))
    return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "`overflow_width` must be in [1, 16]: %d",
                     overflow_width);
  if (debug_level != 0 && debug_level != 1)
    return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "`debug_level` must be 0 or 1: %d", debug_level);
  return TFC_OK;
}

int check_unbounded_shapes(int32_t rows, int32_t cols) {
  // CheckArgumentShapes (unbounded_index_range_coding_kernels.cc:117-139): the 1-D/2-D parts
  // are implied by the C signature; what is left is the column count.
  if (rows < 0 || cols < 3)
    return tfc::fail(TFC_ERR_INVALID_ARGUMENT,
                     "'cdf' should be 2-D and cdf.dim_size(1) >= 3: [%d,%d]", rows, cols);
  return TFC_OK;
}

int run_unbounded_checks(const int32_t* d_index, int64_t index_count, const int32_t* d_cdf,
                         int32_t rows, int32_t cols, const int32_t* d_cdf_size, int precision,
                         cudaStream_t st) {
  CheckRecord* rec = nullptr;
  int rc = alloc_record(&rec, st);
  if (rc) return rc;
  if (index_count > 0) {
    const int blocks = (int)std::min<int64_t>((index_count + 255) / 256, 4 * tfc::num_sms());
    check_index_kernel<<<blocks, 256, 0, st>>>(d_index, index_count, rows, rec);
    TFC_LAUNCHED("check_index_kernel");
  }
  if (rows > 0) {
    check_cdf_kernel<<<(rows + 127) / 128, 128, 0, st>>>(d_cdf, rows, cols, d_cdf_size, true,
                                                         precision, rec);
    TFC_LAUNCHED("check_cdf_kernel");
  }
  return finish_checks(rec, st, rows, cols, precision, 0);
}


// ---- generation-4 unbounded kernels (rc_v4.cuh): launch plumbing -----------------------------
constexpr int kV4TableBudget = 72 * 1024;

// dynamic shared memory a warps-per-stream coder CTA may take (tfc_configure)
size_t coder_smem_max() {
  const int kb = tfc::g_coder_smem_kb.load(std::memory_order_relaxed);
  return kb > 0 ? (size_t)kb * 1024 : (size_t)(227 * 1024 - 1024);
}
constexpr int kCoderSmemAttr = 227 * 1024 - 1024;   // opt-in limit set once per device

int v4_max_puts(int overflow_width) {   // puts (= stream words at most) one symbol can take
  const int ndmax = (32 + overflow_width - 1) / overflow_width;
  const int all1 = (1 << overflow_width) - 1;
  return 1 + ndmax / all1 + 2 + ndmax;
}

bool v4_usable(int32_t rows) {
  static int disabled = -1;
  if (disabled < 0) {
    const char* e = getenv("TFC_CODER_V3");
    disabled = (e && e[0] == '1') ? 1 : 0;
  }
  return !disabled && rows >= 1 && rows <= 1024;
}

// Shared memory reserved for the staged tables: the dense size, at most the budget
// (tfc_configure(TFC_CFG_CODER_TABLE_KB), 72 KB by default), never less than the row records and
// row infos (32 bytes per row), which the kernels always stage.  Ragged rows that do not fit are
// read from the caller's dense table through L1 (stage_tables_v4 decides on the device, where
// the row lengths are).
int v4_table_budget(int32_t rows, int32_t cols) {
  const size_t dense = v4::tables_bytes(rows, (int64_t)rows * cols);
  const int kb = tfc::g_coder_table_kb.load(std::memory_order_relaxed);
  const size_t budget = std::max<size_t>(kb > 0 ? (size_t)kb * 1024 : (size_t)kV4TableBudget,
                                         (size_t)rows * 32);
  return (int)((std::min<size_t>(dense, budget) + 15) & ~size_t(15));
}

// ---- one stream per lane (rc_lanes.cuh): thousands of streams ------------------------------
bool lanes_wanted(int64_t S, int32_t rows) {
  const int mode = tfc::g_coder_lanes.load(std::memory_order_relaxed);
  if (!v4_usable(rows) || mode == 1) return false;
  return mode == 2 || S >= 2048;
}

// warps per CTA: spread the ceil(S / 32) warps over the SMs, at most 16 per CTA (they share one
// staged table); returns the dynamic shared memory of the launch
size_t lanes_geometry(int64_t S, int table_bytes, int* wpc, int64_t* blocks) {
  const int64_t warps = (S + 31) / 32;
  const int sms = tfc::num_sms();
  int w = (int)std::max<int64_t>(1, std::min<int64_t>(lanes::kMaxWarps, (warps + sms - 1) / sms));
  // packed (tfc_configure TFC_CFG_CODER_STREAMS_PER_CTA >= 32): pack / 32 warps share one CTA, so
  // that a latency-bound chain warp (about 0.2 instructions per cycle on its own) has 15
  // neighbours filling the SM's issue slots instead of holding an SM by itself
  const int pack = tfc::g_coder_pack.load(std::memory_order_relaxed);
  if (pack >= 32) w = (int)std::min<int64_t>(std::min(lanes::kMaxWarps, pack / 32), warps);
  const size_t smem_max = 227 * 1024 - 1024;
  while (w > 1 && (size_t)table_bytes + w * lanes::kWarpBytes > smem_max) --w;
  *wpc = w;
  *blocks = (warps + w - 1) / w;
  return (size_t)table_bytes + w * lanes::kWarpBytes;
}

bool bounded_lanes_wanted(int64_t S) {
  const int mode = tfc::g_coder_lanes.load(std::memory_order_relaxed);
  // measured (m = 64, 4096 symbols): warp per stream 14.0 / 5.8 Gsym/s at 4 K streams and
  // 18.5 / 7.5 at 16 K; lane per stream 6.7 / 5.0 and 25.5 / 18.0 (87.9 / 64.4 at 64 K)
  return mode == 2 || (mode == 0 && S >= 8192);
}

// a table shared by all streams is staged in shared memory when it fits in 64 KB
int bounded_table_bytes(int64_t cdf_rows, int chip, int64_t cdf_stream_stride) {
  const int64_t bytes = cdf_rows * chip * 4;
  if (cdf_stream_stride != 0 || bytes > 64 * 1024) return 0;
  return (int)((bytes + 15) & ~int64_t(15));
}

int warps_per_block(int64_t S) {
  // few streams: one warp per block so the streams spread over the SMs; many: pack 8
  const int sms = tfc::num_sms();
  if (S <= 4 * (int64_t)sms) return 1;
  if (S <= 16 * (int64_t)sms) return 4;
  return 8;
}

std::string shape_str(const int64_t* s, int n) {
  std::string r = "[";
  for (int i = 0; i < n; ++i) { if (i) r += ","; r += std::to_string(s[i]); }
  return r + "]";
}

// CheckCdfShape + MergeAxes, producing the device-side broadcast descriptor.
int plan_broadcast(const int64_t* data_shape, int rank, const int64_t* cdf_shape, int cdf_rank,
                   Broadcast* bc, int64_t* count, int32_t* chip, int64_t* cdf_rows) {
  if (cdf_rank != rank + 1)
    return tfc::fail(TFC_ERR_INVALID_ARGUMENT,
                     "`cdf` should have one more axis than `data`: data shape=%s, cdf shape=%s",
                     shape_str(data_shape, rank).c_str(), shape_str(cdf_shape, cdf_rank).c_str());
  if (cdf_shape[cdf_rank - 1] <= 1)
    return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "The last dimension of `cdf` should be > 1: %s",
                     shape_str(cdf_shape, cdf_rank).c_str());
  std::vector<int64_t> md(rank + 1), mc(rank + 2);
  int mr = 0;
  int rc = tfc_merge_axes(data_shape, rank, cdf_shape, cdf_rank, md.data(), mc.data(), &mr);
  if (rc) return rc;
  if (mr > kMaxMerged)
    return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "Irregular broadcast pattern: %s, %s",
                     shape_str(data_shape, rank).c_str(), shape_str(cdf_shape, cdf_rank).c_str());
  bc->nd = mr;
  int64_t n = 1, stride = 1;
  for (int i = mr - 1; i >= 0; --i) {
    bc->dshape[i] = md[i];
    bc->rstride[i] = mc[i] <= 1 ? 0 : stride;
    stride *= mc[i];
    n *= md[i];
  }
  for (int i = mr; i < kMaxMerged; ++i) { bc->dshape[i] = 1; bc->rstride[i] = 0; }
  *count = n;
  *chip = (int32_t)cdf_shape[cdf_rank - 1];
  *cdf_rows = stride;
  return TFC_OK;
}

}  // namespace

// =========================================================================================
// C ABI
// =========================================================================================
extern "C" int tfc_merge_axes(const int64_t* data_shape, int rank, const int64_t* cdf_shape,
                              int cdf_rank, int64_t* merged_data_shape,
                              int64_t* merged_cdf_shape, int* merged_rank) {
  if (cdf_rank != rank + 1)
    return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "`cdf` should have one more axis than `data`");
  // Adjacent axes fuse when both broadcast or both do not; size-1 axes fuse with anything.
  int m = 0;
  merged_data_shape[0] = 1;
  merged_cdf_shape[0] = 1;
  for (int j = 0; j < rank; ++j) {
    const int64_t dj = data_shape[j], cj = cdf_shape[j];
    if (dj != cj && cj != 1)
      return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "Cannot broadcast shape %s to %s",
                       shape_str(cdf_shape, cdf_rank).c_str(),
                       shape_str(data_shape, rank).c_str());
    const bool prev_bcast = merged_cdf_shape[m] == 1;
    const bool this_bcast = cj == 1;
    if (prev_bcast == this_bcast || dj <= 1 || merged_data_shape[m] <= 1) {
      merged_data_shape[m] *= dj;
      merged_cdf_shape[m] *= cj;
    } else {
      ++m;
      merged_data_shape[m] = dj;
      merged_cdf_shape[m] = cj;
    }
  }
  merged_cdf_shape[m + 1] = cdf_shape[cdf_rank - 1];
  *merged_rank = m + 1;
  return TFC_OK;
}

extern "C" int tfc_unbounded_index_range_encode(
    const int32_t* d_data, const int32_t* d_index, int64_t index_stride, int64_t S, int64_t n,
    const int32_t* d_cdf, int32_t rows, int32_t cols, const int32_t* d_cdf_size,
    const int32_t* d_offset, int precision, int overflow_width, int debug_level, uint8_t* d_out,
    int64_t cap, int32_t* d_nbytes, int32_t* d_status, void* stream) {
  int rc = check_attrs(precision, overflow_width, debug_level, true);
  if (rc) return rc;
  rc = check_unbounded_shapes(rows, cols);
  if (rc) return rc;
  if (index_stride != 0 && index_stride != n)
    return tfc::fail(TFC_ERR_INVALID_ARGUMENT,
                     "`data` and `index` should have the same shape (index_stride must be n or 0)");
  if (cap < 0 || cap > 0x7FFFFFF0ll || n < 0 || S < 0)
    return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "bad sizes: S=%lld n=%lld cap=%lld", (long long)S,
                     (long long)n, (long long)cap);
  if (S == 0) return TFC_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (debug_level > 0) {
    rc = run_unbounded_checks(d_index, index_stride ? S * n : n, d_cdf, rows, cols, d_cdf_size,
                              precision, st);
    if (rc) return rc;
  }
  if (lanes_wanted(S, rows)) {
    lanes::Params q{};
    q.data = d_data; q.index = d_index; q.index_stride = index_stride; q.cdf = d_cdf;
    q.cdf_size = d_cdf_size; q.offset = d_offset; q.rows = rows; q.cols = cols;
    q.precision = precision; q.overflow_width = overflow_width; q.out = d_out; q.cap = cap;
    q.nbytes = d_nbytes; q.status = d_status; q.S = S; q.n = n;
    q.table_bytes = v4_table_budget(rows, cols);
    int wpc = 1;
    int64_t blocks = 0;
    const size_t smem = lanes_geometry(S, q.table_bytes, &wpc, &blocks);
    static tfc::PerDeviceOnce attr;
    if (attr.first()) {
      TFC_CUDA(cudaFuncSetAttribute(lanes::encode_unbounded_lanes,
                                    cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - 1024));
    }
    lanes::encode_unbounded_lanes<<<(unsigned)blocks, wpc * 32, smem, st>>>(q);
    TFC_LAUNCHED("encode_unbounded_lanes");
    return TFC_OK;
  }
  if (v4_usable(rows)) {
    v4::EncParamsV4 q{};
    q.data = d_data; q.index = d_index; q.index_stride = index_stride; q.cdf = d_cdf;
    q.cdf_size = d_cdf_size; q.offset = d_offset; q.rows = rows; q.cols = cols;
    q.precision = precision; q.overflow_width = overflow_width; q.out = d_out; q.cap = cap;
    q.nbytes = d_nbytes; q.status = d_status; q.S = S; q.n = n;
    q.table_bytes = v4_table_budget(rows, cols);
    // streams per CTA: 1 unless the caller packs the coder (tfc_configure); the per-stream
    // round buffers shrink until the CTA fits in shared memory
    const int pack = tfc::g_coder_pack.load(std::memory_order_relaxed);
    q.streams_per_cta = (int)std::max<int64_t>(1, std::min<int64_t>(pack > 0 ? std::min(pack, 8) : 1, S));
    const int min_cap = (32 * v4_max_puts(overflow_width) + 31) / 32 * 32;   // one 32-symbol round
    q.put_cap = std::max(v4::kPutCap, min_cap);
    q.max_puts = v4_max_puts(overflow_width);
    const size_t smem_max = coder_smem_max();
    while (q.put_cap / 2 >= std::max(256, min_cap) &&
           (size_t)q.table_bytes + q.streams_per_cta * v4::enc_bytes(q.put_cap) > smem_max)
      q.put_cap /= 2;
    while (q.streams_per_cta > 1 &&
           (size_t)q.table_bytes + q.streams_per_cta * v4::enc_bytes(q.put_cap) > smem_max)
      --q.streams_per_cta;
    q.round = std::max(32, std::min(v4::kRoundMax, q.put_cap / v4_max_puts(overflow_width) / 32 * 32));
    if (q.round * v4_max_puts(overflow_width) > q.put_cap)
      return tfc::fail(TFC_ERR_UNIMPLEMENTED, "overflow_width %d needs larger round buffers", overflow_width);
    q.fixed_round = (tfc::debug_bits() & 128) ? q.round : 0;   // A/B: the worst-case-sized rounds
    const size_t smem = (size_t)q.table_bytes + q.streams_per_cta * v4::enc_bytes(q.put_cap);
    static tfc::PerDeviceOnce attr;
    if (attr.first()) {
      TFC_CUDA(cudaFuncSetAttribute(v4::encode_unbounded_v4<false>,
                                    cudaFuncAttributeMaxDynamicSharedMemorySize, kCoderSmemAttr));
      TFC_CUDA(cudaFuncSetAttribute(v4::encode_unbounded_v4<true>,
                                    cudaFuncAttributeMaxDynamicSharedMemorySize, kCoderSmemAttr));
    }
    const int64_t blocks = (S + q.streams_per_cta - 1) / q.streams_per_cta;
    if (q.streams_per_cta == 1)
      v4::encode_unbounded_v4<false><<<(unsigned)blocks, 96, smem, st>>>(q);
    else
      v4::encode_unbounded_v4<true><<<(unsigned)blocks, 96 * q.streams_per_cta, smem, st>>>(q);
    TFC_LAUNCHED("encode_unbounded_v4");
    return TFC_OK;
  }
  EncodeParams p{};
  p.data32 = d_data; p.index = d_index; p.index_stride = index_stride;
  p.cdf_size = d_cdf_size; p.offset = d_offset; p.cols = cols; p.overflow_width = overflow_width;
  p.cdf = d_cdf; p.S = S; p.n = n; p.precision = precision;
  p.out = d_out; p.cap = cap; p.nbytes = d_nbytes; p.status = d_status;
  p.rows = rows;
  const int64_t dense = (int64_t)rows * cols * 4 + (int64_t)rows * 16 + 16;
  const int budget = (int)std::min<int64_t>(dense, kTableBudget);
  const size_t smem = (size_t)4 * kSuper * sizeof(uint32_t) + (size_t)budget;
  static tfc::PerDeviceOnce enc_attr_set;
  if (enc_attr_set.first()) {
    TFC_CUDA(cudaFuncSetAttribute(encode_unbounded_kernel,
                                  cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)(4 * kSuper * sizeof(uint32_t)) + kTableBudget));
  }
  encode_unbounded_kernel<<<(unsigned)S, 128, smem, st>>>(p, budget);
  TFC_LAUNCHED("encode_kernel<unbounded>");
  return TFC_OK;
}

extern "C" int tfc_unbounded_index_range_decode(
    const uint8_t* d_enc, int64_t cap, const int32_t* d_nbytes, const int32_t* d_index,
    int64_t index_stride, int64_t S, int64_t n, const int32_t* d_cdf, int32_t rows, int32_t cols,
    const int32_t* d_cdf_size, const int32_t* d_offset, int precision, int overflow_width,
    int debug_level, int32_t* d_out, int32_t* d_status, void* stream) {
  int rc = check_attrs(precision, overflow_width, debug_level, true);
  if (rc) return rc;
  rc = check_unbounded_shapes(rows, cols);
  if (rc) return rc;
  if (index_stride != 0 && index_stride != n)
    return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "index_stride must be n or 0");
  if (cap < 0 || cap > 0x7FFFFFF0ll || n < 0 || S < 0)
    return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "bad sizes: S=%lld n=%lld cap=%lld", (long long)S,
                     (long long)n, (long long)cap);
  if (S == 0 || n == 0) return TFC_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (debug_level > 0) {
    rc = run_unbounded_checks(d_index, index_stride ? S * n : n, d_cdf, rows, cols, d_cdf_size,
                              precision, st);
    if (rc) return rc;
  }
  if (lanes_wanted(S, rows)) {
    lanes::Params q{};
    q.enc = d_enc; q.cap = cap; q.nbytes = const_cast<int32_t*>(d_nbytes); q.index = d_index;
    q.index_stride = index_stride; q.cdf = d_cdf; q.cdf_size = d_cdf_size; q.offset = d_offset;
    q.rows = rows; q.cols = cols; q.precision = precision; q.overflow_width = overflow_width;
    q.values = d_out; q.status = d_status; q.S = S; q.n = n;
    q.max_digits = (32 + overflow_width - 1) / overflow_width;
    q.table_bytes = v4_table_budget(rows, cols);
    int wpc = 1;
    int64_t blocks = 0;
    const size_t smem = lanes_geometry(S, q.table_bytes, &wpc, &blocks);
    static tfc::PerDeviceOnce attr;
    if (attr.first()) {
      TFC_CUDA(cudaFuncSetAttribute(lanes::decode_unbounded_lanes,
                                    cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - 1024));
    }
    lanes::decode_unbounded_lanes<<<(unsigned)blocks, wpc * 32, smem, st>>>(q);
    TFC_LAUNCHED("decode_unbounded_lanes");
    return TFC_OK;
  }
  if (v4_usable(rows)) {
    v4::DecParamsV4 q{};
    q.enc = d_enc; q.cap = cap; q.nbytes = d_nbytes; q.index = d_index;
    q.index_stride = index_stride; q.cdf = d_cdf; q.cdf_size = d_cdf_size; q.offset = d_offset;
    q.rows = rows; q.cols = cols; q.precision = precision; q.overflow_width = overflow_width;
    q.out = d_out; q.status = d_status; q.S = S; q.n = n;
    const int mp = v4_max_puts(overflow_width);
    q.max_digits = (32 + overflow_width - 1) / overflow_width;
    q.table_bytes = v4_table_budget(rows, cols);
    const int sms = tfc::num_sms();
    q.streams_per_cta = S <= sms ? 1 : (S <= 2 * (int64_t)sms ? 2 : 4);
    const int pack = tfc::g_coder_pack.load(std::memory_order_relaxed);
    if (pack > 0) q.streams_per_cta = (int)std::max<int64_t>(1, std::min<int64_t>(std::min(pack, 8), S));
    // rounds as long as shared memory allows (fewer barriers), shorter when streams are packed
    const size_t smem_max = coder_smem_max();
    auto geometry = [&](int round) {
      q.round = round;
      q.round_words = (round * mp + 3) / 4 * 4;
      q.lin_words = (2 * q.round_words + 16 + 7) / 8 * 8;
      return (size_t)q.table_bytes +
             (size_t)q.streams_per_cta * v4::dec_pair_bytes(q.round, q.lin_words);
    };
    int round = std::max(32, std::min(v4::kRoundMax, (v4::kLinMax - 16) / 2 / mp / 32 * 32));
    while (round > 32 && geometry(round) > smem_max) round -= 32;
    while (q.streams_per_cta > 1 && geometry(round) > smem_max) --q.streams_per_cta;
    const size_t smem = geometry(round);
    if (smem > smem_max)
      return tfc::fail(TFC_ERR_UNIMPLEMENTED, "overflow_width %d needs a larger stream window", overflow_width);
    static tfc::PerDeviceOnce attr;
    if (attr.first()) {
      TFC_CUDA(cudaFuncSetAttribute(v4::decode_unbounded_v4,
                                    cudaFuncAttributeMaxDynamicSharedMemorySize, kCoderSmemAttr));
    }
    const int64_t blocks = (S + q.streams_per_cta - 1) / q.streams_per_cta;
    v4::decode_unbounded_v4<<<(unsigned)blocks, 64 * q.streams_per_cta, smem, st>>>(q);
    TFC_LAUNCHED("decode_unbounded_v4");
    return TFC_OK;
  }
  DecodeParams p{};
  p.enc = d_enc; p.cap = cap; p.nbytes = d_nbytes; p.index = d_index;
  p.index_stride = index_stride; p.cdf_size = d_cdf_size; p.offset = d_offset;
  p.rows = rows; p.cols = cols; p.overflow_width = overflow_width; p.out32 = d_out;
  p.cdf = d_cdf; p.S = S; p.n = n; p.precision = precision; p.status = d_status;
  const int64_t dense = (int64_t)rows * cols * 4 + (int64_t)rows * 16 + 16;
  const int budget = (int)std::min<int64_t>(dense, kTableBudget);
  p.table_budget = budget;
  static tfc::PerDeviceOnce attr_set;
  if (attr_set.first()) {
    TFC_CUDA(cudaFuncSetAttribute(decode_unbounded_kernel2,
                                  cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  kTableBudget + 16 + 8 * kDecScratch));
  }
  const int wpb = warps_per_block(S);
  const int64_t blocks = (S + wpb - 1) / wpb;
  decode_unbounded_kernel2<<<(unsigned)blocks, wpb * 32,
                             (size_t)((budget + 15) & ~15) + (size_t)wpb * kDecScratch, st>>>(p);
  TFC_LAUNCHED("decode_unbounded_kernel");
  return TFC_OK;
}

extern "C" int tfc_range_encode(const int16_t* d_data, int64_t S, const int64_t* data_shape,
                                int rank, const int32_t* d_cdf, const int64_t* cdf_shape,
                                int cdf_rank, int64_t cdf_stream_stride, int precision,
                                int debug_level, uint8_t* d_out, int64_t cap, int32_t* d_nbytes,
                                int32_t* d_status, void* stream) {
  int rc = check_attrs(precision, 1, debug_level, false);
  if (rc) return rc;
  EncodeParams p{};
  int64_t n = 0, cdf_rows = 0;
  rc = plan_broadcast(data_shape, rank, cdf_shape, cdf_rank, &p.bc, &n, &p.chip, &cdf_rows);
  if (rc) return rc;
  if (cap < 0 || cap > 0x7FFFFFF0ll || S < 0)
    return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "bad sizes: S=%lld cap=%lld", (long long)S,
                     (long long)cap);
  if (S == 0) return TFC_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (debug_level > 0) {
    if (p.chip <= 2)
      return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "CDF size should be > 2: %d", p.chip);
    CheckRecord* rec = nullptr;
    rc = alloc_record(&rec, st);
    if (rc) return rc;
    const int64_t total_rows = cdf_stream_stride ? cdf_rows + (S - 1) * (cdf_stream_stride / p.chip)
                                                 : cdf_rows;
    check_cdf_kernel<<<(unsigned)((total_rows + 127) / 128), 128, 0, st>>>(
        d_cdf, total_rows, p.chip, nullptr, false, precision, rec);
    TFC_LAUNCHED("check_cdf_kernel");
    if (S * n > 0) {
      const int blocks = (int)std::min<int64_t>((S * n + 255) / 256, 4 * tfc::num_sms());
      check_data16_kernel<<<blocks, 256, 0, st>>>(d_data, S * n, p.chip, rec);
      TFC_LAUNCHED("check_data16_kernel");
    }
    rc = finish_checks(rec, st, 0, 0, precision, p.chip);
    if (rc) return rc;
  }
  if (bounded_lanes_wanted(S)) {
    lanes::BoundedParams q{};
    q.data = d_data; q.cdf = d_cdf; q.cdf_stream_stride = cdf_stream_stride; q.cdf_rows = cdf_rows;
    q.chip = p.chip; q.precision = precision; q.bc = p.bc; q.out = d_out; q.cap = cap;
    q.nbytes = d_nbytes; q.status = d_status; q.S = S; q.n = n;
    q.table_bytes = bounded_table_bytes(cdf_rows, p.chip, cdf_stream_stride);
    int wpc = 1;
    int64_t blocks = 0;
    const size_t smem = lanes_geometry(S, q.table_bytes, &wpc, &blocks);
    static tfc::PerDeviceOnce attr;
    if (attr.first()) {
      TFC_CUDA(cudaFuncSetAttribute(lanes::encode_bounded_lanes,
                                    cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - 1024));
    }
    lanes::encode_bounded_lanes<<<(unsigned)blocks, wpc * 32, smem, st>>>(q);
    TFC_LAUNCHED("encode_bounded_lanes");
    return TFC_OK;
  }
  p.data16 = d_data; p.cdf_stream_stride = cdf_stream_stride; p.cdf = d_cdf;
  p.S = S; p.n = n; p.precision = precision;
  p.out = d_out; p.cap = cap; p.nbytes = d_nbytes; p.status = d_status;
  const int wpb = warps_per_block(S);
  const size_t smem = (size_t)wpb * (2 * kChunk + 4) * sizeof(uint32_t);
  encode_kernel<false><<<(unsigned)((S + wpb - 1) / wpb), wpb * 32, smem, st>>>(p);
  TFC_LAUNCHED("encode_kernel<bounded>");
  return TFC_OK;
}

extern "C" int tfc_range_decode(const uint8_t* d_enc, int64_t cap, const int32_t* d_nbytes,
                                int64_t S, const int64_t* data_shape, int rank,
                                const int32_t* d_cdf, const int64_t* cdf_shape, int cdf_rank,
                                int64_t cdf_stream_stride, int precision, int debug_level,
                                int16_t* d_out, int32_t* d_status, void* stream) {
  int rc = check_attrs(precision, 1, debug_level, false);
  if (rc) return rc;
  DecodeParams p{};
  int64_t n = 0, cdf_rows = 0;
  rc = plan_broadcast(data_shape, rank, cdf_shape, cdf_rank, &p.bc, &n, &p.chip, &cdf_rows);
  if (rc) return rc;
  if (cap < 0 || cap > 0x7FFFFFF0ll || S < 0)
    return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "bad sizes: S=%lld cap=%lld", (long long)S,
                     (long long)cap);
  if (S == 0 || n == 0) return TFC_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (debug_level > 0) {
    if (p.chip <= 2)
      return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "CDF size should be > 2: %d", p.chip);
    CheckRecord* rec = nullptr;
    rc = alloc_record(&rec, st);
    if (rc) return rc;
    const int64_t total_rows = cdf_stream_stride ? cdf_rows + (S - 1) * (cdf_stream_stride / p.chip)
                                                 : cdf_rows;
    check_cdf_kernel<<<(unsigned)((total_rows + 127) / 128), 128, 0, st>>>(
        d_cdf, total_rows, p.chip, nullptr, false, precision, rec);
    TFC_LAUNCHED("check_cdf_kernel");
    rc = finish_checks(rec, st, 0, 0, precision, p.chip);
    if (rc) return rc;
  }
  if (bounded_lanes_wanted(S)) {
    lanes::BoundedParams q{};
    q.cdf = d_cdf; q.cdf_stream_stride = cdf_stream_stride; q.cdf_rows = cdf_rows;
    q.chip = p.chip; q.precision = precision; q.bc = p.bc; q.enc = d_enc; q.cap = cap;
    q.nbytes = const_cast<int32_t*>(d_nbytes); q.values = d_out; q.status = d_status;
    q.S = S; q.n = n;
    q.table_bytes = bounded_table_bytes(cdf_rows, p.chip, cdf_stream_stride);
    int wpc = 1;
    int64_t blocks = 0;
    const size_t smem = lanes_geometry(S, q.table_bytes, &wpc, &blocks);
    static tfc::PerDeviceOnce attr;
    if (attr.first()) {
      TFC_CUDA(cudaFuncSetAttribute(lanes::decode_bounded_lanes,
                                    cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - 1024));
    }
    lanes::decode_bounded_lanes<<<(unsigned)blocks, wpc * 32, smem, st>>>(q);
    TFC_LAUNCHED("decode_bounded_lanes");
    return TFC_OK;
  }
  p.enc = d_enc; p.cap = cap; p.nbytes = d_nbytes; p.cdf_stream_stride = cdf_stream_stride;
  p.out16 = d_out; p.cdf = d_cdf; p.S = S; p.n = n; p.precision = precision; p.status = d_status;
  const int wpb = warps_per_block(S);
  decode_bounded_kernel<<<(unsigned)((S + wpb - 1) / wpb), wpb * 32, 0, st>>>(p);
  TFC_LAUNCHED("decode_bounded_kernel");
  return TFC_OK;
}

// Debug: the occupancy trace of the warps-per-stream coder kernels (rc_v4.cuh g_trace):
// out4 = {encode ns, encode streams, decode ns, decode streams} summed since the last reset.
// Synchronises the device.
extern "C" int tfc_debug_coder_trace(uint64_t* h_out4, int reset) {
  TFC_CUDA(cudaDeviceSynchronize());
  unsigned long long v[4] = {0, 0, 0, 0};
  TFC_CUDA(cudaMemcpyFromSymbol(v, v4::g_trace, sizeof(v)));
  if (h_out4) for (int i = 0; i < 4; ++i) h_out4[i] = v[i];
  if (reset) {
    unsigned long long z[4] = {0, 0, 0, 0};
    TFC_CUDA(cudaMemcpyToSymbol(v4::g_trace, z, sizeof(z)));
  }
  return TFC_OK;
}
