// Unbounded-index range coder for MANY streams (included by range_coder.cu): one stream per
// LANE.  The v4 kernels (rc_v4.cuh) give each stream two or three warps and shorten the serial
// chain of one image-sized stream; with thousands of independent streams (BASELINE configs[3]:
// 1K-64K streams of 4096 symbols) the chains themselves are the parallelism, so here every
// thread runs the plain scalar recurrence of RangeEncoder / RangeDecoder
// (range_coder.cc:38-301, :303-377) and the op loops of
// unbounded_index_range_coding_kernels.cc:189-251, :308-368 for its own stream:
//   * tables staged in shared memory pre-shifted to precision 16 (stage_tables_v4), shared by
//     all the warps of the CTA;
//   * symbols / indexes / decoded values move through a per-warp 32 x 32 shared-memory tile so
//     that global memory sees 128-byte rows (the tensors are stream-major [S, n]: a lane
//     reading its own stream directly would touch 32 different lines per load);
//   * encoded bytes are stored by each lane into its own slab row (2-byte stores, merged in
//     L2); the decoder reads its string through an 8-byte big-endian register window.
// Results are bit-identical to the v4 kernels and to the reference.
#pragma once

namespace lanes {

constexpr int kTile = 32;               // symbols per tile row
constexpr int kLd = 33;                 // padded row length (conflict-free lane-owns-row reads)
constexpr int kMaxWarps = 16;
constexpr size_t kWarpBytes = (size_t)2 * kTile * kLd * sizeof(int32_t);

struct Params {
  const int32_t* data;        // encode: symbols [S, n]
  const int32_t* index;
  int64_t index_stride;       // n, or 0 when the index is shared by all streams
  const int32_t* cdf;
  const int32_t* cdf_size;
  const int32_t* offset;
  int32_t rows, cols, precision, overflow_width;
  uint8_t* out;               // encode: slab [S, cap]
  const uint8_t* enc;         // decode: slab [S, cap]
  int64_t cap;
  int32_t* nbytes;            // encode: written; decode: read
  int32_t* values;            // decode: symbols [S, n]
  int32_t* status;
  int64_t S, n;
  int32_t table_bytes;
  int32_t max_digits;
};

// ---- encoder state of one stream (range_coder.cc:38-301) ------------------------------------
// [lo, lo + span1] is the interval; `held` is non-zero while it straddles 2^32: its low 16 bits
// are the word whose final value (itself or itself - 1) is not known yet, the bits above count
// the filler bytes that follow it (the reference's delay_).
struct Enc {
  uint32_t lo, span1, pos, cap;
  uint64_t held;
  uint8_t* out;
  bool even;

  __device__ __forceinline__ void byte(uint32_t v) {
    if (pos < cap) out[pos] = (uint8_t)v;
    ++pos;
  }
  __device__ __forceinline__ void word(uint32_t w) {
    if (even && (pos & 1u) == 0 && pos + 2u <= cap) {
      *reinterpret_cast<uint16_t*>(out + pos) = (uint16_t)(((w & 0xFFu) << 8) | ((w >> 8) & 0xFFu));
      pos += 2u;
    } else {
      byte(w >> 8);
      byte(w);
    }
  }
  // c0, c1: CDF entries scaled to precision 16
  __device__ __forceinline__ void put(uint32_t c0, uint32_t c1) {
    const uint64_t span = (uint64_t)span1 + 1u;
    const uint32_t a = (uint32_t)((span * c0) >> 16);
    const uint32_t b = (uint32_t)((span * c1) >> 16) - 1u;
    lo += a;
    const bool wrapped = lo < a;
    span1 = b - a;
    if ((uint32_t)(lo + span1) < lo) {            // straddles 2^32 (range_coder.cc:169-210)
      if ((span1 >> 16) == 0) {
        lo <<= 16;
        span1 = (span1 << 16) | 0xFFFFu;
        held += 0x20000u;
      }
      return;
    }
    if (held != 0) {                              // resolved (range_coder.cc:215-247)
      if (!wrapped) --held;
      word((uint32_t)held);
      const uint32_t fill = wrapped ? 0u : 0xFFu;
      for (uint64_t k = held >> 16; k > 0; --k) byte(fill);
      held = 0;
    }
    if ((span1 >> 16) == 0) {                     // range_coder.cc:249-265
      const uint32_t hi = lo >> 16;
      lo <<= 16;
      span1 = (span1 << 16) | 0xFFFFu;
      if ((uint32_t)(lo + span1) >= lo) word(hi); else held = (uint64_t)hi + 1u;
    }
  }
  __device__ __forceinline__ void flush() {       // range_coder.cc:268-301
    uint32_t w;
    if (held != 0) w = (uint32_t)held & 0xFFFFu;
    else if (lo != 0) w = ((lo - 1u) >> 16) + 1u;
    else return;
    byte(w >> 8);
    if (w & 0xFFu) byte(w);
  }
};

// A lane walks its own stream as a small state machine, ONE interval update per trip of the
// loop: phase 0 codes the next symbol (or its escape marker), phase 1 the digit count in unary
// groups, phase 2 the digits.  The lanes of the warp are therefore only synchronised at tile
// boundaries: a lane inside an escape (about 11 updates) does not hold the other 31 up, the
// way a per-symbol loop with the escape expansion in its body does (measured: 44 % of the
// 32-symbol steps of the Gaussian workload hold at least one escape).
// The next tile's symbols and indexes are fetched into registers before the chain runs.
__global__ void __launch_bounds__(kMaxWarps * 32, 1) encode_unbounded_lanes(const Params p) {
  v4::Tables t;
  v4::stage_tables_v4(p.cdf, p.rows, p.cols, p.cdf_size, p.offset, p.precision, v4::smem_v4,
                      p.table_bytes, &t);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, wpc = blockDim.x >> 5;
  const int64_t s0 = ((int64_t)blockIdx.x * wpc + warp) * 32;
  if (s0 >= p.S) return;
  int32_t* td = reinterpret_cast<int32_t*>(v4::smem_v4 + p.table_bytes + (size_t)warp * kWarpBytes);
  int32_t* ti = td + kTile * kLd;
  const int64_t s = s0 + lane;
  const bool active = s < p.S;
  const int rows_here = (int)min((int64_t)32, p.S - s0);
  Enc e;
  e.lo = 0; e.span1 = 0xFFFFFFFFu; e.pos = 0; e.cap = (uint32_t)p.cap; e.held = 0;
  e.out = p.out + (active ? s : s0) * p.cap;
  e.even = (reinterpret_cast<uintptr_t>(e.out) & 1) == 0;
  const int ow = p.overflow_width, dsh = 16 - ow;
  const uint32_t all1 = (1u << ow) - 1u;
  int32_t pd[kTile], pi[kTile];
  auto fetch = [&](int64_t i0) {
    const bool col = i0 + lane < p.n;
#pragma unroll
    for (int r = 0; r < kTile; ++r) {
      const bool on = col && r < rows_here;
      pd[r] = on ? __ldg(p.data + (s0 + r) * p.n + i0 + lane) : 0;
      pi[r] = on ? __ldg(p.index + (s0 + r) * p.index_stride + i0 + lane) : 0;
    }
  };
  fetch(0);
  for (int64_t i0 = 0; i0 < p.n; i0 += kTile) {
#pragma unroll
    for (int r = 0; r < kTile; ++r) {
      td[r * kLd + lane] = pd[r];
      ti[r * kLd + lane] = pi[r];
    }
    __syncwarp();
    if (i0 + kTile < p.n) fetch(i0 + kTile);
    if (active) {
      const int cnt = (int)min((int64_t)kTile, p.n - i0);
      int j = 0, phase = 0, nd = 0, q = 0;
      uint32_t x = 0, u = 0;
      while (j < cnt || phase != 0) {
        uint32_t c0, c1;
        if (phase == 0) {
          const int32_t row = min(max(ti[lane * kLd + j], 0), p.rows - 1);
          const int4 ri = t.rowinfo[row];                    // {start, len, offset, entry 0}
          const uint4 rr = t.rowrec[row];
          const int32_t esc = ri.y - 2;                      // escape symbol == #direct values
          const int32_t v = td[lane * kLd + j] - ri.z;
          const bool direct = v >= 0 && v < esc;
          ++j;
          if (direct && rr.z != 0u) {                        // 5-entry row: {e0, c1, c2, c3, top}
            c0 = v == 0 ? (uint32_t)ri.w : ((v == 1 ? rr.x : rr.y) >> 16);
            c1 = (v == 0 ? rr.x : (v == 1 ? rr.y : rr.z)) >> 16;
          } else {
            const int32_t k = direct ? v : esc;
            c0 = t.at(ri.x + k);
            c1 = t.at(ri.x + k + 1);
          }
          if (!direct) {       // unbounded_index_range_coding_kernels.cc:213-247
            x = v < 0 ? (uint32_t)(-2 * v - 1) : (uint32_t)(2 * (v - esc));
            nd = 0;
            while (((uint64_t)x >> (nd * ow)) != 0) ++nd;
            u = (uint32_t)nd;
            q = 0;
            phase = 1;
          }
        } else if (phase == 1) {                             // digit count, all1 per group
          if (u >= all1) {
            c0 = all1 << dsh; c1 = (all1 + 1u) << dsh;
            u -= all1;
          } else {
            c0 = u << dsh; c1 = (u + 1u) << dsh;
            phase = nd > 0 ? 2 : 0;
          }
        } else {                                             // digits, least significant first
          const uint32_t dg = (x >> (q * ow)) & all1;
          c0 = dg << dsh; c1 = (dg + 1u) << dsh;
          if (++q == nd) phase = 0;
        }
        e.put(c0, c1);
      }
    }
    __syncwarp();
  }
  if (active) {
    e.flush();
    p.nbytes[s] = (int32_t)e.pos;
    if (p.status) p.status[s] = e.pos > e.cap ? TFC_STREAM_CAPACITY : TFC_STREAM_OK;
  }
}

// ---- decoder state of one stream (range_coder.cc:303-377) -----------------------------------
struct Dec {
  uint32_t rem;       // code - lo
  uint32_t span1;
  uint32_t pos, len;
  uint64_t win;       // next stream bytes, big-endian, most significant first
  int avail;          // 16-bit words left in win
  const uint8_t* src;
  bool vec, corrupt;

  __device__ __forceinline__ void refill() {
    if (vec && pos + 8u <= len) {
      const uint2 v = __ldg(reinterpret_cast<const uint2*>(src + pos));
      win = ((uint64_t)__byte_perm(v.x, 0, 0x0123) << 32) | (uint64_t)__byte_perm(v.y, 0, 0x0123);
    } else {
      win = 0;                                   // zeros past the end (range_coder.cc:368-377)
      for (uint32_t k = 0; k < 8u; ++k)
        if (pos + k < len) win |= (uint64_t)__ldg(src + pos + k) << (56 - 8 * (int)k);
    }
    pos += 8u;
    avail = 4;
  }
  __device__ __forceinline__ uint32_t next16() {
    if (avail == 0) refill();
    const uint32_t w = (uint32_t)(win >> 48);
    win <<= 16;
    --avail;
    return w;
  }
  __device__ __forceinline__ void narrow(uint32_t a, uint32_t b) {
    rem -= a;
    span1 = b - a;
    if ((span1 >> 16) == 0) {
      span1 = (span1 << 16) | 0xFFFFu;
      rem = (rem << 16) | next16();
    }
  }
  // smallest k in [1, len-1] with floor(span * row[k] / 2^16) > rem; returns k - 1
  __device__ __forceinline__ int32_t get(const v4::Tables& t, int start, int len) {
    const uint64_t span = (uint64_t)span1 + 1u;
    int first = 1, count = len - 1;
    while (count > 0) {
      const int step = count >> 1;
      if (((span * t.at(start + first + step)) >> 16) <= (uint64_t)rem) {
        first += step + 1;
        count -= step + 1;
      } else {
        count = step;
      }
    }
    if (first > len - 1) { corrupt = true; first = len - 1; }   // ran past the row
    const uint32_t a = (uint32_t)((span * t.at(start + first - 1)) >> 16);
    const uint32_t b = (uint32_t)((span * t.at(start + first)) >> 16) - 1u;
    narrow(a, b);
    return first - 1;
  }
  // uniform alphabet of 2^ow digits: largest d with floor(span * d / 2^ow) <= rem
  __device__ __forceinline__ uint32_t digit(int dsh, uint32_t all1) {
    const uint64_t span = (uint64_t)span1 + 1u;
    uint32_t lo = 0, hi = all1;
    while (lo < hi) {
      const uint32_t mid = (lo + hi + 1u) >> 1;
      if (((span * ((uint64_t)mid << dsh)) >> 16) <= (uint64_t)rem) lo = mid; else hi = mid - 1u;
    }
    const uint32_t a = (uint32_t)((span * ((uint64_t)lo << dsh)) >> 16);
    const uint32_t b = (uint32_t)((span * ((uint64_t)(lo + 1u) << dsh)) >> 16) - 1u;
    narrow(a, b);
    return lo;
  }
};

// Same structure as the encoder: one interval search per trip of the loop (symbol, digit-count
// group or digit), lanes synchronised only at tile boundaries.  A symbol of a 5-entry row is
// found with three multiplies (the row record holds c1..c3 scaled so that
// floor(span * c / 2^16) is one IMAD.HI); everything else is the binary search of Dec::get.
__global__ void __launch_bounds__(kMaxWarps * 32, 1) decode_unbounded_lanes(const Params p) {
  v4::Tables t;
  v4::stage_tables_v4(p.cdf, p.rows, p.cols, p.cdf_size, p.offset, p.precision, v4::smem_v4,
                      p.table_bytes, &t);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, wpc = blockDim.x >> 5;
  const int64_t s0 = ((int64_t)blockIdx.x * wpc + warp) * 32;
  if (s0 >= p.S) return;
  int32_t* tv = reinterpret_cast<int32_t*>(v4::smem_v4 + p.table_bytes + (size_t)warp * kWarpBytes);
  int32_t* ti = tv + kTile * kLd;
  const int64_t s = s0 + lane;
  const bool active = s < p.S;
  const int rows_here = (int)min((int64_t)32, p.S - s0);
  Dec d;
  d.src = p.enc + (active ? s : s0) * p.cap;
  d.len = active ? (uint32_t)max(min((int64_t)p.nbytes[s], p.cap), (int64_t)0) : 0u;
  d.pos = 0; d.win = 0; d.avail = 0;
  d.corrupt = active && (int64_t)p.nbytes[s] > p.cap;   // a truncated string
  d.vec = (reinterpret_cast<uintptr_t>(d.src) & 7) == 0;
  d.span1 = 0xFFFFFFFFu;
  d.rem = 0;
  if (active) {
    const uint32_t hi = d.next16();
    d.rem = (hi << 16) | d.next16();
  }
  const int ow = p.overflow_width, dsh = 16 - ow;
  const uint32_t all1 = (1u << ow) - 1u;
  int32_t pi[kTile];
  auto fetch = [&](int64_t i0) {
    const bool col = i0 + lane < p.n;
#pragma unroll
    for (int r = 0; r < kTile; ++r)
      pi[r] = (col && r < rows_here) ? __ldg(p.index + (s0 + r) * p.index_stride + i0 + lane) : 0;
  };
  fetch(0);
  for (int64_t i0 = 0; i0 < p.n; i0 += kTile) {
    const bool col = i0 + lane < p.n;
#pragma unroll
    for (int r = 0; r < kTile; ++r) ti[r * kLd + lane] = pi[r];
    __syncwarp();
    if (i0 + kTile < p.n) fetch(i0 + kTile);
    if (active) {
      const int cnt = (int)min((int64_t)kTile, p.n - i0);
      int j = 0, phase = 0, nd = 0, q = 0;
      int32_t esc = 0, off = 0;
      uint32_t x = 0;
      while (j < cnt) {
        bool done = false;         // symbol j is complete: x / esc / off give its value
        if (phase == 0) {
          const int32_t row = min(max(ti[lane * kLd + j], 0), p.rows - 1);
          const int4 ri = t.rowinfo[row];
          const uint4 rr = t.rowrec[row];
          esc = ri.y - 2;
          off = ri.z;
          int32_t v = -1;
          if (rr.z != 0u && d.span1 != 0xFFFFFFFFu) {
            const uint32_t span = d.span1 + 1u;
            const uint32_t t1 = __umulhi(span, rr.x), t2 = __umulhi(span, rr.y),
                           t3 = __umulhi(span, rr.z);
            if (t3 > d.rem) {      // a direct value; entry 0 is 0 in this class (rc_v4.cuh step)
              const bool p1 = t1 <= d.rem, p2 = t2 <= d.rem;
              const uint32_t a = p2 ? t2 : (p1 ? t1 : 0u);
              const uint32_t b = (p2 ? t3 : (p1 ? t2 : t1)) - 1u;
              d.narrow(a, b);
              v = p2 ? 2 : (p1 ? 1 : 0);
            }
          }
          if (v < 0) v = d.get(t, ri.x, ri.y);
          if (v == esc) {          // unbounded_index_range_coding_kernels.cc:336-360
            phase = 1; nd = 0; q = 0; x = 0;
          } else {
            tv[lane * kLd + j] = v + off;
            ++j;
          }
        } else if (phase == 1) {   // digit count, groups of all1
          const uint32_t u = d.digit(dsh, all1);
          nd += (int)u;
          if (u != all1 || nd > p.max_digits) {
            if (nd > p.max_digits) { d.corrupt = true; nd = p.max_digits; }
            if (nd > 0) phase = 2; else done = true;
          }
        } else {                   // digits, least significant first
          const uint32_t dg = d.digit(dsh, all1);
          if (q * ow < 32) x |= dg << (q * ow);
          if (++q == nd) done = true;
        }
        if (done) {
          const int32_t v = (x & 1u) ? -(int32_t)(x >> 1) - 1 : (int32_t)(x >> 1) + esc;
          tv[lane * kLd + j] = v + off;
          ++j;
          phase = 0;
        }
      }
    }
    __syncwarp();
    if (col) {
#pragma unroll 8
      for (int r = 0; r < rows_here; ++r) p.values[(s0 + r) * p.n + i0 + lane] = tv[r * kLd + lane];
    }
    __syncwarp();
  }
  if (active && p.status) p.status[s] = d.corrupt ? TFC_STREAM_CORRUPT : TFC_STREAM_OK;
}


// ---- bounded RangeEncode / RangeDecode (range_coding_kernels.cc:178-380), one stream per lane ---
// data int16 in [0, chip - 1), cdf rows of `chip` entries picked per element through the merged
// broadcast pattern (the same row for every stream at element i unless the CDF is per stream).
// The table is staged in shared memory (pre-shifted to precision 16) when it is shared by the
// streams and fits; otherwise entries come from global memory through L1.
struct BoundedParams {
  const int16_t* data;        // encode: [S, n]
  const int32_t* cdf;
  int64_t cdf_stream_stride;  // elements between the tables of consecutive streams (0: shared)
  int64_t cdf_rows;           // rows of one table
  int32_t chip, precision;
  Broadcast bc;
  uint8_t* out;
  const uint8_t* enc;
  int64_t cap;
  int32_t* nbytes;
  int16_t* values;            // decode: [S, n]
  int32_t* status;
  int64_t S, n;
  int32_t table_bytes;        // > 0: shared table staged in shared memory
};

struct BoundedTable {
  const uint32_t* smem;       // staged, pre-shifted (or null)
  const int32_t* gmem;        // this lane's table in global memory
  int sh;
  __device__ __forceinline__ uint32_t at(int64_t i) const {
    return smem ? smem[i] : ((uint32_t)__ldg(gmem + i) << sh);
  }
};

__device__ __forceinline__ const uint32_t* stage_bounded(const BoundedParams& p) {
  if (p.table_bytes <= 0) return nullptr;
  uint32_t* tbl = reinterpret_cast<uint32_t*>(v4::smem_v4);
  const int64_t total = p.cdf_rows * p.chip;
  const int sh = 16 - p.precision;
  for (int64_t i = threadIdx.x; i < total; i += blockDim.x) tbl[i] = (uint32_t)__ldg(p.cdf + i) << sh;
  __syncthreads();
  return tbl;
}

__global__ void __launch_bounds__(kMaxWarps * 32, 1) encode_bounded_lanes(const BoundedParams p) {
  const uint32_t* staged = stage_bounded(p);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, wpc = blockDim.x >> 5;
  const int64_t s0 = ((int64_t)blockIdx.x * wpc + warp) * 32;
  if (s0 >= p.S) return;
  int32_t* td = reinterpret_cast<int32_t*>(v4::smem_v4 + p.table_bytes + (size_t)warp * kWarpBytes);
  const int64_t s = s0 + lane;
  const bool active = s < p.S;
  const int rows_here = (int)min((int64_t)32, p.S - s0);
  BoundedTable t{staged, p.cdf + (active ? s : s0) * p.cdf_stream_stride, 16 - p.precision};
  Enc e;
  e.lo = 0; e.span1 = 0xFFFFFFFFu; e.pos = 0; e.cap = (uint32_t)p.cap; e.held = 0;
  e.out = p.out + (active ? s : s0) * p.cap;
  e.even = (reinterpret_cast<uintptr_t>(e.out) & 1) == 0;
  for (int64_t i0 = 0; i0 < p.n; i0 += kTile) {
    const bool col = i0 + lane < p.n;
#pragma unroll 8
    for (int r = 0; r < rows_here; ++r)
      td[r * kLd + lane] = col ? (int32_t)__ldg(p.data + (s0 + r) * p.n + i0 + lane) : 0;
    __syncwarp();
    if (active) {
      const int cnt = (int)min((int64_t)kTile, p.n - i0);
      for (int j = 0; j < cnt; ++j) {
        const int64_t row = broadcast_row(p.bc, i0 + j) * p.chip;
        const int32_t v = min(max(td[lane * kLd + j], 0), p.chip - 2);
        e.put(t.at(row + v), t.at(row + v + 1));
      }
    }
    __syncwarp();
  }
  if (active) {
    e.flush();
    p.nbytes[s] = (int32_t)e.pos;
    if (p.status) p.status[s] = e.pos > e.cap ? TFC_STREAM_CAPACITY : TFC_STREAM_OK;
  }
}

__global__ void __launch_bounds__(kMaxWarps * 32, 1) decode_bounded_lanes(const BoundedParams p) {
  const uint32_t* staged = stage_bounded(p);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, wpc = blockDim.x >> 5;
  const int64_t s0 = ((int64_t)blockIdx.x * wpc + warp) * 32;
  if (s0 >= p.S) return;
  int32_t* tv = reinterpret_cast<int32_t*>(v4::smem_v4 + p.table_bytes + (size_t)warp * kWarpBytes);
  const int64_t s = s0 + lane;
  const bool active = s < p.S;
  const int rows_here = (int)min((int64_t)32, p.S - s0);
  BoundedTable t{staged, p.cdf + (active ? s : s0) * p.cdf_stream_stride, 16 - p.precision};
  Dec d;
  d.src = p.enc + (active ? s : s0) * p.cap;
  d.len = active ? (uint32_t)max(min((int64_t)p.nbytes[s], p.cap), (int64_t)0) : 0u;
  d.pos = 0; d.win = 0; d.avail = 0;
  d.corrupt = active && (int64_t)p.nbytes[s] > p.cap;   // a truncated string
  d.vec = (reinterpret_cast<uintptr_t>(d.src) & 7) == 0;
  d.span1 = 0xFFFFFFFFu;
  d.rem = 0;
  if (active) {
    const uint32_t hi = d.next16();
    d.rem = (hi << 16) | d.next16();
  }
  for (int64_t i0 = 0; i0 < p.n; i0 += kTile) {
    const bool col = i0 + lane < p.n;
    if (active) {
      const int cnt = (int)min((int64_t)kTile, p.n - i0);
      for (int j = 0; j < cnt; ++j) {
        const int64_t row = broadcast_row(p.bc, i0 + j) * p.chip;
        // smallest k in [1, chip-1] with floor(span * row[k] / 2^16) > rem  (range_coder.cc:323-340)
        const uint64_t span = (uint64_t)d.span1 + 1u;
        int first = 1, count = p.chip - 1;
        while (count > 0) {
          const int step = count >> 1;
          if (((span * t.at(row + first + step)) >> 16) <= (uint64_t)d.rem) {
            first += step + 1;
            count -= step + 1;
          } else {
            count = step;
          }
        }
        if (first > p.chip - 1) { d.corrupt = true; first = p.chip - 1; }
        const uint32_t a = (uint32_t)((span * t.at(row + first - 1)) >> 16);
        const uint32_t b = (uint32_t)((span * t.at(row + first)) >> 16) - 1u;
        d.narrow(a, b);
        tv[lane * kLd + j] = first - 1;
      }
    }
    __syncwarp();
    if (col) {
#pragma unroll 8
      for (int r = 0; r < rows_here; ++r)
        p.values[(s0 + r) * p.n + i0 + lane] = (int16_t)tv[r * kLd + lane];
    }
    __syncwarp();
  }
  if (active && p.status) p.status[s] = d.corrupt ? TFC_STREAM_CORRUPT : TFC_STREAM_OK;
}

}  // namespace lanes
