// Unbounded-index range coder kernels, second generation (included by range_coder.cu).
//
// encode: one CTA (4 warps) per stream.  Warps 1-3 stage 512-symbol super-chunks — symbol and
//         index fetch, offset shift, escape mapping and both CDF lookups, i.e. everything that
//         does not depend on the coder state — into a double-buffered shared-memory ring while
//         lane 0 of warp 0 runs the 32-bit state machine over the previous super-chunk.
// decode: one warp per stream, state replicated in all lanes.  Each lane owns one entry of a
//         32-wide window of the row (placed around the row's mode at staging time) and
//         computes its own a = (span * cdf) >> precision; the symbol is the first lane whose
//         a exceeds code - lo (one ballot), and the new interval is read from that lane and
//         its left neighbour (two shuffles) — the reference's binary search plus its two
//         multiplications collapse into one multiply per lane.  Encoded bytes are held 128 at
//         a time in registers (4 per lane) and handed out by shuffle.
// Tables are staged ragged in shared memory by both kernels.
#pragma once

namespace {

constexpr int kSuper = 512;   // symbols per encode super-chunk

extern __shared__ __align__(16) uint8_t rc_smem[];

struct StagedTables {
  const int32_t* tbl;   // ragged rows (shared) or the dense table (global)
  const int4* meta;     // {start, len, window base, offset} per row (shared) or nullptr
  bool staged;
};

// Copies the rows of cdf[rows, cols] (lengths cdf_size[]) into shared memory back to back and
// fills meta[].  Returns false (nothing staged) if it does not fit in `budget` bytes.
// Must be called by all threads of the block.
__device__ bool stage_tables(const int32_t* __restrict__ cdf, int rows, int cols,
                             const int32_t* __restrict__ cdf_size,
                             const int32_t* __restrict__ offset, uint8_t* smem, int budget,
                             int* s_total, StagedTables* out) {
  int4* meta = reinterpret_cast<int4*>(smem);
  int32_t* tbl = reinterpret_cast<int32_t*>(smem + (size_t)rows * sizeof(int4));
  const int64_t meta_bytes = (int64_t)rows * (int64_t)sizeof(int4);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  bool staged = meta_bytes + 16 <= budget;
  if (staged) {
    if (threadIdx.x == 0) {
      int acc = 0;
      for (int r = 0; r < rows; ++r) {
        const int len = min(max(__ldg(cdf_size + r), 0), cols);
        meta[r] = make_int4(acc, len, 0, __ldg(offset + r));
        acc += len;
      }
      *s_total = acc;
    }
    __syncthreads();
    staged = meta_bytes + (int64_t)(*s_total) * 4 <= budget;
  }
  if (staged) {
    for (int r = warp; r < rows; r += nwarps) {
      const int start = meta[r].x, len = meta[r].y;
      const int32_t* src = cdf + (int64_t)r * cols;
      int best = 0, bestk = 1;
      for (int k = lane; k < len; k += 32) {
        const int32_t c = __ldg(src + k);
        tbl[start + k] = c;
        if (k + 1 < len) {
          const int pm = __ldg(src + k + 1) - c;
          if (pm > best) { best = pm; bestk = k + 1; }
        }
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const int ob = __shfl_xor_sync(0xFFFFFFFFu, best, o);
        const int ok = __shfl_xor_sync(0xFFFFFFFFu, bestk, o);
        if (ob > best || (ob == best && ok < bestk)) { best = ob; bestk = ok; }
      }
      if (lane == 0) meta[r].z = len <= 32 ? 0 : min(max(bestk - 16, 0), len - 32);
    }
  }
  __syncthreads();
  out->staged = staged;
  out->tbl = staged ? tbl : cdf;
  out->meta = staged ? meta : nullptr;
  return staged;
}

// ------------------------------------------------------------------------------------------
// encode
// ------------------------------------------------------------------------------------------
template <bool STAGED>
__device__ __forceinline__ void stage_symbols(const EncodeParams& p, const StagedTables& t,
                                              const int32_t* __restrict__ d32,
                                              const int32_t* __restrict__ idx, int64_t base,
                                              int cnt, uint32_t* packed, uint32_t* extra,
                                              int tid, int nthreads) {
  for (int j = tid; j < cnt; j += nthreads) {
    const int32_t r = __ldg(idx + base + j);
    int start, len, off;
    if (STAGED) {
      const int4 m = t.meta[r];
      start = m.x; len = m.y; off = m.w;
    } else {
      start = r * p.cols; len = __ldg(p.cdf_size + r); off = __ldg(p.offset + r);
    }
    const int32_t top = len - 2;
    int32_t v = __ldg(d32 + base + j) - off;
    uint32_t x = 0;
    if (v < 0) { x = (uint32_t)(-2 * v - 1) + 1u; v = top; }
    else if (v >= top) { x = (uint32_t)(2 * (v - top)) + 1u; v = top; }
    const uint32_t c0 = (uint32_t)t.tbl[start + v];
    const uint32_t c1 = (uint32_t)t.tbl[start + v + 1];
    packed[j] = (c0 & 0xFFFFu) | ((c1 - 1u) << 16);
    extra[j] = x;   // 0: plain symbol; otherwise zig-zagged excess + 1
  }
}

// Rare paths of the encoder, kept out of line so that the per-symbol loop stays a few dozen
// instructions (it runs in a single warp per SM: an instruction-cache miss is fully exposed).
struct EncState { uint32_t lo, span1, pos; uint64_t held; };

__device__ __noinline__ EncState enc_slow(EncState s, bool wrapped, uint8_t* out, uint32_t cap) {
  // called after the interval update when it straddles 2^32 or a carry is pending
  auto byte = [&](uint32_t b) { if (s.pos < cap) out[s.pos] = (uint8_t)b; ++s.pos; };
  if (s.lo + s.span1 < s.lo) {
    if ((s.span1 >> 16) == 0) {
      s.lo <<= 16;
      s.span1 = (s.span1 << 16) | 0xFFFFu;
      s.held += 0x20000;
    }
    return s;
  }
  if (s.held != 0) {
    if (!wrapped) --s.held;
    byte(((uint32_t)s.held >> 8) & 0xFFu);
    byte((uint32_t)s.held & 0xFFu);
    const uint32_t fill = wrapped ? 0x00u : 0xFFu;
    for (uint64_t k = s.held >> 16; k != 0; --k) byte(fill);
    s.held = 0;
  }
  if ((s.span1 >> 16) == 0) {
    const uint32_t hi = s.lo >> 16;
    s.lo <<= 16;
    s.span1 = (s.span1 << 16) | 0xFFFFu;
    if (s.lo + s.span1 >= s.lo) { byte(hi >> 8); byte(hi & 0xFFu); } else s.held = hi + 1;
  }
  return s;
}

// One encoder step.  The straight-line path (interval stays >= 2^16 wide, no carry pending)
// has a single, rarely taken branch; everything else (renormalisation + byte emission, carry
// bookkeeping) is behind it.
__device__ __forceinline__ void enc_put(EncState& s, uint32_t c0, uint32_t c1, int prec,
                                        uint8_t* out, uint32_t cap) {
  const uint32_t a = (uint32_t)(((uint64_t)s.span1 * c0 + c0) >> prec);
  const uint32_t b = (uint32_t)(((uint64_t)s.span1 * c1 + c1) >> prec) - 1u;
  s.lo += a;
  s.span1 = b - a;
  const bool wrapped = s.lo < a;
  // rare := straddles 2^32  |  carry pending  |  needs renormalisation
  if ((s.lo + s.span1 < s.lo) | (s.held != 0) | ((s.span1 >> 16) == 0)) {
    if ((s.lo + s.span1 < s.lo) | (s.held != 0)) {
      s = enc_slow(s, wrapped, out, cap);
      return;
    }
    const uint32_t hi = s.lo >> 16;
    s.lo <<= 16;
    s.span1 = (s.span1 << 16) | 0xFFFFu;
    if (s.lo + s.span1 >= s.lo) {
      if (s.pos + 2 <= cap) { out[s.pos] = (uint8_t)(hi >> 8); out[s.pos + 1] = (uint8_t)hi; }
      s.pos += 2;
    } else {
      s.held = hi + 1;
    }
  }
}

__device__ __noinline__ EncState enc_escape(EncState s, uint32_t x, int w, uint8_t* out,
                                            uint32_t cap) {
  const uint32_t all1 = (1u << w) - 1u;
  int nd = 0;
  while (nd * w < 32 && (x >> (nd * w)) != 0) ++nd;
  uint32_t u = (uint32_t)nd;
  for (; u >= all1; u -= all1) enc_put(s, all1, all1 + 1, w, out, cap);
  enc_put(s, u, u + 1, w, out, cap);
  for (int j = 0; j < nd; ++j) {
    const uint32_t d = (x >> (j * w)) & all1;
    enc_put(s, d, d + 1, w, out, cap);
  }
  return s;
}

__global__ void __launch_bounds__(128) encode_unbounded_kernel(const EncodeParams p,
                                                               int table_budget) {
  __shared__ int s_total;
  // [packed 2 x kSuper][extra 2 x kSuper][tables ...]
  uint32_t* packed = reinterpret_cast<uint32_t*>(rc_smem);
  uint32_t* extra = packed + 2 * kSuper;
  uint8_t* tsm = reinterpret_cast<uint8_t*>(extra + 2 * kSuper);
  StagedTables t;
  stage_tables(p.cdf, p.rows, p.cols, p.cdf_size, p.offset, tsm, table_budget, &s_total, &t);

  const int64_t s = blockIdx.x;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int32_t* idx = p.index + s * p.index_stride;
  const int32_t* d32 = p.data32 + s * p.n;
  uint8_t* out = p.out + s * p.cap;
  const uint32_t cap = (uint32_t)p.cap;
  const int prec = p.precision, ow = p.overflow_width;
  EncState e{0u, 0xFFFFFFFFu, 0u, 0ull};
  const int64_t nchunks = (p.n + kSuper - 1) / kSuper;

  auto stage = [&](int64_t k, int tid, int nthreads) {
    const int64_t base = k * kSuper;
    const int cnt = (int)min((int64_t)kSuper, p.n - base);
    uint32_t* pk = packed + (k & 1) * kSuper;
    uint32_t* ex = extra + (k & 1) * kSuper;
    if (t.staged) stage_symbols<true>(p, t, d32, idx, base, cnt, pk, ex, tid, nthreads);
    else stage_symbols<false>(p, t, d32, idx, base, cnt, pk, ex, tid, nthreads);
  };

  if (nchunks > 0) stage(0, threadIdx.x, 128);
  __syncthreads();
  for (int64_t k = 0; k < nchunks; ++k) {
    if (warp == 0) {
      if (lane == 0) {
        const int cnt = (int)min((int64_t)kSuper, p.n - k * kSuper);
        const uint32_t* pk = packed + (k & 1) * kSuper;
        const uint32_t* ex = extra + (k & 1) * kSuper;
        uint32_t w = pk[0], x = ex[0];
#pragma unroll 1
        for (int j = 0; j < cnt; ++j) {
          const uint32_t cw = w, cx = x;
          const int jn = min(j + 1, kSuper - 1);
          w = pk[jn];            // next symbol's words are in flight while this one is coded
          x = ex[jn];
          enc_put(e, cw & 0xFFFFu, (cw >> 16) + 1u, prec, out, cap);
          if (cx != 0) e = enc_escape(e, cx - 1u, ow, out, cap);
        }
      }
    } else if (k + 1 < nchunks) {
      stage(k + 1, threadIdx.x - 32, 96);
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    // Finalize (range_coder.cc:268-301)
    uint32_t wd = 0;
    bool any = true;
    if (e.held != 0) wd = (uint32_t)e.held & 0xFFFFu;
    else if (e.lo != 0) wd = ((e.lo - 1) >> 16) + 1;
    else any = false;
    if (any) {
      if (e.pos < cap) out[e.pos] = (uint8_t)(wd >> 8);
      ++e.pos;
      if (wd & 0xFFu) { if (e.pos < cap) out[e.pos] = (uint8_t)wd; ++e.pos; }
    }
    p.nbytes[s] = (int32_t)e.pos;
    if (p.status) p.status[s] = e.pos > cap ? TFC_STREAM_CAPACITY : TFC_STREAM_OK;
  }
}

// ------------------------------------------------------------------------------------------
// decode
// ------------------------------------------------------------------------------------------
struct FastDecoder {
  uint32_t lo, span1, code;
  uint32_t buf;                    // the next two 16-bit words of the stream, first one on top
  uint32_t bufcnt;                 // words left in buf (2 or 1)
  uint32_t widx;                   // index of the next 32-bit stream word to put into buf
  const uint8_t* src;
  uint32_t len;
  uint32_t cur, ahead;             // this lane's 4 bytes of the current / next 128-byte chunk
  uint32_t chunk;                  // 32-bit-word index at which the current chunk starts
  bool corrupt;
  int lane;

  __device__ __forceinline__ uint32_t load4(uint32_t at) const {  // big-endian, zero past end
    uint32_t v = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      v <<= 8;
      if (at + k < len) v |= (uint32_t)__ldg(src + at + k);
    }
    return v;
  }
  __device__ __forceinline__ uint32_t word32(uint32_t wi) {   // wi >= chunk, uniform
    if (wi - chunk >= 32u) {
      chunk += 32u;
      cur = ahead;
      ahead = load4(4u * (chunk + 32u + lane));
    }
    return __shfl_sync(0xFFFFFFFFu, cur, (wi - chunk) & 31);
  }
  __device__ __forceinline__ void init(const uint8_t* s, uint32_t n, int ln) {
    src = s; len = n; lane = ln; lo = 0; span1 = 0xFFFFFFFFu; corrupt = false;
    chunk = 0;
    cur = load4(4u * lane);
    ahead = load4(128u + 4u * lane);
    code = word32(0);
    buf = word32(1);
    bufcnt = 2;
    widx = 2;
  }
  __device__ __forceinline__ void renorm() {
    lo <<= 16;
    span1 = (span1 << 16) | 0xFFFFu;
    code = (code << 16) | (buf >> 16);
    buf <<= 16;
    if (--bufcnt == 0) { buf = word32(widx); ++widx; bufcnt = 2; }
  }
  // a = floor(span * c / 2^prec) mod 2^32 ; span = span1 + 1
  __device__ __forceinline__ uint32_t scale(uint32_t c, int prec) const {
    return (uint32_t)(((uint64_t)span1 * c + c) >> prec);
  }
  __device__ __forceinline__ void narrow_a(uint32_t a0, uint32_t a1) {
    lo += a0;
    span1 = a1 - 1u - a0;
    if ((span1 >> 16) == 0) renorm();
  }
  // is span * c > target, i.e. floor(span*c/2^prec) > code - lo  (c == 2^prec always is)
  __device__ __forceinline__ bool above(uint32_t c, int prec) const {
    return (c >> prec) != 0 || scale(c, prec) > code - lo;
  }
  __device__ __forceinline__ uint32_t digit(int w) {   // uniform alphabet of 2^w digits
    const uint64_t span = (uint64_t)span1 + 1u;
    const uint64_t tgt = (((uint64_t)(uint32_t)(code - lo) + 1u) << w) - 1u;
    const uint64_t d = tgt / span;
    if (d >> w) corrupt = true;
    const uint32_t dd = (uint32_t)min(d, (uint64_t)((1u << w) - 1u));
    narrow_a(scale(dd, w), scale(dd + 1u, w));
    return dd;
  }
};

// 32-ary ballot search for the smallest k in [lo, hi] with span*row[k] > target
template <typename RowPtr>
__device__ __forceinline__ int wide_search2(const FastDecoder& d, RowPtr row, int lo, int hi,
                                            int prec, int lane) {
  while (hi - lo + 1 > 32) {
    const int step = (hi - lo + 1 + 31) >> 5;
    const int k = min(lo + (lane + 1) * step - 1, hi);
    const uint32_t b = __ballot_sync(0xFFFFFFFFu, d.above((uint32_t)row[k], prec));
    if (b == 0) return hi + 1;
    const int f = __ffs(b) - 1;
    const int nlo = lo + f * step;
    hi = min(nlo + step - 1, hi);
    lo = nlo;
  }
  const int k = lo + lane;
  const uint32_t b =
      __ballot_sync(0xFFFFFFFFu, k <= hi && d.above((uint32_t)row[min(k, hi)], prec));
  if (b == 0) return hi + 1;
  return lo + __ffs(b) - 1;
}

// Rare decode paths, out of line (see the encoder note on instruction-cache footprint).
struct DecCore { uint32_t lo, span1, code; };

template <bool STAGED>
__device__ __forceinline__ uint32_t table_at(const DecodeParams& p, int i) {
  if (STAGED) return (uint32_t)reinterpret_cast<const int32_t*>(rc_smem + (size_t)p.rows * 16)[i];
  return (uint32_t)__ldg(p.cdf + i);
}

template <bool STAGED>
__device__ __forceinline__ int4 row_meta(const DecodeParams& p, int32_t r) {
  if (STAGED) return reinterpret_cast<const int4*>(rc_smem)[r];
  const int len = __ldg(p.cdf_size + r);
  return make_int4(r * p.cols, len, len <= 32 ? 0 : min(max(len / 2 - 16, 0), len - 32),
                   __ldg(p.offset + r));
}

// symbol search outside the window: returns k (1..len-1), or -(len-1) for a corrupt stream
template <bool STAGED>
__device__ __noinline__ int decode_slow_search(const DecodeParams& p, FastDecoder d, int start,
                                               int len, int wb, int f, int lane) {
  const int32_t* row = STAGED
      ? reinterpret_cast<const int32_t*>(rc_smem + (size_t)p.rows * 16) + start
      : p.cdf + start;
  int k;
  if (f == 0) k = wb == 0 ? len : wide_search2(d, row, 1, wb, p.precision, lane);
  else k = (wb + 32 <= len - 1) ? wide_search2(d, row, wb + 32, len - 1, p.precision, lane) : len;
  if (k > len - 1) k = -(len - 1);   // negative: ran past the row (corrupt stream)
  return k;
}

// escape tail (unbounded_index_range_coding_kernels.cc:340-360); returns the decoded value
struct EscResult { FastDecoder d; int32_t value; };
__device__ __noinline__ EscResult decode_escape(FastDecoder d, int w, int32_t top) {
  const uint32_t all1 = (1u << w) - 1u;
  int nd = 0;
  uint32_t u;
  do { u = d.digit(w); nd += (int)u; } while (u == all1 && nd < 64 && !d.corrupt);
  uint32_t x = 0;
  for (int q = 0; q < nd; ++q) {
    const uint32_t dg = d.digit(w);
    if (q * w < 32) x |= dg << (q * w);
  }
  EscResult r;
  r.d = d;
  r.value = (x & 1u) ? -(int32_t)(x >> 1) - 1 : (int32_t)(x >> 1) + top;
  return r;
}

constexpr int kDecScratch = 32 * 32 * 4 + 32 * 16;   // per-warp window matrix + row info

template <bool STAGED>
__device__ void decode_stream2(const DecodeParams& p, int64_t s, int lane, uint8_t* scratch) {
  uint32_t* s_win = reinterpret_cast<uint32_t*>(scratch);            // [symbol][lane]
  int4* s_gm = reinterpret_cast<int4*>(scratch + 32 * 32 * 4);       // {start, len, wb, off}
  FastDecoder d;
  d.init(p.enc + s * p.cap, tfc::clamp_stream_len(p.nbytes[s], p.cap), lane);
  d.corrupt = (int64_t)p.nbytes[s] > p.cap;   // a truncated string
  const int32_t* idx = p.index + s * p.index_stride;
  int32_t* out = p.out32 + s * p.n;
  const int prec = p.precision, w = p.overflow_width;

  int32_t myr_next = lane < p.n ? __ldg(idx + lane) : 0;
  for (int64_t base = 0; base < p.n; base += 32) {
    const int cnt = (int)min((int64_t)32, p.n - base);
    // ---- phase A (state independent, all lanes in parallel): row info of the group's 32
    //      symbols and, for each, the 32-entry CDF window one lane per entry
    const int32_t myr = (lane < cnt) ? myr_next : 0;
    if (base + 32 + lane < p.n) myr_next = __ldg(idx + base + 32 + lane);
    s_gm[lane] = row_meta<STAGED>(p, myr);
    __syncwarp();
#pragma unroll 8
    for (int q = 0; q < 32; ++q) {
      const int4 mq = s_gm[q];
      const int e = mq.z + lane;
      s_win[q * 32 + lane] = e < mq.y ? table_at<STAGED>(p, mq.x + e) : 0xFFFFFFFFu;
    }
    __syncwarp();
    // ---- phase B: the serial chain
    int32_t mine = 0;
    uint32_t wv = s_win[lane];
    int4 m = s_gm[0];
#pragma unroll 1
    for (int j = 0; j < cnt; ++j) {
      const int jn = min(j + 1, 31);
      const uint32_t nwv = s_win[jn * 32 + lane];   // next symbol's operands, state independent
      const int4 mn = s_gm[jn];
      const uint32_t a = d.scale(wv, prec);
      const bool hit = wv != 0xFFFFFFFFu && ((wv >> prec) != 0 || a > d.code - d.lo);
      const uint32_t b = __ballot_sync(0xFFFFFFFFu, hit);
      const int f = __ffs(b) - 1;
      int k = m.z + f;
      if (f >= 1) {
        const uint32_t a0 = __shfl_sync(0xFFFFFFFFu, a, f - 1);
        const uint32_t a1 = __shfl_sync(0xFFFFFFFFu, a, f);
        d.lo += a0;
        d.span1 = a1 - 1u - a0;
      } else {
        k = decode_slow_search<STAGED>(p, d, m.x, m.y, m.z, f, lane);
        if (k < 0) { d.corrupt = true; k = -k; }
        const uint32_t a0 = d.scale(table_at<STAGED>(p, m.x + k - 1), prec);
        const uint32_t a1 = d.scale(table_at<STAGED>(p, m.x + k), prec);
        d.lo += a0;
        d.span1 = a1 - 1u - a0;
      }
      if ((d.span1 >> 16) == 0) d.renorm();
      int32_t v = k - 1;
      if (v == m.y - 2) { const EscResult er = decode_escape(d, w, v); d = er.d; v = er.value; }
      mine = (lane == j) ? v + m.w : mine;
      m = mn;
      wv = nwv;
    }
    if (lane < cnt) out[base + lane] = mine;
    __syncwarp();
  }
  if (lane == 0 && p.status) p.status[s] = d.corrupt ? TFC_STREAM_CORRUPT : TFC_STREAM_OK;
}

__global__ void __launch_bounds__(256, 1) decode_unbounded_kernel2(const DecodeParams p) {
  __shared__ int s_total;
  StagedTables t;
  stage_tables(p.cdf, p.rows, p.cols, p.cdf_size, p.offset, rc_smem, p.table_budget, &s_total,
               &t);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t s = (int64_t)blockIdx.x * (blockDim.x >> 5) + warp;
  if (s >= p.S) return;
  uint8_t* scratch = rc_smem + ((p.table_budget + 15) & ~15) + warp * kDecScratch;
  if (t.staged) decode_stream2<true>(p, s, lane, scratch);
  else decode_stream2<false>(p, s, lane, scratch);
}

}  // namespace
