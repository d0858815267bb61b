// Unbounded-index range coder kernels, generation 4 (included by range_coder.cu).
//
// The coder state is a serial chain per stream, so everything here is about making that chain
// short and keeping everything else off it:
//   * all CDF entries are pre-shifted to precision 16 when the tables are staged in shared
//     memory (floor(span*c / 2^p) == floor(span*(c << (16-p)) / 2^16)), so the per-symbol code
//     has no variable shifts;
//   * decode: one warp runs the chain for a stream with the state REPLICATED in all lanes (no
//     shuffles, no ballots, no divergence; lane 0 stores).  A second "helper" warp per stream
//     turns the next round's indexes into per-symbol records {c1, c2, c3, meta} (rows of up to
//     4 symbols are decided by three independent multiplies + compares — no search), keeps a
//     ring of big-endian stream words filled in shared memory, and writes the previous
//     round's symbols back coalesced (adding the row offset).  Longer rows and escapes take an
//     out-of-line binary search over the staged table.
//   * encode: see the second half of this file.
// Bit-exact with RangeEncoder/RangeDecoder (range_coder.cc:38-377) and the op loops of
// unbounded_index_range_coding_kernels.cc:189-251, :308-368.
#pragma once

#include <type_traits>

namespace v4 {

constexpr int kRing = 8192;            // stream ring: 16-bit words per stream
constexpr int kRoundMax = 256;         // symbols per round (upper bound)
constexpr int kMaxRows = 65535;

extern __shared__ __align__(16) uint8_t smem_v4[];

// Occupancy trace (tfc_debug_coder_trace): nanoseconds between a stream's first and last
// instruction, summed per kernel kind — how long coder CTAs really HOLD their SM when other
// kernels share the GPU (end-of-phase events on the stream cannot tell waiting for an SM from
// running on it).  [0] encode ns, [1] encode streams, [2] decode ns, [3] decode streams.
__device__ unsigned long long g_trace[4];
__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

struct Tables {          // shared-memory image of the CDF tables
  const uint4* rowrec;   // {c1<<16, c2<<16, c3<<16, row} (precision 16) for 5-entry rows, else c=0
  const int4* rowinfo;   // {start, len, offset, entry 0 << (16 - precision)}
  const uint32_t* tbl;   // ragged rows, entries << (16 - precision); null if they did not fit
  const int32_t* gcdf;   // the caller's dense table (used when tbl is null; start = row*cols)
  int sh;
  __device__ __forceinline__ uint32_t at(int i) const {
    return tbl ? tbl[i] : ((uint32_t)__ldg(gcdf + i) << sh);
  }
};

__host__ __device__ inline size_t tables_bytes(int64_t rows, int64_t total_entries) {
  return (size_t)rows * 32 + (size_t)total_entries * 4;
}

// All threads of the block.  The ragged rows are staged if they fit in `budget` bytes (the
// host cannot know: the row lengths live on the device); row records always are.
__device__ void stage_tables_v4(const int32_t* __restrict__ cdf, int rows, int cols,
                                const int32_t* __restrict__ cdf_size,
                                const int32_t* __restrict__ offset, int precision,
                                uint8_t* smem, int budget, Tables* t) {
  __shared__ int s_total;
  uint4* rowrec = reinterpret_cast<uint4*>(smem);
  int4* rowinfo = reinterpret_cast<int4*>(smem + (size_t)rows * 16);
  uint32_t* tbl = reinterpret_cast<uint32_t*>(smem + (size_t)rows * 32);
  const int sh = 16 - precision;
  if (threadIdx.x < 32) {       // warp 0: exclusive scan of the row lengths
    const int lane = threadIdx.x;
    int carry = 0;
    for (int r0 = 0; r0 < rows; r0 += 32) {
      const int r = r0 + lane;
      const int len = r < rows ? min(max(__ldg(cdf_size + r), 2), cols) : 0;
      int incl = len;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(0xFFFFFFFFu, incl, o);
        if (lane >= o) incl += v;
      }
      if (r < rows)   // .w: entry 0 at precision 16 (0 in a well-formed table)
        rowinfo[r] = make_int4(carry + incl - len, len, __ldg(offset + r),
                               (int)((uint32_t)__ldg(cdf + (int64_t)r * cols) << (16 - precision)));
      carry += __shfl_sync(0xFFFFFFFFu, incl, 31);
    }
    if (lane == 0) s_total = carry;
  }
  __syncthreads();
  const bool fits = tables_bytes(rows, s_total) <= (size_t)budget;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  for (int r = warp; r < rows; r += nwarps) {
    int4 ri = rowinfo[r];
    const int32_t* src = cdf + (int64_t)r * cols;
    if (fits) {
      for (int k = lane; k < ri.y; k += 32) tbl[ri.x + k] = (uint32_t)__ldg(src + k) << sh;
    } else if (lane == 0) {
      rowinfo[r].x = r * cols;
    }
    if (lane == 0) {
      // fast class: rows of exactly 5 entries (3 coded values + escape; every row of a
      // SymmetricConditional table with scale < ~0.35).  c' = c << 16 so that
      // floor(span * c / 2^16) is one IMAD.HI.  Every other row gets c3' = 0, which makes the
      // fast path's escape test (t3 <= rem) fire and sends the symbol to the general path.
      uint32_t c1 = 0, c2 = 0, c3 = 0;
      if (ri.y == 5) {
        c1 = (uint32_t)__ldg(src + 1) << sh;
        c2 = (uint32_t)__ldg(src + 2) << sh;
        c3 = (uint32_t)__ldg(src + 3) << sh;
        if (c1 > 0xFFFFu || c2 > 0xFFFFu || c3 > 0xFFFFu) c1 = c2 = c3 = 0;  // malformed
      }
      // rows outside the fast class carry their staged position instead (z == 0 marks them)
      if (c3 == 0) rowrec[r] = make_uint4((uint32_t)(fits ? ri.x : r * cols), (uint32_t)ri.y, 0u, (uint32_t)r);
      else rowrec[r] = make_uint4(c1 << 16, c2 << 16, c3 << 16, (uint32_t)r);
    }
  }
  __syncthreads();
  t->rowrec = rowrec; t->rowinfo = rowinfo; t->tbl = fits ? tbl : nullptr;
  t->gcdf = cdf; t->sh = sh;
}

__device__ __forceinline__ uint32_t scale16(uint32_t span1, uint32_t c) {
  return (uint32_t)(((uint64_t)span1 * c + c) >> 16);
}

__device__ __forceinline__ void pair_barrier(int id) {
  asm volatile("bar.sync %0, 64;" ::"r"(id) : "memory");
}

// ------------------------------------------------------------------------------------------
// decode
// ------------------------------------------------------------------------------------------
struct DecState {
  uint32_t rem;      // code - lo  (all the decoder ever needs of the two)
  uint32_t span1;    // span - 1
  uint32_t w0;       // the next stream word
  uint32_t wa;       // shared-space address of the word after w0 (linear window)
  bool corrupt;

  __device__ __forceinline__ void narrow(uint32_t a, uint32_t b) {
    rem -= a;
    span1 = b - a;
    if ((span1 >> 16) == 0) {
      span1 = (span1 << 16) | 0xFFFFu;
      rem = (rem << 16) | w0;
      uint16_t v;
      asm volatile("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(wa));
      w0 = v;
      wa += 2;
    }
  }
  // largest d in [0, all1] with floor(span * d / 2^w) <= rem, then narrow to [d, d+1)
  __device__ __forceinline__ uint32_t digit(int sh, uint32_t all1) {
    uint32_t lo = 0, hi = all1;
    while (lo < hi) {
      const uint32_t mid = (lo + hi + 1) >> 1;
      if (scale16(span1, mid << sh) <= rem) lo = mid; else hi = mid - 1;
    }
    const uint32_t c1 = (lo + 1) << sh;
    narrow(scale16(span1, lo << sh), (c1 >> 16) ? span1 : scale16(span1, c1) - 1u);
    return lo;
  }
};

// Escape tail (unbounded_index_range_coding_kernels.cc:336-360): digits of the uniform 2^ow
// alphabet; returns the zig-zag decoded value relative to `top` (= max_value of the row).
__device__ __forceinline__ int32_t decode_escape(DecState& d, int ow, int max_digits, int32_t top) {
  const uint32_t all1 = (1u << ow) - 1u;
  const int sh = 16 - ow;
  int nd = 0;
  uint32_t u;
  do { u = d.digit(sh, all1); nd += (int)u; } while (u == all1 && nd <= max_digits);
  if (nd > max_digits) { d.corrupt = true; nd = max_digits; }
  uint32_t x = 0;
  for (int q = 0; q < nd; ++q) {
    const uint32_t dg = d.digit(sh, all1);
    if (q * ow < 32) x |= dg << (q * ow);
  }
  return (x & 1u) ? -(int32_t)(x >> 1) - 1 : (int32_t)(x >> 1) + top;
}

// General symbol (any row), called by the whole chain warp: the decoder state is replicated in
// every lane, so each lane evaluates ONE candidate entry of the row and a ballot picks the
// bracket — a 32-ary search, one step for rows of up to 33 entries (every EntropyBottleneck
// row) — then the escape tail (unbounded_index_range_coding_kernels.cc:336-360).  Returns the
// value before the row offset.
__device__ __forceinline__ int32_t decode_general(DecState& d, const Tables& tb, int start,
                                                  int len, int ow, int max_digits, int lane) {
  // smallest k in [lo, hi] with span * row[k] > target; row[len - 1] = 2^16 always qualifies
  int lo = 1, hi = len - 1;
  while (hi - lo + 1 > 32) {
    const int step = (hi - lo + 1 + 31) >> 5;
    const int k = min(lo + (lane + 1) * step - 1, hi);
    const uint32_t c = tb.at(start + k);
    const uint32_t m = __ballot_sync(0xFFFFFFFFu, (c >> 16) != 0 || scale16(d.span1, c) > d.rem);
    if (m == 0) { lo = hi; break; }
    const int f = __ffs(m) - 1;
    const int nlo = lo + f * step;
    hi = min(nlo + step - 1, hi);
    lo = nlo;
  }
  uint32_t a, b1;     // floor(span*row[k-1]/2^16), floor(span*row[k]/2^16) (mod 2^32)
  bool top;           // row[k] == 2^16
  {
    const int k = min(lo + lane, hi);
    const uint32_t c = tb.at(start + k);
    const uint32_t tk = scale16(d.span1, c);
    const bool over = (c >> 16) != 0;
    const uint32_t m = __ballot_sync(0xFFFFFFFFu, over || tk > d.rem);
    int f = __ffs(m) - 1;
    if (m == 0) { d.corrupt = true; f = hi - lo; }     // ran past the row: malformed input
    f = min(f, hi - lo);
    b1 = __shfl_sync(0xFFFFFFFFu, tk, f);
    top = __shfl_sync(0xFFFFFFFFu, (int)over, f) != 0;
    const uint32_t prev = __shfl_sync(0xFFFFFFFFu, tk, max(f - 1, 0));
    a = prev;
    if (f == 0) a = scale16(d.span1, tb.at(start + lo - 1));
    lo += f;
  }
  d.narrow(a, top ? d.span1 : b1 - 1u);
  int32_t v = lo - 1;
  if (v == len - 2) v = decode_escape(d, ow, max_digits, v);
  return v;
}

struct DecParamsV4 {
  const uint8_t* enc;
  int64_t cap;
  const int32_t* nbytes;
  const int32_t* index;
  int64_t index_stride;
  const int32_t* cdf;
  const int32_t* cdf_size;
  const int32_t* offset;
  int32_t rows, cols;
  int32_t precision, overflow_width;
  int32_t* out;
  int32_t* status;
  int64_t S, n;
  int32_t round;          // symbols per round (multiple of 32, <= kRoundMax)
  int32_t max_digits;     // bound on the escape digits of one symbol
  int32_t round_words;    // bound on the stream words one round can consume
  int32_t table_bytes;    // shared-memory budget of the staged tables (16-byte multiple)
  int32_t streams_per_cta;
  int32_t lin_words;      // words of one linear window (>= 2 * round_words + 16, multiple of 8)
};

constexpr int kRecPad = 16;    // the chain prefetches records up to 15 past the round's end
constexpr int kLinMax = 4608;  // words of one linear stream window (>= 2 * round words + 16)

// per-stream shared memory of the decoder for rounds of `round` symbols (multiple of 32) and
// windows of `lin` words (multiple of 8)
__host__ __device__ inline size_t dec_pair_bytes(int round, int lin) {
  return (size_t)2 * (round + kRecPad) * 16   // records
         + (size_t)2 * round * 4              // values decoded by the general path
         + (size_t)2 * round                  // one code byte per symbol
         + (size_t)2 * lin * 2 + 32;          // two linear stream windows, ctrl[8]
}

// Stream words reach the chain through per-round LINEAR windows (no index masking on the
// chain): the window of round r+1 is filled while round r runs and starts at the stream
// position round r started at (the only position known by then); a round consumes at most
// p.round_words words, so 2 * round_words + 8 words cover every position round r+1 can need.
__global__ void __launch_bounds__(512, 1) decode_unbounded_v4(const DecParamsV4 p) {
  const unsigned long long trace_t0 = global_ns();
  Tables t;
  stage_tables_v4(p.cdf, p.rows, p.cols, p.cdf_size, p.offset, p.precision, smem_v4,
                  p.table_bytes, &t);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int spc = p.streams_per_cta;       // warps [0, spc): chains, [spc, 2 spc): helpers
  const bool helper = warp >= spc;
  const int pair = helper ? warp - spc : warp;
  const int64_t s = (int64_t)blockIdx.x * p.streams_per_cta + pair;
  if (s >= p.S) return;
  const int kRmax = p.round, kLmax = p.lin_words;
  uint8_t* mine = smem_v4 + p.table_bytes + (size_t)pair * dec_pair_bytes(p.round, p.lin_words);
  uint4* recs = reinterpret_cast<uint4*>(mine);                      // [2][kRmax + kRecPad]
  int32_t* vals = reinterpret_cast<int32_t*>(recs + 2 * (kRmax + kRecPad));  // [2][kRmax]
  uint8_t* codes = reinterpret_cast<uint8_t*>(vals + 2 * kRmax);             // [2][kRmax]
  uint16_t* lin = reinterpret_cast<uint16_t*>(codes + 2 * kRmax);            // [2][kLmax]
  volatile uint32_t* ctrl = reinterpret_cast<volatile uint32_t*>(lin + 2 * kLmax);
  // ctrl[r & 1]: stream position (in words) at which round r starts
  const int bar = 1 + pair;
  const int R = p.round;
  // Window of round r+1: starts at the stream position round r started at, rounded down to 8
  // words (16-byte loads), and must reach the last word round r+1 can touch.  Worst case that
  // is 2 * round_words + 16 words (every symbol an escape with all its digits); the helper
  // fills only what streams without escape bursts need — a word per symbol — and the chain,
  // which checks its read position once per group of symbols, completes the window itself in
  // the rare case it gets near the end of the filled part.
  const uint32_t win_max = 2u * (uint32_t)p.round_words + 16u;                 // <= kLmax
  const uint32_t win_typ = min(win_max, (2u * (uint32_t)p.round + 72u + 7u) & ~7u);
  const uint8_t* src = p.enc + s * p.cap;
  const uint32_t len = tfc::clamp_stream_len(p.nbytes[s], p.cap);
  const bool oversize = (int64_t)p.nbytes[s] > p.cap;   // a truncated string: reported corrupt
  const bool vec = (reinterpret_cast<uintptr_t>(src) & 15) == 0;
  // dst[k] = stream word first + k for k in [k0, k1) (big-endian pairs, zeros past the end);
  // first, k0 and k1 are multiples of 8
  auto fill = [&](uint16_t* dst, uint32_t first, uint32_t k0, uint32_t k1) {
#pragma unroll 2
    for (uint32_t c = k0 / 8u + lane; c < k1 / 8u; c += 32) {
      const uint32_t at = 2u * (first + 8u * c);
      if (vec && at + 16u <= len) {
        uint4 v = __ldg(reinterpret_cast<const uint4*>(src + at));
        v.x = __byte_perm(v.x, 0, 0x2301); v.y = __byte_perm(v.y, 0, 0x2301);
        v.z = __byte_perm(v.z, 0, 0x2301); v.w = __byte_perm(v.w, 0, 0x2301);
        *reinterpret_cast<uint4*>(dst + 8u * c) = v;
      } else {
        for (uint32_t e = 0; e < 8u; ++e) {
          const uint32_t b0 = at + 2u * e;
          uint32_t v = 0;
          if (b0 < len) v = (uint32_t)__ldg(src + b0) << 8;
          if (b0 + 1 < len) v |= (uint32_t)__ldg(src + b0 + 1);
          dst[8u * c + e] = (uint16_t)v;
        }
      }
    }
  };
  const int64_t rounds = (p.n + R - 1) / R;
  const int32_t* idx = p.index + s * p.index_stride;

  if (helper) {
    int32_t* out = p.out + s * p.n;
    auto build = [&](int64_t r) {
      const int64_t base = r * R;
      const int cnt = (int)min((int64_t)R, p.n - base);
      uint4* dst = recs + (r & 1) * (kRmax + kRecPad);
      int slow = 0;       // records of this round that are not in the fast class
      uint32_t longest = 5;   // entries of the longest row of this round (fast class: 5)
#pragma unroll 4
      for (int j = lane; j < cnt; j += 32) {
        const uint4 q = t.rowrec[__ldg(idx + base + j)];
        dst[j] = q;
        slow += q.z == 0u;
        longest = max(longest, q.z == 0u ? q.y : 5u);
      }
      slow = __reduce_add_sync(0xFFFFFFFFu, slow);
      longest = __reduce_max_sync(0xFFFFFFFFu, longest);
      if (lane == 0) { ctrl[2 + (r & 1)] = (uint32_t)slow; ctrl[4 + (r & 1)] = longest; }
    };
    auto flush = [&](int64_t r) {
      const int64_t base = r * R;
      const int cnt = (int)min((int64_t)R, p.n - base);
      const int32_t* ev = vals + (r & 1) * kRmax;
      const uint8_t* cb = codes + (r & 1) * kRmax;
#pragma unroll 4
      for (int j = lane; j < cnt; j += 32) {
        const int32_t off = t.rowinfo[__ldg(idx + base + j)].z;
        const uint32_t c = cb[j];
        out[base + j] = (c == 0xFFu ? ev[j] : (int32_t)c) + off;
      }
    };
    fill(lin, 0, 0, win_typ);
    build(0);
    pair_barrier(bar);
    for (int64_t r = 0; r < rounds; ++r) {
      // the chain is working on round r: prepare r+1, flush r-1
      const uint32_t pos = ctrl[r & 1];
      if (r + 1 < rounds) {
        build(r + 1);
        fill(lin + ((r + 1) & 1) * kLmax, pos & ~7u, 0, win_typ);
      }
      if (r > 0) flush(r - 1);
      pair_barrier(bar);
    }
    flush(rounds - 1);
    return;
  }

  // ---- the chain (state replicated in every lane of this warp) ----
  if (lane == 0) ctrl[0] = 2;
  pair_barrier(bar);
  DecState d;
  d.corrupt = oversize;
  // fast-path state: span modulo 2^32 (0 stands for 2^32 and always takes the general path);
  // `wa` is the shared-space address of the word after w0
  const uint32_t lin_s = (uint32_t)__cvta_generic_to_shared(lin);
  auto lds16 = [](uint32_t a) {
    uint16_t v;
    asm volatile("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(a));
    return (uint32_t)v;
  };
  uint32_t rem = (lds16(lin_s) << 16) | lds16(lin_s + 2), span = 0;
  uint32_t w0 = lds16(lin_s + 4), wa = lin_s + 6;
  uint32_t win_first = 0;      // stream position of the current window's word 0
  uint32_t p_cur = 2;          // stream position of w0 at the start of the current round
  const int ow = p.overflow_width;
  for (int64_t r = 0; r < rounds; ++r) {
    const int cnt = (int)min((int64_t)R, p.n - r * R);
    const uint4* rc = recs + (r & 1) * (kRmax + kRecPad);
    int32_t* ev = vals + (r & 1) * kRmax;
    uint8_t* cb = codes + (r & 1) * kRmax;
    const uint32_t lin_r = lin_s + (uint32_t)(r & 1) * (kLmax * 2);
    // reads stay inside the part of the window the helper filled; past `lim` the chain warp
    // completes the window itself (escape bursts only)
    uint32_t lim = lin_r + 2u * (win_typ - 16u);
    auto guard = [&]() {
      if (wa > lim) {
        fill(lin + (r & 1) * kLmax, win_first, win_typ, win_max);
        __syncwarp();
        lim = 0xFFFFFFFFu;
      }
    };

    // One symbol of a 5-entry row: three multiplies, compares and selects, no branch.  `ok`
    // goes (and stays) false when a symbol needs the general path — escape, any other row
    // (c3' == 0), or span == 2^32; the group is then replayed one symbol at a time.
    auto step = [&](const uint4& q, uint8_t* dst, bool& ok) {
      const uint32_t t3 = __umulhi(span, q.z);
      const uint32_t t1 = __umulhi(span, q.x);
      const uint32_t t2 = __umulhi(span, q.y);
      ok = ok && (t3 > rem);
      const bool p1 = t1 <= rem, p2 = t2 <= rem;
      const uint32_t a = p2 ? t2 : (p1 ? t1 : 0u);
      const uint32_t hi = p2 ? t3 : (p1 ? t2 : t1);
      rem -= a;
      span = hi - a;
      *dst = (uint8_t)(p2 ? 2 : (p1 ? 1 : 0));
      // renormalisation without a branch: the word after w0 is read unconditionally (the
      // window is padded) and everything is a select
      const bool rn = span <= 65536u;
      const uint32_t wn = lds16(wa);
      span = rn ? (span << 16) : span;
      rem = rn ? (rem * 65536u + w0) : rem;
      w0 = rn ? wn : w0;
      wa += rn ? 2u : 0u;
    };

    int j = 0;
    const uint4* rp = rc;
    uint8_t* cp = cb;
    // one symbol at a time: fast step for 5-entry rows, the warp-wide search otherwise
    auto careful = [&](int stop) {
      for (; j < stop; ++j, ++rp, ++cp) {
        guard();
        const uint4 q = rp[0];
        const uint32_t k_rem = rem, k_span = span, k_w0 = w0, k_wa = wa;
        int start = (int)q.x, len = (int)q.y;
        if (q.z != 0u) {
          bool ok = true;
          step(q, cp, ok);
          if (ok) continue;
          const int4 ri = t.rowinfo[q.w];     // escape (or span == 2^32) on a fast-class row
          start = ri.x; len = ri.y;
        } else if (len <= 33) {
          // Rows of up to 33 entries (every EntropyBottleneck row, Gaussian rows up to
          // sigma ~ 5): lane l tests entry l + 1, one ballot finds the symbol, two shuffles
          // fetch its bounds; the renormalisation is all selects.
          const uint32_t span1 = span - 1u;
          const int k = min(1 + lane, len - 1);
          const uint32_t ck = t.at(start + k), cq = t.at(start + k - 1);
          const uint32_t tk = scale16(span1, ck), tq = scale16(span1, cq);
          const bool over = (ck >> 16) != 0u;
          const uint32_t mk = __ballot_sync(0xFFFFFFFFu, over || tk > rem);
          int f = __ffs(mk) - 1;
          if (mk == 0u) { d.corrupt = true; f = len - 2; }
          const uint32_t a = __shfl_sync(0xFFFFFFFFu, tq, f);
          const uint32_t b1 = __shfl_sync(0xFFFFFFFFu, tk, f);
          const bool top = __shfl_sync(0xFFFFFFFFu, (int)over, f) != 0;
          const uint32_t ns1 = (top ? span1 : b1 - 1u) - a;
          const bool rn = (ns1 >> 16) == 0u;
          const uint32_t wn = lds16(wa);
          rem -= a;
          span = (rn ? ((ns1 << 16) | 0xFFFFu) : ns1) + 1u;
          rem = rn ? ((rem << 16) | w0) : rem;
          w0 = rn ? wn : w0;
          wa += rn ? 2u : 0u;
          int32_t v = f;
          if (v == len - 2) {
            d.rem = rem; d.span1 = span - 1u; d.w0 = w0; d.wa = wa;
            v = decode_escape(d, ow, p.max_digits, v);
            rem = d.rem; span = d.span1 + 1u; w0 = d.w0; wa = d.wa;
          }
          ev[j] = v;
          *cp = 0xFF;
          continue;
        }
        d.rem = k_rem; d.span1 = k_span - 1u; d.w0 = k_w0; d.wa = k_wa;
        const int32_t v = decode_general(d, t, start, len, ow, p.max_digits, lane);
        rem = d.rem; span = d.span1 + 1u; w0 = d.w0; wa = d.wa;
        ev[j] = v;
        *cp = 0xFF;
      }
    };
    // A round with many general symbols whose rows all have <= 33 entries (every
    // EntropyBottleneck stream; Gaussian streams of a trained model at moderate scales): ONE
    // branch-free body for every symbol, whatever its class — lane l tests entry l + 1 of the
    // row, a ballot finds the symbol, two shuffles fetch its bounds.  Everything that does not
    // depend on the coder state (record -> row info -> the lane's two table entries) is loaded
    // one to three symbols ahead, so the serial chain is multiply, compare, ballot, shuffle,
    // select.
    auto uniform = [&](auto staged) {
      constexpr bool kStaged = decltype(staged)::value;     // tables in shared memory
      const uint32_t last_row = (uint32_t)(p.rows - 1);
      auto row_of = [&](int jj) {     // records past the round's end are uninitialised: clamp
        return (int)min(reinterpret_cast<const uint32_t*>(rc + jj)[3], last_row);
      };
      auto entry = [&](int i) {
        return kStaged ? t.tbl[i] : ((uint32_t)__ldg(t.gcdf + i) << t.sh);
      };
      uint32_t span1 = span - 1u;
      uint32_t ck;
      int len;
      auto entries = [&](const int4& ri) {
        len = ri.y;
        ck = entry(ri.x + min(1 + lane, len - 1));
      };
      entries(t.rowinfo[row_of(0)]);
      int4 ri1 = t.rowinfo[row_of(1)];
      int row2 = row_of(2);
      while (j < cnt) {
        guard();
        const int stop = min(j + 8, cnt);
        for (; j < stop; ++j) {
          const uint32_t c_k = ck;
          const int c_len = len;
          entries(ri1);                       // symbol j + 1
          ri1 = t.rowinfo[row2];              // symbol j + 2
          row2 = row_of(j + 3);               // symbol j + 3 (records are padded past the round)
          const uint32_t wn = lds16(wa);
          // lane l holds T_l = floor(span * row[l + 1] / 2^16) (nondecreasing in l; the entry
          // 2^16 stands for span itself).  With P_l = (T_l > rem) the symbol is the first lane
          // with P, its lower bound the largest T without P (0 if none: row[0] = 0) and its
          // upper bound the smallest T with P — two warp reductions, no ballot / find-first.
          const bool over = (c_k >> 16) != 0u;
          const uint32_t tk = scale16(span1, c_k);
          const bool pl = over || tk > rem;
          const uint32_t hi1 = over ? span1 : tk - 1u;       // upper bound - 1
          const uint32_t a = __reduce_max_sync(0xFFFFFFFFu, pl ? 0u : tk);
          const uint32_t b1 = __reduce_min_sync(0xFFFFFFFFu, pl ? hi1 : 0xFFFFFFFFu);
          const uint32_t below = __ballot_sync(0xFFFFFFFFu, !pl);
          const uint32_t ns1 = b1 - a;
          const bool rn = (ns1 >> 16) == 0u;
          rem -= a;
          span1 = rn ? ((ns1 << 16) | 0xFFFFu) : ns1;
          rem = rn ? ((rem << 16) | w0) : rem;
          w0 = rn ? wn : w0;
          wa += rn ? 2u : 0u;
          int32_t v = __popc(below);
          if (v >= c_len - 2) {
            if (v > c_len - 2) { d.corrupt = true; v = c_len - 2; }   // ran past the row
            d.rem = rem; d.span1 = span1; d.w0 = w0; d.wa = wa;
            v = decode_escape(d, ow, p.max_digits, v);
            rem = d.rem; span1 = d.span1; w0 = d.w0; wa = d.wa;
            guard();
          }
          ev[j] = v;
          cb[j] = 0xFF;
        }
      }
      span = span1 + 1u;
      rp = rc + cnt;
      cp = cb + cnt;
    };
    if (ctrl[2 + (r & 1)] * 8u > (uint32_t)cnt) {
      // too many general symbols for speculation in groups of four to pay
      if (ctrl[4 + (r & 1)] > 33u) careful(cnt);
      else if (t.tbl != nullptr) uniform(std::true_type{});
      else uniform(std::false_type{});
    }
    // Speculation in groups of kGroup symbols.  Per group the loop pays the window guard, the
    // class test of its records, the state save and three branches — on a lone warp each taken
    // branch is ~10 cycles of pipeline refill — so 8 symbols per group instead of 4 take the
    // per-symbol overhead from ~35 to ~18 cycles (the sampled SASS of the 4-group loop: 111
    // instructions per group, a quarter of them loop bookkeeping); a group that fails (an
    // escape: ~14 % of the 8-groups on the codec's streams) is replayed symbol by symbol.
    constexpr int kGroup = 8;
    while (j < cnt) {
      if (j + kGroup <= cnt) {
        uint4 q[kGroup];
#pragma unroll
        for (int i = 0; i < kGroup; ++i) q[i] = rp[i];
        bool ok = true;
        uint32_t s_rem, s_span, s_w0, s_wa;
        do {
          guard();
          s_rem = rem; s_span = span; s_w0 = w0; s_wa = wa;
          uint32_t all = q[0].z;
#pragma unroll
          for (int i = 1; i < kGroup; ++i) all = min(all, q[i].z);
          if (all == 0u) { ok = false; break; }
#pragma unroll
          for (int i = 0; i < kGroup; ++i) {
            step(q[i], cp + i, ok);
            q[i] = rp[kGroup + i];         // records are padded past the round's end (kRecPad)
          }
          if (!ok) break;
          j += kGroup; rp += kGroup; cp += kGroup;
        } while (j + kGroup <= cnt);
        if (ok) continue;
        rem = s_rem; span = s_span; w0 = s_w0; wa = s_wa;   // replay this group carefully
      }
      careful(min(j + kGroup, cnt));
    }
    // P_{r+1}: stream position of w0; the next window starts at P_r (see the kernel comment)
    const uint32_t p_next = win_first + (wa - lin_r) / 2u - 1u;
    if (lane == 0) ctrl[(r + 1) & 1] = p_next;
    pair_barrier(bar);
    const uint32_t lin_n = lin_s + (uint32_t)((r + 1) & 1) * (kLmax * 2);
    win_first = p_cur & ~7u;
    wa = lin_n + 2u * (p_next + 1u - win_first);
    p_cur = p_next;
  }
  if (lane == 0 && p.status) p.status[s] = d.corrupt ? TFC_STREAM_CORRUPT : TFC_STREAM_OK;
  if (lane == 0) {
    atomicAdd(&g_trace[2], global_ns() - trace_t0);
    atomicAdd(&g_trace[3], 1ull);
  }
}


// ------------------------------------------------------------------------------------------
// encode
// ------------------------------------------------------------------------------------------
// One CTA (3 warps) per stream, three pipeline stages one round apart:
//   warp 2  stager : symbols + indexes -> "puts" {c0, c1} at precision 16 (escape tails expanded,
//                    warp scan for the positions); nothing here depends on the coder state;
//   warp 0  span   : the only truly serial recurrence, span' = f(span, c0, c1): two multiplies,
//                    a subtract, a compare and a select per put; emits a = floor(span*c0/2^16)
//                    and the renormalisation flag;
//   warp 1  low    : lo += a with immediate carry propagation into the words already produced
//                    (the reference delays the carry, range_coder.cc:169-265; the bytes are the
//                    same big-number sum), one 16-bit word per flag; flushes coalesced.
// A round is as many 32-symbol groups as are CERTAIN to fit in the put buffer (a group is
// started while put_cap - fill >= 32 * max_puts), not a fixed number of symbols sized for the
// worst case: a symbol takes ~1.02 puts on the codec's streams and up to 11 in the worst case,
// and when streams are packed into a CTA the buffers shrink to 512 puts — fixed rounds were then
// 32 symbols long, one barrier of the three warps per 32 symbols (y encode of 64 streams:
// 21.7 ms at 8 per CTA against 13.5 at 2; see profiles/r2z_encoder_variable_rounds.txt).  The
// stager announces its last round in shared memory; the other two warps follow one and two
// rounds behind.
constexpr int kPutCap = 2048;          // puts per round (upper bound)

struct EncParamsV4 {
  const int32_t* data;
  const int32_t* index;
  int64_t index_stride;
  const int32_t* cdf;
  const int32_t* cdf_size;
  const int32_t* offset;
  int32_t rows, cols;
  int32_t precision, overflow_width;
  uint8_t* out;
  int64_t cap;
  int32_t* nbytes;
  int32_t* status;
  int64_t S, n;
  int32_t round;          // symbols per round: round * max_puts <= put_cap, multiple of 32
  int32_t table_bytes;
  int32_t put_cap;        // puts per round (<= kPutCap, multiple of 32)
  int32_t streams_per_cta;
  int32_t max_puts;       // puts one symbol can take at most (escape with all its digits)
  int32_t fixed_round;    // > 0 (experiments): rounds of exactly this many symbols, as before
};

__host__ __device__ inline size_t enc_bytes(int put_cap) {
  const size_t b = (size_t)2 * put_cap * 8      // puts
                   + (size_t)2 * put_cap * 8    // {a, renormalisation flag} per put
                   + (size_t)put_cap * 2 + 4    // words of one round + the pending one
                   + 64;                        // counts[4], span, ...
  return (b + 15) & ~size_t(15);
}

__device__ __forceinline__ void trio_barrier(int id) {
  asm volatile("bar.sync %0, 96;" ::"r"(id) : "memory");
}

__device__ __forceinline__ void put_byte(uint8_t* out, uint32_t cap, uint32_t at, uint32_t b) {
  if (at < cap) out[at] = (uint8_t)b;
}

// streams_per_cta streams share one CTA (and one staged copy of the tables); each stream's three
// warps synchronise on their own named barrier, so streams never wait for each other.
template <bool MULTI>
__global__ void __launch_bounds__(MULTI ? 768 : 96) encode_unbounded_v4(const EncParamsV4 p) {
  const unsigned long long trace_t0 = global_ns();
  Tables t;
  stage_tables_v4(p.cdf, p.rows, p.cols, p.cdf_size, p.offset, p.precision, smem_v4,
                  p.table_bytes, &t);
  const int slot = MULTI ? (int)(threadIdx.x / 96) : 0;
  const int64_t s = MULTI ? (int64_t)blockIdx.x * p.streams_per_cta + slot : (int64_t)blockIdx.x;
  if (s >= p.S) return;
  const int kCap = p.put_cap;
  uint8_t* mine = smem_v4 + p.table_bytes + (size_t)slot * enc_bytes(kCap);
  uint2* puts = reinterpret_cast<uint2*>(mine);                                  // [2][kCap]
  uint2* avals = reinterpret_cast<uint2*>(mine + 2 * kCap * 8);                  // [2][kCap]
  uint16_t* words = reinterpret_cast<uint16_t*>(avals + 2 * kCap);               // [kCap + 2]
  volatile uint32_t* counts = reinterpret_cast<volatile uint32_t*>(words + kCap + 2);  // [4]
  volatile uint32_t* fin = counts + 4;                                           // final span1
  volatile int32_t* last_round = reinterpret_cast<volatile int32_t*>(counts + 5);  // -1: not yet

  const int warp = (int)((threadIdx.x - slot * 96) >> 5), lane = threadIdx.x & 31;
  if (warp == 2 && lane == 0) last_round[0] = -1;     // read by the others after the first barrier
  const int32_t* idx = p.index + s * p.index_stride;
  const int32_t* data = p.data + s * p.n;
  uint8_t* out = p.out + s * p.cap;
  const uint32_t cap = (uint32_t)p.cap;
  const int ow = p.overflow_width;
  const uint32_t all1 = (1u << ow) - 1u;
  const int dsh = 16 - ow;

  uint32_t span1 = 0xFFFFFFFFu;          // warp 0
  uint32_t lo = 0, total = 0;            // warp 1: low word, words flushed so far
  uint32_t last = 0, have_last = 0;      // warp 1: the newest word, not flushed yet (words[0])
  // warp 2: symbols and indexes are fetched two 32-symbol groups ahead of their use
  int32_t pf_idx[2] = {0, 0}, pf_dat[2] = {0, 0};
  if (warp == 2) {
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      const int64_t at = (int64_t)k * 32 + lane;
      if (at < p.n) { pf_idx[k] = __ldg(idx + at); pf_dat[k] = __ldg(data + at); }
    }
  }

  int64_t staged = 0;                    // warp 2: symbols turned into puts so far
  const uint32_t group_puts = 32u * (uint32_t)p.max_puts;
  for (int64_t r = 0;; ++r) {
    const int64_t L = last_round[0];     // the stager's last round, -1 while it is not known
    if (warp == 2) {
      if (L < 0) {
        // ---- stager: 32-symbol groups for as long as the next one is certain to fit ----
        uint2* dst = puts + (r & 1) * kCap;
        uint32_t fill = 0;
        const int64_t stop = p.fixed_round > 0 ? staged + p.fixed_round : p.n;
        while (staged < p.n && staged < stop && fill + group_puts <= (uint32_t)kCap) {
          const int64_t j = staged + lane;
          uint32_t c0 = 0, c1 = 0, x = 0;
          int np = 0, nd = 0;
          const int32_t cur_idx = pf_idx[0], cur_dat = pf_dat[0];
          pf_idx[0] = pf_idx[1]; pf_dat[0] = pf_dat[1];
          {
            const int64_t at = j + 64;
            if (at < p.n) { pf_idx[1] = __ldg(idx + at); pf_dat[1] = __ldg(data + at); }
          }
          if (j < p.n) {
            const int4 ri = t.rowinfo[cur_idx];
            const int32_t top = ri.y - 2;
            int32_t v = cur_dat - ri.z;
            np = 1;
            bool esc = false;
            if (v < 0) { x = (uint32_t)(-2 * v - 1); v = top; esc = true; }
            else if (v >= top) { x = (uint32_t)(2 * (v - top)); v = top; esc = true; }
            c0 = t.at(ri.x + v);
            c1 = t.at(ri.x + v + 1);
            if (esc) {
              while (nd * ow < 32 && (x >> (nd * ow)) != 0) ++nd;
              np += (int)((uint32_t)nd / all1) + 1 + nd;
            }
          }
          int incl = np;
#pragma unroll
          for (int o = 1; o < 32; o <<= 1) {
            const int u = __shfl_up_sync(0xFFFFFFFFu, incl, o);
            if (lane >= o) incl += u;
          }
          uint32_t at = fill + (uint32_t)(incl - np);
          if (np > 0) {
            dst[at++] = make_uint2(c0, c1);
            if (np > 1) {
              uint32_t u = (uint32_t)nd;
              for (; u >= all1; u -= all1) dst[at++] = make_uint2(all1 << dsh, (all1 + 1u) << dsh);
              dst[at++] = make_uint2(u << dsh, (u + 1u) << dsh);
              for (int q = 0; q < nd; ++q) {
                const uint32_t dg = (x >> (q * ow)) & all1;
                dst[at++] = make_uint2(dg << dsh, (dg + 1u) << dsh);
              }
            }
          }
          fill += (uint32_t)__shfl_sync(0xFFFFFFFFu, incl, 31);
          staged += 32;
        }
        if (lane == 0) {
          counts[r & 3] = fill;
          if (staged >= p.n) last_round[0] = (int32_t)r;
        }
      }
    } else if (warp == 0) {
      if (r >= 1 && (L < 0 || r - 1 <= L)) {
        // ---- span chain (replicated in all lanes; lane 0 stores) ----
        const int64_t q = r - 1;
        const uint32_t cnt = counts[q & 3];
        const uint2* src = puts + (q & 1) * kCap;
        uint2* av = avals + (q & 1) * kCap;
        // puts are fetched two iterations ahead so that the shared-memory latency is off the
        // span recurrence (reads past the round's end land in the neighbouring buffers and
        // are never used)
        uint2 cur = src[0], nxt = src[1];
#pragma unroll 4
        for (uint32_t i = 0; i < cnt; ++i) {
          const uint2 nn = src[i + 2];
          const uint32_t a = scale16(span1, cur.x);
          const uint32_t b1 = scale16(span1, cur.y);       // b + 1 (mod 2^32)
          const uint32_t ns = b1 - 1u - a;
          const bool rn = (ns >> 16) == 0;
          span1 = rn ? ((ns << 16) | 0xFFFFu) : ns;
          av[i] = make_uint2(a, rn ? 1u : 0u);
          cur = nxt;
          nxt = nn;
        }
        if (q == L && lane == 0) fin[0] = span1;       // L is settled by the time q reaches it
      }
    } else {
      if (r >= 2 && (L < 0 || r - 2 <= L)) {
        // ---- low chain + emission (replicated; lane 0 does the rare global fix-ups) ----
        const int64_t q = r - 2;
        const uint32_t cnt = counts[q & 3];
        const uint2* av = avals + (q & 1) * kCap;
        // The newest word of the stream stays pending (register `last`, mirrored in
        // words[nw - 1]; words[0] at the start of a round) because a carry out of the 32-bit
        // window lands in it.  Groups of 32 puts (one flag word) run branch-free: add, carry
        // into `last`, predicated stores.  A carry that has to ripple further (last == 0xFFFF,
        // probability 2^-16) marks the group; it is rolled back and replayed carefully.
        uint32_t nw = have_last;           // words of this round, the pending one included
        for (uint32_t g = 0; g < cnt; g += 32) {
          const uint32_t m = min(32u, cnt - g);
          const uint32_t lo0 = lo, nw0 = nw, last0 = last;
          bool slow = false;
          uint2 a2 = av[g], a1 = av[g + 1];
#pragma unroll 4
          for (uint32_t k = 0; k < m; ++k) {
            const uint2 an = av[g + k + 2];
            const uint32_t a = a2.x;
            lo += a;
            const bool c = lo < a;
            last += c ? 1u : 0u;
            slow |= (last >> 16) != 0u;
            if (c && nw > 0u) words[nw - 1] = (uint16_t)last;   // (no carry before the first word)
            const bool emit = a2.y != 0u;
            const uint32_t w = lo >> 16;
            if (emit) words[nw] = (uint16_t)w;
            last = emit ? w : last;
            nw += emit ? 1u : 0u;
            lo = emit ? (lo << 16) : lo;
            a2 = a1;
            a1 = an;
          }
          if (!slow) continue;
          lo = lo0; nw = nw0;
          __syncwarp();
          if (nw0 > 0) words[nw0 - 1] = (uint16_t)last0;
          __syncwarp();
          for (uint32_t k = 0; k < m; ++k) {
            const uint32_t ak = av[g + k].x;
            const bool emit_k = av[g + k].y != 0u;
            lo += ak;
            if (lo < ak) {
              // carry out of the window: +1 into the words already produced
              int32_t kk = (int32_t)nw - 1;
              bool carry = true;
              __syncwarp();
              while (carry && kk >= 0) {
                const uint32_t wv = (uint32_t)words[kk] + 1u;
                __syncwarp();
                words[kk] = (uint16_t)wv;
                carry = (wv >> 16) != 0;
                --kk;
              }
              if (carry && lane == 0) {
                int64_t gw = (int64_t)total - 1;     // flushed words live in global memory
                while (gw >= 0) {
                  const uint32_t at = 2u * (uint32_t)gw;
                  uint32_t wv = 0;
                  if (at < cap) wv = (uint32_t)out[at] << 8;
                  if (at + 1 < cap) wv |= out[at + 1];
                  wv += 1u;
                  put_byte(out, cap, at, (wv >> 8) & 0xFFu);
                  put_byte(out, cap, at + 1, wv & 0xFFu);
                  if ((wv >> 16) == 0) break;
                  --gw;
                }
              }
              __syncwarp();
            }
            if (emit_k) {
              words[nw] = (uint16_t)(lo >> 16);
              ++nw;
              lo <<= 16;
            }
          }
          __syncwarp();
          last = nw > 0 ? (uint32_t)words[nw - 1] : 0u;
        }
        __syncwarp();
        // flush all but the newest word (everything after the last round)
        const uint32_t keep = ((L < 0 || q < L) && nw > 0) ? 1u : 0u;
        const uint32_t nflush = nw - keep;
        for (uint32_t k = lane; k < nflush; k += 32) {
          const uint32_t wv = words[k];
          put_byte(out, cap, 2u * (total + k), wv >> 8);
          put_byte(out, cap, 2u * (total + k) + 1u, wv & 0xFFu);
        }
        total += nflush;
        __syncwarp();
        if (keep) words[0] = (uint16_t)last;
        have_last = keep;
        __syncwarp();
      }
    }
    if (MULTI) trio_barrier(1 + slot); else __syncthreads();
    const int64_t L2 = last_round[0];
    if (L2 >= 0 && r >= L2 + 2) break;
  }

  if (warp == 1) {
    // Finalize (range_coder.cc:268-301)
    __threadfence_block();
    if (lane == 0) {
      const uint32_t fs = fin[0];
      uint32_t nbytes = 2u * total;
      auto rd = [&](uint32_t g) {
        const uint32_t at = 2u * g;
        uint32_t wv = 0;
        if (at < cap) wv = (uint32_t)out[at] << 8;
        if (at + 1 < cap) wv |= out[at + 1];
        return wv;
      };
      if (p.n > 0 && lo + fs < lo) {
        // state 1: the pending word is the last one that is not 0xFFFF; it gets the carry and
        // everything after it (zeros) is dropped
        int64_t g = (int64_t)total - 1;
        if (2u * total > cap) {
          // the stream did not fit: the words are not all there to look at; report the size
          // of the unfinalised stream (an upper bound) — the caller retries with more room
          g = -1;
        } else {
          while (g >= 0 && rd((uint32_t)g) == 0xFFFFu) --g;
        }
        if (g >= 0) {
          const uint32_t wv = rd((uint32_t)g) + 1u;
          put_byte(out, cap, 2u * (uint32_t)g, (wv >> 8) & 0xFFu);
          nbytes = 2u * (uint32_t)g + 1u;
          if (wv & 0xFFu) { put_byte(out, cap, nbytes, wv & 0xFFu); ++nbytes; }
        }
      } else if (p.n > 0 && lo != 0) {
        const uint32_t mid = ((lo - 1u) >> 16) + 1u;
        put_byte(out, cap, nbytes, (mid >> 8) & 0xFFu);
        ++nbytes;
        if (mid & 0xFFu) { put_byte(out, cap, nbytes, mid & 0xFFu); ++nbytes; }
      }
      p.nbytes[s] = (int32_t)nbytes;
      if (p.status) p.status[s] = (nbytes > cap || 2u * total > cap) ? TFC_STREAM_CAPACITY
                                                                    : TFC_STREAM_OK;
      atomicAdd(&g_trace[0], global_ns() - trace_t0);
      atomicAdd(&g_trace[1], 1ull);
    }
  }
}

}  // namespace v4
