// TEST INFRASTRUCTURE ONLY — the "port" checker: an independent CPU restatement of the
// reference's entropy-coding arithmetic.  It is written from the algorithm, not from the
// reference text, and is cross-checked in tests/ against (i) the golden bit strings of
// python/layers/entropy_models_test.py:179-199, :509-533, :544-574, :585-615 and (ii)
// oracle/_ref/libtfc_ref.so (the reference's own code) wherever that was built.
//
// Algorithm being restated (citations are /root/reference/tensorflow_compression/cc/kernels/):
//   encoder step      range_coder.cc:38-266     decoder step  range_coder.cc:309-377
//   encoder flush     range_coder.cc:268-301    op loops      see oracle_api.h
//   pmf -> cdf        range_coding_helper_kernels.cc:97-201 (depends on libstdc++ std::sort
//                     order among equal keys and on glibc log2; both are used here as-is)
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <vector>

#include "oracle_api.h"

extern "C" const char* orc_kind(void) { return "port"; }

namespace {

// ---- encoder ---------------------------------------------------------------------------
// The interval is [lo, lo + span] with span stored minus one, both 32 bit.  `held` is zero
// while the interval does not straddle 2^32; otherwise its low 16 bits hold the word whose
// final value (itself or itself-1) is not yet known and the bits above count, in steps of
// 0x20000, pairs of filler bytes that follow it.
struct Enc {
  uint32_t lo = 0, span1 = 0xFFFFFFFFu;
  uint64_t held = 0;
  std::vector<uint8_t>* sink;
  void word(uint32_t w) { sink->push_back(uint8_t(w >> 8)); sink->push_back(uint8_t(w)); }
  void put(uint32_t c0, uint32_t c1, int prec) {
    const uint64_t span = uint64_t(span1) + 1;
    const uint32_t a = uint32_t((span * c0) >> prec);
    const uint32_t b = uint32_t((span * c1) >> prec) - 1;
    lo += a;
    const bool wrapped = lo < a;
    span1 = b - a;
    if (uint32_t(lo + span1) < lo) {  // still straddling 2^32
      if ((span1 >> 16) == 0) {
        lo <<= 16;
        span1 = (span1 << 16) | 0xFFFFu;
        held += 0x20000;
      }
      return;
    }
    if (held != 0) {  // the straddle resolved: above (wrapped) or below 2^32
      if (!wrapped) --held;
      word(uint32_t(held));
      sink->insert(sink->end(), size_t(held >> 16), wrapped ? uint8_t(0) : uint8_t(0xFF));
      held = 0;
    }
    if ((span1 >> 16) == 0) {
      const uint32_t hi = lo >> 16;
      lo <<= 16;
      span1 = (span1 << 16) | 0xFFFFu;
      if (uint32_t(lo + span1) >= lo) word(hi); else held = hi + 1;
    }
  }
  void flush() {
    uint32_t w;
    if (held != 0) w = uint32_t(held) & 0xFFFFu;
    else if (lo != 0) w = ((lo - 1) >> 16) + 1;
    else return;
    sink->push_back(uint8_t(w >> 8));
    if (w & 0xFF) sink->push_back(uint8_t(w));
  }
};

// ---- decoder ---------------------------------------------------------------------------
struct Dec {
  uint32_t lo = 0, span1 = 0xFFFFFFFFu, code = 0;
  const uint8_t* p;
  const uint8_t* end;
  Dec(const uint8_t* b, int64_t n) : p(b), end(b + n) { pull(); pull(); }
  void pull() {
    for (int k = 0; k < 2; ++k) { code <<= 8; if (p != end) code |= *p++; }
  }
  // smallest k >= 1 with span*cdf[k] > target; returns k-1
  int32_t get(const int32_t* cdf, int64_t len, int prec) {
    const uint64_t span = uint64_t(span1) + 1;
    const uint64_t target = ((uint64_t(uint32_t(code - lo)) + 1) << prec) - 1;
    int64_t first = 1, count = len - 1;
    while (count > 0) {
      const int64_t step = count / 2;
      if (span * uint64_t(cdf[first + step]) <= target) { first += step + 1; count -= step + 1; }
      else count = step;
    }
    const uint32_t a = uint32_t((span * uint64_t(cdf[first - 1])) >> prec);
    const uint32_t b = uint32_t((span * uint64_t(cdf[first])) >> prec) - 1;
    lo += a;
    span1 = b - a;
    if ((span1 >> 16) == 0) { lo <<= 16; span1 = (span1 << 16) | 0xFFFFu; pull(); }
    return int32_t(first - 1);
  }
  // uniform alphabet of 2^w digits (cdf = 0,1,2,...,2^w)
  uint32_t digit(int w) {
    const uint64_t span = uint64_t(span1) + 1;
    const uint64_t target = ((uint64_t(uint32_t(code - lo)) + 1) << w) - 1;
    const uint64_t d = target / span;  // smallest k with span*k > target is d+1
    const uint32_t a = uint32_t((span * d) >> w);
    const uint32_t b = uint32_t((span * (d + 1)) >> w) - 1;
    lo += a;
    span1 = b - a;
    if ((span1 >> 16) == 0) { lo <<= 16; span1 = (span1 << 16) | 0xFFFFu; pull(); }
    return uint32_t(d);
  }
};

int64_t deliver(const std::vector<uint8_t>& s, uint8_t* out, int64_t cap) {
  if (int64_t(s.size()) > cap) return -1;
  if (!s.empty()) std::memcpy(out, s.data(), s.size());
  return int64_t(s.size());
}

}  // namespace

extern "C" int64_t orc_ui_encode(const int32_t* data, const int32_t* index, int64_t n,
                                 const int32_t* cdf, int32_t rows, int32_t cols,
                                 const int32_t* cdf_size, const int32_t* offset, int precision,
                                 int overflow_width, uint8_t* out, int64_t cap) {
  (void)rows;
  std::vector<uint8_t> sink;
  Enc e;
  e.sink = &sink;
  const int w = overflow_width;
  const uint32_t all1 = (1u << w) - 1;
  for (int64_t i = 0; i < n; ++i) {
    const int32_t r = index[i];
    const int32_t esc = cdf_size[r] - 2;  // escape symbol == number of direct values
    const int32_t* row = cdf + int64_t(r) * cols;
    const int32_t v = data[i] - offset[r];
    if (v >= 0 && v < esc) { e.put(row[v], row[v + 1], precision); continue; }
    // out of range: zig-zag the excess, then <digit count in unary-ish base 2^w - 1>,
    // then the digits least-significant first
    const uint32_t x = v < 0 ? uint32_t(-2 * v - 1) : uint32_t(2 * (v - esc));
    e.put(row[esc], row[esc + 1], precision);
    int nd = 0;
    while ((x >> (nd * w)) != 0) ++nd;
    uint32_t u = nd;
    for (; u >= all1; u -= all1) e.put(all1, all1 + 1, w);
    e.put(u, u + 1, w);
    for (int j = 0; j < nd; ++j) { const uint32_t d = (x >> (j * w)) & all1; e.put(d, d + 1, w); }
  }
  e.flush();
  return deliver(sink, out, cap);
}

extern "C" int orc_ui_decode(const uint8_t* enc, int64_t nbytes, const int32_t* index,
                             int64_t n, const int32_t* cdf, int32_t rows, int32_t cols,
                             const int32_t* cdf_size, const int32_t* offset, int precision,
                             int overflow_width, int32_t* out) {
  (void)rows;
  Dec d(enc, nbytes);
  const int w = overflow_width;
  const uint32_t all1 = (1u << w) - 1;
  for (int64_t i = 0; i < n; ++i) {
    const int32_t r = index[i];
    const int32_t esc = cdf_size[r] - 2;
    int32_t v = d.get(cdf + int64_t(r) * cols, esc + 2, precision);
    if (v == esc) {
      int nd = 0;
      uint32_t u;
      do { u = d.digit(w); nd += int(u); } while (u == all1);
      uint32_t x = 0;
      for (int j = 0; j < nd; ++j) x |= d.digit(w) << (j * w);
      v = (x & 1) ? -int32_t(x >> 1) - 1 : int32_t(x >> 1) + esc;
    }
    out[i] = v + offset[r];
  }
  return 0;
}

extern "C" int64_t orc_rc_encode(const int16_t* data, const int64_t* row_of, int64_t n,
                                 const int32_t* cdf, int32_t chip, int precision,
                                 uint8_t* out, int64_t cap) {
  std::vector<uint8_t> sink;
  Enc e;
  e.sink = &sink;
  for (int64_t i = 0; i < n; ++i) {
    const int32_t* row = cdf + (row_of ? row_of[i] : 0) * int64_t(chip);
    e.put(row[data[i]], row[data[i] + 1], precision);
  }
  e.flush();
  return deliver(sink, out, cap);
}

extern "C" int orc_rc_decode(const uint8_t* enc, int64_t nbytes, const int64_t* row_of,
                             int64_t n, const int32_t* cdf, int32_t chip, int precision,
                             int16_t* out) {
  Dec d(enc, nbytes);
  for (int64_t i = 0; i < n; ++i)
    out[i] = int16_t(d.get(cdf + (row_of ? row_of[i] : 0) * int64_t(chip), chip, precision));
  return 0;
}

#include "batch.inc"

// ---- pmf -> quantized cdf --------------------------------------------------------------
namespace {
struct Slot { int32_t* v; double mass; double key; };

double steal_cost(const Slot& s) {  // cost of lowering *v by one; never lower a 1
  if (*s.v <= 1) return std::numeric_limits<double>::infinity();
  return s.mass * (std::log2(double(*s.v)) - std::log2(double(*s.v - 1)));
}
double give_gain(const Slot& s) {  // gain of raising *v by one
  if (*s.v < 1) return -std::numeric_limits<double>::infinity();
  return s.mass * (std::log2(double(*s.v + 1)) - std::log2(double(*s.v)));
}

void quantize_row(const float* pmf, int32_t n, int precision, int32_t* c) {
  const int32_t total = 1 << precision;
  int32_t sum = 0;
  for (int32_t i = 0; i < n; ++i) {
    const float scaled = pmf[i] * float(total);
    c[i] = std::max(int32_t(std::rint(scaled)), 1);
    sum += c[i];
  }
  if (sum != total) {
    const bool over = sum > total;
    std::vector<Slot> q(n);
    for (int32_t i = 0; i < n; ++i) {
      q[i].v = c + i;
      q[i].mass = pmf[i];
      q[i].key = over ? steal_cost(q[i]) : give_gain(q[i]);
    }
    if (over) std::sort(q.begin(), q.end(), [](const Slot& a, const Slot& b) { return a.key < b.key; });
    else std::sort(q.begin(), q.end(), [](const Slot& a, const Slot& b) { return a.key > b.key; });
    for (int32_t todo = over ? sum - total : total - sum; todo > 0; --todo) {
      Slot head = q[0];
      *head.v += over ? -1 : 1;
      head.key = over ? steal_cost(head) : give_gain(head);
      // head goes back in front of the first entry that is strictly worse than it
      size_t j = 1;
      while (j < q.size() && !(over ? head.key < q[j].key : head.key > q[j].key)) {
        q[j - 1] = q[j];
        ++j;
      }
      q[j - 1] = head;
    }
  }
  for (int32_t i = 1; i < n; ++i) c[i] += c[i - 1];
}
}  // namespace

extern "C" void orc_pmf_to_cdf(const float* pmf, int64_t rows, int32_t n, int precision,
                               int32_t* cdf, int threads) {
  orc_parallel_for(rows, threads, [&](int64_t r) {
    int32_t* c = cdf + r * int64_t(n + 1);
    c[0] = 0;
    quantize_row(pmf + r * int64_t(n), n, precision, c + 1);
  });
}
