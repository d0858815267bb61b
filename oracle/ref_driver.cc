// TEST INFRASTRUCTURE ONLY — the "reference" checker.
//
// Thin C drivers around the reference's own arithmetic, compiled from the sources where
// they lie under /root/reference (see oracle/Makefile; nothing of the reference is copied
// into this repository):
//   * RangeEncoder / RangeDecoder  = cc/kernels/range_coder.{h,cc}, compiled unmodified
//     against oracle/shim/ (absl::Span, tf integer aliases, CHECK macros);
//   * PmfToCdfOp::PenaltyItem/GainItem/PerShard = range_coding_helper_kernels.cc:97-201,
//     which the Makefile extracts with sed into oracle/_ref/pershard.inc and which is
//     #included below inside a struct providing the `precision_` member it reads.
// The op loops around them are OpKernel methods (need TensorFlow headers to compile), so
// the loops of unbounded_index_range_coding_kernels.cc:203-250, :325-365 and
// range_coding_kernels.cc:247-269, :362-372 are re-driven here on raw pointers.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <functional>
#include <iterator>
#include <limits>
#include <numeric>
#include <string>
#include <vector>

#include "absl/types/span.h"
#include "tensorflow/core/platform/logging.h"
#include "tensorflow/core/platform/types.h"
#include "tensorflow_compression/cc/kernels/range_coder.h"

#include "oracle_api.h"

namespace {
using tensorflow::int32;
using tensorflow::int64;
using tensorflow::uint32;
struct RefPmfToCdf {
  int precision_;
#include "pershard.inc"
};
}  // namespace

using tensorflow_compression::RangeDecoder;
using tensorflow_compression::RangeEncoder;

extern "C" const char* orc_kind(void) { return "reference"; }

static int64_t emit(const std::string& s, uint8_t* out, int64_t cap) {
  if ((int64_t)s.size() > cap) return -1;
  std::memcpy(out, s.data(), s.size());
  return (int64_t)s.size();
}

extern "C" int64_t orc_ui_encode(const int32_t* data, const int32_t* index, int64_t n,
                                 const int32_t* cdf, int32_t rows, int32_t cols,
                                 const int32_t* cdf_size, const int32_t* offset, int precision,
                                 int overflow_width, uint8_t* out, int64_t cap) {
  (void)rows;
  RangeEncoder enc;
  std::string sink;
  const uint32_t top = (1u << overflow_width) - 1;
  for (int64_t i = 0; i < n; ++i) {
    const int32_t r = index[i];
    const int32_t esc = cdf_size[r] - 2;
    int32_t v = data[i] - offset[r];
    uint32_t extra = 0;
    if (v < 0) { extra = -2 * v - 1; v = esc; }
    else if (v >= esc) { extra = 2 * (v - esc); v = esc; }
    const int32_t* row = cdf + (int64_t)r * cols;
    enc.Encode(row[v], row[v + 1], precision, &sink);
    if (v == esc) {
      int32_t nd = 0;
      while ((extra >> (nd * overflow_width)) != 0) ++nd;
      uint32_t u = nd;
      while (u >= top) { enc.Encode(top, top + 1, overflow_width, &sink); u -= top; }
      enc.Encode(u, u + 1, overflow_width, &sink);
      for (int32_t j = 0; j < nd; ++j) {
        const uint32_t d = (extra >> (j * overflow_width)) & top;
        enc.Encode(d, d + 1, overflow_width, &sink);
      }
    }
  }
  enc.Finalize(&sink);
  return emit(sink, out, cap);
}

extern "C" int orc_ui_decode(const uint8_t* encp, int64_t nbytes, const int32_t* index,
                             int64_t n, const int32_t* cdf, int32_t rows, int32_t cols,
                             const int32_t* cdf_size, const int32_t* offset, int precision,
                             int overflow_width, int32_t* out) {
  (void)rows;
  const std::string src(reinterpret_cast<const char*>(encp), (size_t)nbytes);
  RangeDecoder dec(src);
  const uint32_t top = (1u << overflow_width) - 1;
  std::vector<int32_t> digits((1 << overflow_width) + 1);
  std::iota(digits.begin(), digits.end(), 0);
  for (int64_t i = 0; i < n; ++i) {
    const int32_t r = index[i];
    const int32_t esc = cdf_size[r] - 2;
    const int32_t* row = cdf + (int64_t)r * cols;
    int32_t v = dec.Decode(absl::Span<const int32_t>(row, esc + 2), precision);
    if (v == esc) {
      int32_t nd = 0;
      uint32_t u;
      do { u = dec.Decode(digits, overflow_width); nd += u; } while (u == top);
      uint32_t extra = 0;
      for (int32_t j = 0; j < nd; ++j) {
        const uint32_t d = dec.Decode(digits, overflow_width);
        extra |= d << (j * overflow_width);
      }
      v = extra >> 1;
      if (extra & 1) v = -v - 1; else v += esc;
    }
    out[i] = v + offset[r];
  }
  return 0;
}

extern "C" int64_t orc_rc_encode(const int16_t* data, const int64_t* row_of, int64_t n,
                                 const int32_t* cdf, int32_t chip, int precision,
                                 uint8_t* out, int64_t cap) {
  RangeEncoder enc;
  std::string sink;
  for (int64_t i = 0; i < n; ++i) {
    const int32_t* row = cdf + (row_of ? row_of[i] : 0) * (int64_t)chip;
    const int64_t v = data[i];
    enc.Encode(row[v], row[v + 1], precision, &sink);
  }
  enc.Finalize(&sink);
  return emit(sink, out, cap);
}

extern "C" int orc_rc_decode(const uint8_t* encp, int64_t nbytes, const int64_t* row_of,
                             int64_t n, const int32_t* cdf, int32_t chip, int precision,
                             int16_t* out) {
  const std::string src(reinterpret_cast<const char*>(encp), (size_t)nbytes);
  RangeDecoder dec(src);
  for (int64_t i = 0; i < n; ++i) {
    const int32_t* row = cdf + (row_of ? row_of[i] : 0) * (int64_t)chip;
    out[i] = (int16_t)dec.Decode(absl::Span<const int32_t>(row, (size_t)chip), precision);
  }
  return 0;
}

#include "batch.inc"

extern "C" void orc_pmf_to_cdf(const float* pmf, int64_t rows, int32_t n, int precision,
                               int32_t* cdf, int threads) {
  RefPmfToCdf op{precision};
  orc_parallel_for(rows, threads, [&](int64_t r) {
    int32_t* c = cdf + r * (int64_t)(n + 1);
    c[0] = 0;
    op.PerShard(absl::Span<const float>(pmf + r * (int64_t)n, (size_t)n),
                absl::Span<int32_t>(c + 1, (size_t)n));
  });
}
