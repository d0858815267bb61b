/* TEST INFRASTRUCTURE ONLY — never linked into or imported by the product path.
 *
 * C API shared by the two CPU checkers of the entropy-coding hot path:
 *   oracle/_ref/libtfc_ref.so   the reference's own range_coder.cc + PerShard, compiled from
 *                               /root/reference by oracle/Makefile ("reference")
 *   oracle/libtfc_port.so       an independent restatement (oracle/rc_port.cc, "port")
 * Both export exactly these symbols so tests can swap one for the other.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load them.
 */
#ifndef TFC_ORACLE_API_H_
#define TFC_ORACLE_API_H_
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* which implementation: "reference" or "port" */
const char* orc_kind(void);

/* UnboundedIndexRangeEncode op loop (unbounded_index_range_coding_kernels.cc:189-251).
 * Returns the number of bytes of the encoded string; -1 if it would exceed cap. */
int64_t orc_ui_encode(const int32_t* data, const int32_t* index, int64_t n,
                      const int32_t* cdf, int32_t rows, int32_t cols,
                      const int32_t* cdf_size, const int32_t* offset,
                      int precision, int overflow_width, uint8_t* out, int64_t cap);

/* UnboundedIndexRangeDecode op loop (unbounded_index_range_coding_kernels.cc:308-368). */
int orc_ui_decode(const uint8_t* enc, int64_t nbytes, const int32_t* index, int64_t n,
                  const int32_t* cdf, int32_t rows, int32_t cols,
                  const int32_t* cdf_size, const int32_t* offset,
                  int precision, int overflow_width, int32_t* out);

/* RangeEncode / RangeDecode op loops (range_coding_kernels.cc:234-271, 346-374).  The CDF
 * broadcast is resolved by the caller: element i uses row row_of[i] of cdf[rows, chip]
 * (row_of == NULL: every element uses row 0). */
int64_t orc_rc_encode(const int16_t* data, const int64_t* row_of, int64_t n,
                      const int32_t* cdf, int32_t chip, int precision,
                      uint8_t* out, int64_t cap);
int orc_rc_decode(const uint8_t* enc, int64_t nbytes, const int64_t* row_of, int64_t n,
                  const int32_t* cdf, int32_t chip, int precision, int16_t* out);

/* PmfToQuantizedCdf (range_coding_helper_kernels.cc:62-94, 152-201): pmf[rows, n] ->
 * cdf[rows, n+1].  threads<=1: serial. */
void orc_pmf_to_cdf(const float* pmf, int64_t rows, int32_t n, int precision,
                    int32_t* cdf, int threads);

/* Batched forms used for the timed CPU baseline: S independent streams of n symbols each,
 * spread over `threads` host threads (one stream is always coded by one thread, as in the
 * reference where tf.map_fn runs one op per image).  index_stride = n (per-stream indexes)
 * or 0 (indexes shared by all streams).  out is [S, cap]; nbytes[S]. */
void orc_ui_encode_batch(int64_t S, const int32_t* data, const int32_t* index,
                         int64_t index_stride, int64_t n, const int32_t* cdf, int32_t rows,
                         int32_t cols, const int32_t* cdf_size, const int32_t* offset,
                         int precision, int overflow_width, uint8_t* out, int64_t cap,
                         int64_t* nbytes, int threads);
void orc_ui_decode_batch(int64_t S, const uint8_t* enc, int64_t cap, const int64_t* nbytes,
                         const int32_t* index, int64_t index_stride, int64_t n,
                         const int32_t* cdf, int32_t rows, int32_t cols,
                         const int32_t* cdf_size, const int32_t* offset, int precision,
                         int overflow_width, int32_t* out, int threads);
#ifdef __cplusplus
}
#endif
#endif
