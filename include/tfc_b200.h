/* tfc_b200.h — C ABI of libtfc_b200.so, the B200 (sm_100a) implementation of the
 * tensorflow_compression learned-image-codec hot path.
 *
 * This header is the drop-in boundary.  The reference reaches its native code
 * through a TF custom-op library (`load_op_library("libtensorflow_compression.so")`,
 * python/ops/range_coding_ops.py:26-28); this library exports the same five ops — batched, on
 * device buffers — as plain `extern "C"` functions, plus the layer-level kernels the reference
 * gets from stock TF ops (conv2d, conv2d_backprop_input, GDN's 1x1 conv, erfc, ...).
 * Citations are relative to /root/reference/tensorflow_compression/.
 *
 * Conventions
 *   - every pointer named d_* is a DEVICE pointer; h_* is a HOST pointer;
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream); all work is
 *     stream-ordered and the functions are re-entrant (no global mutable state);
 *   - the return value is a tfc_status; tfc_last_error() gives the message of the calling
 *     thread's last failure, using the reference's own wording where the reference has one
 *     (cc/kernels/range_coding_kernels.cc:137-176, unbounded_index_range_coding_kernels.cc:
 *     58-139, range_coding_kernels_util.cc:58-60);
 *   - the library never allocates device memory behind the caller's back; scratch space is
 *     passed in and sized by the matching *_workspace() function;
 *   - there is no CPU fallback: without a CUDA device every compute entry point returns
 *     TFC_ERR_CUDA.
 */
#ifndef TFC_B200_H_
#define TFC_B200_H_
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
  TFC_OK = 0,
  TFC_ERR_INVALID_ARGUMENT = 1, /* == tf.errors.InvalidArgumentError in the reference */
  TFC_ERR_CUDA = 2,             /* CUDA runtime failure (message has cudaGetErrorString) */
  TFC_ERR_CAPACITY = 3,         /* an encode output slab was too small; nbytes[] holds need */
  TFC_ERR_DECODE = 4,           /* corrupt stream (reference: CHECK-abort, range_coder.cc:347) */
  TFC_ERR_UNIMPLEMENTED = 5     /* == NotImplementedError (signal_conv.py:393-401) */
} tfc_status;

const char* tfc_last_error(void);
const char* tfc_version(void);
/* number of kernels this library has launched since load (bench.py's gpu_launches) */
int64_t tfc_launch_count(void);

/* Process-wide scheduling knobs (no effect on results).  The range coder is latency bound (one
 * serial chain per image) and leaves most of each SM idle; a caller that overlaps it with the
 * transforms of another batch on a second CUDA stream packs several coder streams into one CTA
 * and keeps the persistent transform kernels off the SMs the coder occupies.
 *   TFC_CFG_SM_BUDGET              SMs (= CTAs) a persistent transform kernel may use; 0 = all
 *   TFC_CFG_CODER_STREAMS_PER_CTA  1..8 coder streams per CTA; 0 = automatic (1 while S <= #SMs)
 *                                  values of 32..512 additionally pack value / 32 WARPS of the
 *                                  lane-per-stream kernels (32 streams each) into one CTA; the
 *                                  warps-per-stream kernels use min(value, 8)
 *   TFC_CFG_TILES_PER_CTA          0 = persistent transform kernels (one CTA per SM walks the
 *                                  tile list); k > 0 = k consecutive tiles per CTA and as many
 *                                  CTAs as that takes, so the hardware scheduler fills whatever
 *                                  SMs the coder leaves free
 *   TFC_CFG_CONV_CLUSTER           thread-block cluster size of the implicit-GEMM kernel (its CTAs
 *                                  share each weight tile by TMA multicast): 0 = automatic (currently
 *                                  off: no gain measured), 1 = off, 2, 4 (where the shape allows)
 *   TFC_CFG_CODER_LANES            unbounded-index coder with one stream per LANE (many short
 *                                  streams) instead of 2-3 warps per stream (few long ones):
 *                                  0 = automatic (unbounded-index: from 2048 streams; bounded: from 8192), 1 = never,
 *                                  2 = always */
#define TFC_CFG_SM_BUDGET 1
#define TFC_CFG_CODER_STREAMS_PER_CTA 2
#define TFC_CFG_TILES_PER_CTA 3
#define TFC_CFG_CONV_CLUSTER 4
#define TFC_CFG_CODER_LANES 5
/*   TFC_CFG_CONV_STAGES            stages of the implicit-GEMM kernel's TMA operand ring: 0 = automatic
 *                                  (4), 2, 3, 4.  Three stages take 172 KB instead of 216 KB of shared
 *                                  memory per CTA, which leaves room on the SM for a one-stream
 *                                  range-coder CTA of another batch (co-residency)
 *   TFC_CFG_CODER_TABLE_KB         shared-memory budget (KB) of the range coder's staged CDF tables:
 *                                  0 = automatic (72); tables that do not fit are read through L1
 *                                  from the caller's dense table (rows of 5 entries — the fast path —
 *                                  live in the 32-byte row records, which are always staged)
 *   TFC_CFG_CODER_SMEM_KB          dynamic shared memory (KB) a warps-per-stream coder CTA may take:
 *                                  0 = automatic (226), else 16..226; the per-stream round buffers
 *                                  shrink (shorter rounds) until the CTA fits
 *   TFC_CFG_CONV_PAIR              implicit-GEMM kernel on CTA pairs: one tcgen05.mma.cta_group::2 of
 *                                  M = 256 covers two adjacent 128-pixel tiles, the two CTAs of the
 *                                  cluster each hold half of the weight tile (half the shared-memory
 *                                  operand reads per FLOP).  0 = automatic, 1 = off, 2 = on where the
 *                                  shape allows (even tile count per phase, filters a multiple of 32) */
#define TFC_CFG_CONV_STAGES 6
#define TFC_CFG_CODER_TABLE_KB 7
#define TFC_CFG_CODER_SMEM_KB 8
#define TFC_CFG_CONV_PAIR 9
int tfc_configure(int key, int value);

/* Debug aid (no reference counterpart): phase stamps of the thin-input GDN layer's kernel
 * (csrc/conv_thin.cu, TFC_A0_STAMPS): clock64() of CTA 0's phase boundaries over its tiles 8..15 of
 * the most recent launch, 8 x 64 values (tools/a0_trace.py names the slots).  Copies min(n, 512)
 * values to host memory; returns a cudaError_t.  Call after synchronising the launch's stream. */
int tfc_debug_a0_trace(unsigned long long* h_dst, int n);

/* Debug aid (no reference counterpart): how long the warps-per-stream coder kernels held their
 * SMs.  h_out4 = {encode ns, encode streams, decode ns, decode streams}, each stream's time from
 * its CTA's first instruction to its last, summed since the last reset.  Synchronises the
 * device.  tools/pipe_gantt.py uses it to tell "waiting for an SM" from "running on it". */
int tfc_debug_coder_trace(uint64_t* h_out4, int reset);

/* ------------------------------------------------------------------------------------------
 * Entropy coding — replaces the five custom ops of cc/ops/range_coding_ops.cc
 * ---------------------------------------------------------------------------------------- */

/* Per-stream status bits written to d_status[s]. */
#define TFC_STREAM_OK 0
#define TFC_STREAM_CAPACITY 1 /* output did not fit in cap; d_nbytes[s] is the needed size */
#define TFC_STREAM_CORRUPT 2  /* decoder search ran past the end of a CDF */

/* UnboundedIndexRangeEncode (cc/ops/range_coding_ops.cc:125-135; kernel
 * cc/kernels/unbounded_index_range_coding_kernels.cc:160-251), batched: S independent streams
 * of n symbols, one encoded string per stream — exactly what EntropyModel.compress builds with
 * tf.map_fn over the batch (python/layers/entropy_models.py:234-244).
 *   d_data   [S, n] int32      d_index [S, n] (index_stride = n) or [n] shared (index_stride = 0)
 *   d_cdf    [rows, cols]      d_cdf_size, d_offset [rows]
 *   d_out    [S, cap] bytes    d_nbytes [S] int32 length of each string
 *   d_status [S] int32 (may be NULL)
 * debug_level 1 runs the reference's CheckArgumentValues (:107-115) on the device and
 * synchronises the stream to report it; 0 (what the layers pass, entropy_models.py:239) does
 * not synchronise. */
int tfc_unbounded_index_range_encode(const int32_t* d_data, const int32_t* d_index,
                                     int64_t index_stride, int64_t S, int64_t n,
                                     const int32_t* d_cdf, int32_t rows, int32_t cols,
                                     const int32_t* d_cdf_size, const int32_t* d_offset,
                                     int precision, int overflow_width, int debug_level,
                                     uint8_t* d_out, int64_t cap, int32_t* d_nbytes,
                                     int32_t* d_status, void* stream);

/* UnboundedIndexRangeDecode (cc/ops/range_coding_ops.cc:203-216; kernel :280-368).
 *   d_enc [S, cap] bytes, d_nbytes [S]; reads past the end of a string yield zeros
 *   (range_coder.cc:368-377).  d_out [S, n] int32. */
int tfc_unbounded_index_range_decode(const uint8_t* d_enc, int64_t cap, const int32_t* d_nbytes,
                                     const int32_t* d_index, int64_t index_stride, int64_t S,
                                     int64_t n, const int32_t* d_cdf, int32_t rows, int32_t cols,
                                     const int32_t* d_cdf_size, const int32_t* d_offset,
                                     int precision, int overflow_width, int debug_level,
                                     int32_t* d_out, int32_t* d_status, void* stream);

/* RangeEncode / RangeDecode (cc/ops/range_coding_ops.cc:29-35, :91-103; kernels
 * cc/kernels/range_coding_kernels.cc:178-380), batched over S streams that share one
 * data shape and one cdf shape.  `data_shape[rank]` and `cdf_shape[rank+1]` are the
 * per-stream tensor shapes as the op sees them; the broadcast rules, axis merging and error
 * texts are those of CheckCdfShape (:137-151) and MergeAxes (range_coding_kernels_util.cc:
 * 36-92).  cdf_stream_stride = number of int32 between the cdf of consecutive streams
 * (0: all streams share one cdf tensor).
 *   d_data [S, prod(data_shape)] int16 ; d_out/d_nbytes/d_status as above. */
int tfc_range_encode(const int16_t* d_data, int64_t S, const int64_t* data_shape, int rank,
                     const int32_t* d_cdf, const int64_t* cdf_shape, int cdf_rank,
                     int64_t cdf_stream_stride, int precision, int debug_level, uint8_t* d_out,
                     int64_t cap, int32_t* d_nbytes, int32_t* d_status, void* stream);
int tfc_range_decode(const uint8_t* d_enc, int64_t cap, const int32_t* d_nbytes, int64_t S,
                     const int64_t* data_shape, int rank, const int32_t* d_cdf,
                     const int64_t* cdf_shape, int cdf_rank, int64_t cdf_stream_stride,
                     int precision, int debug_level, int16_t* d_out, int32_t* d_status,
                     void* stream);

/* PmfToQuantizedCdf (cc/ops/range_coding_ops.cc:244-257; kernel
 * cc/kernels/range_coding_helper_kernels.cc:53-204): d_pmf [rows, n] float32 ->
 * d_cdf [rows, n+1] int32, one device thread per row replaying libstdc++'s std::sort
 * decision sequence so that ties resolve exactly as in the reference build.
 * Workspace: tfc_pmf_to_quantized_cdf_workspace(rows, n) bytes of device memory. */
size_t tfc_pmf_to_quantized_cdf_workspace(int64_t rows, int32_t n);
int tfc_pmf_to_quantized_cdf(const float* d_pmf, int64_t rows, int32_t n, int precision,
                             int32_t* d_cdf, void* d_workspace, size_t workspace_bytes,
                             void* stream);

/* The same op over rows of DIFFERENT lengths in one launch (the table rebuild of
 * python/layers/entropy_models.py:131-146 builds row r from pmf[r, :len_r] ++ [tail_r] and
 * zero-pads to the longest row): row r reads d_n[r] entries (clamped to [2, max_n]) at
 * d_pmf + r * pmf_stride and writes d_n[r] + 1 entries at d_cdf + r * cdf_stride, zeros behind
 * them up to cdf_stride.  The lengths stay on the device.  One warp per row in shared memory:
 * max_n up to ~7000. */
int tfc_pmf_to_quantized_cdf_ragged(const float* d_pmf, int64_t rows, int32_t pmf_stride,
                                    const int32_t* d_n, int32_t max_n, int precision,
                                    int32_t* d_cdf, int32_t cdf_stride, void* stream);

/* The same op for callers that cannot allocate device memory (code under CUDA graph capture,
 * tight serving loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * It takes no workspace: the scratch of the
 * thread-per-row kernel lives in a file-scope `static __device__` array of
 * tfc_pmf_to_quantized_cdf_static_capacity() = 8 MiB that every GPU loading this library holds
 * until exit.  Limit: round_up(rows, 32) * n <= 524 288 sort items of 16 bytes (e.g. 8192 rows x
 * 64 entries, 352 rows x 1480); anything larger returns TFC_ERR_INVALID_ARGUMENT — use
 * tfc_pmf_to_quantized_cdf.  Calls no cudaMalloc / cudaFree / allocating library routine; the
 * first call on a device looks up the array's device address (cudaGetSymbolAddress) and uploads
 * the log2 table.  All calls share the one array (nothing in it needs to start at zero: every
 * item is written before it is read), so calls that OVERLAP on the GPU race: issue them on one
 * stream at a time.  Results are identical to tfc_pmf_to_quantized_cdf. */
size_t tfc_pmf_to_quantized_cdf_static_capacity(void);
int tfc_pmf_to_quantized_cdf_static(const float* d_pmf, int64_t rows, int32_t n, int precision,
                                    int32_t* d_cdf, void* stream);

/* Host-side helper shared by the bounded ops: the reference's MergeAxes
 * (range_coding_kernels_util.cc:36-92).  Writes the merged shapes (at most rank entries, and
 * one more for the cdf) and returns their count through *merged_rank. */
int tfc_merge_axes(const int64_t* data_shape, int rank, const int64_t* cdf_shape, int cdf_rank,
                   int64_t* merged_data_shape, int64_t* merged_cdf_shape, int* merged_rank);


/* ------------------------------------------------------------------------------------------
 * Entropy models — the elementwise half of python/layers/entropy_models.py.  All tensors are
 * channels_last: `count` elements with the channel index fastest (channel = i % channels).
 * ---------------------------------------------------------------------------------------- */

/* Per-channel density-network parameters of EntropyBottleneck (entropy_models.py:461-489),
 * already reparameterised (softplus(matrix), tanh(factor)), packed per channel as
 *   layer 0..L: matrix[f_{i+1}][f_i], bias[f_{i+1}], then (i < L) factor[f_{i+1}]
 * with f = (1,) + filters + (1,) and L = len(filters).  `filters` must be uniform (all layers
 * the same width <= 4, at most 4 layers; (3,3,3) is the default and what both models use). */

/* EntropyBottleneck call(training=False) (entropy_models.py:148-187 with :609-614, :626-658):
 *   d_values[i] = floor(x + (0.5 - median_c)) + median_c
 *   d_lik[i]    = max(|sigmoid(s*upper) - sigmoid(s*lower)|, likelihood_bound)
 * mode 0: quantise as above; mode 1: x is used as given (training=True path after the caller
 * has added its uniform noise, :605-607).  d_values may be NULL in mode 1. */
int tfc_eb_call(const float* d_x, int64_t count, int32_t channels, const float* d_params,
                int32_t num_filters, int32_t filter_width, const float* d_medians,
                float likelihood_bound, int mode, float* d_values, float* d_lik, void* stream);

/* Gradient of the likelihood above in mode 1 (the training graph, examples/bls2017.py:156-171;
 * TF derives it through entropy_models.py:626-658 and the lower_bound rule of
 * ops/math_ops.py:61-74): d_grad_x[i] = d_grad_lik[i] * d lik / d x, d_grad_params [channels,
 * stride] = sum over the elements of each channel of d_grad_lik * d lik / d params, params in
 * the packed effective layout of tfc_eb_call.  num_filters >= 1. */
int tfc_eb_likelihood_backward(const float* d_x, int64_t count, int32_t channels,
                               const float* d_params, int32_t num_filters, int32_t filter_width,
                               float likelihood_bound, const float* d_grad_lik, float* d_grad_x,
                               float* d_grad_params, void* stream);

/* _quantize(mode="symbols") (entropy_models.py:609-618): int32(floor(x + (0.5 - median_c))) */
int tfc_eb_symbols(const float* d_x, int64_t count, int32_t channels, const float* d_medians,
                   int32_t* d_symbols, void* stream);
/* _dequantize (entropy_models.py:620-624): float(symbol) + median_c.
 * out_bf16 != 0 writes bfloat16 (the form the hyper-synthesis kernels consume). */
int tfc_eb_dequantize(const int32_t* d_symbols, int64_t count, int32_t channels,
                      const float* d_medians, void* d_out, int out_bf16, void* stream);

/* Logits of the cumulative (entropy_models.py:402-436) at arbitrary points, used on the host
 * side to build the PMF tables (:540-559): d_x [channels, samples] -> d_logits same shape. */
int tfc_eb_logits_cumulative(const float* d_x, int32_t channels, int32_t samples,
                             const float* d_params, int32_t num_filters, int32_t filter_width,
                             float* d_logits, void* stream);

/* GaussianConditional call(training=False), mean=None (entropy_models.py:887-932, :965-969):
 *   d_values = floor(y + 0.5);  s = max(sigma, scale_bound)
 *   d_lik    = max(Phi((.5-|v|)/s) - Phi((-.5-|v|)/s), likelihood_bound), Phi(t)=.5*erfc(-t/sqrt2)
 * mode 1: y used as given (noise path). */
int tfc_gaussian_call(const float* d_y, const float* d_sigma, int64_t count, float scale_bound,
                      float likelihood_bound, int mode, float* d_values, float* d_lik,
                      void* stream);

/* Gradient of the likelihood above in mode 1 (the training graph of examples/bmshj2018.py:
 * 201-282; TF derives it through entropy_models.py:916-932, :965-969 and the two lower_bound
 * ops, ops/math_ops.py:61-74, on the scale and on the likelihood):
 *   d_grad_y[i]     = g * d lik / d y        d_grad_sigma[i] = g * d lik / d sigma
 * with g = d_grad_lik[i] passed where lik >= likelihood_bound or g < 0, and the scale gradient
 * passed where sigma >= scale_bound or it is negative.  A bound <= 0 is disabled. */
int tfc_gaussian_likelihood_backward(const float* d_y, const float* d_sigma,
                                     const float* d_grad_lik, int64_t count, float scale_bound,
                                     float likelihood_bound, float* d_grad_y,
                                     float* d_grad_sigma, void* stream);

/* Compress-side fusion (entropy_models.py:221 + :869-881 + :895-907):
 *   d_symbols = int32(floor(y + 0.5))              (d_y / d_symbols may be NULL)
 *   d_indexes = (levels-1) - #{k < levels-1 : max(sigma, scale_bound) <= table[k]}
 * d_table: `levels` ascending float32 scales on the device. */
int tfc_gaussian_symbols_indexes(const float* d_y, const float* d_sigma, int64_t count,
                                 const float* d_table, int32_t levels, float scale_bound,
                                 int32_t* d_symbols, int32_t* d_indexes, void* stream);

/* _dequantize for mean=None: float(symbol); out_bf16 as above. */
int tfc_symbols_to_float(const int32_t* d_symbols, int64_t count, void* d_out, int out_bf16,
                         void* stream);

/* abs() + cast feeding the hyper-analysis transform (examples/bmshj2018.py:304). */
int tfc_abs_to_bf16(const float* d_x, int64_t count, void* d_out_bf16, void* stream);


/* ------------------------------------------------------------------------------------------
 * Transforms — SignalConv2D and GDN (python/layers/signal_conv.py, python/layers/gdn.py)
 * ---------------------------------------------------------------------------------------- */

/* TFC_U8: 8-bit images at the two ends of the codec.  As an INPUT dtype (thin-input kernel,
 * cin == 3) a byte b stands for b / 255, the scaling of read_png (examples/bmshj2018.py:53-57);
 * as an OUTPUT dtype (thin-output kernel, cout <= 4) the result is quantize_image
 * (examples/bmshj2018.py:60-63): saturate_cast<uint8>(round(x * 255)), round half to even.  Other
 * kernels return TFC_ERR_UNIMPLEMENTED for it. */
typedef enum { TFC_F32 = 0, TFC_BF16 = 1, TFC_U8 = 2 } tfc_dtype;
typedef enum {
  TFC_ACT_NONE = 0,
  TFC_ACT_RELU = 1, /* tf.nn.relu (examples/bmshj2018.py:153,157,183,187) */
  TFC_ACT_GDN = 2,  /* y = x * rsqrt(beta + gamma^T x^2)   (gdn.py:171-202) */
  TFC_ACT_IGDN = 3  /* y = x * sqrt (beta + gamma^T x^2)   (gdn.py inverse=True) */
} tfc_activation;

/* One 2-D signal convolution in its normalised form, channels_last:
 *   out[n,i,j,o] = act( bias[o] + sum_{a,b,c} U[n, i*sd + a - pad_top, j*sd + b - pad_left, c]
 *                                              * w[a,b,c,o] )
 * where U is `in` zero-stuffed by su (U[su*i', su*j'] = in[i',j'], zero elsewhere and outside).
 * The host side (python/layers/signal_conv.py mirror) reduces every supported combination of
 * corr / strides_down / strides_up / same_zeros to this form: a true convolution is a
 * correlation with the flipped kernel, and tf.nn.conv2d_backprop_input with explicit padding p
 * (signal_conv.py:629-695) is the zero-stuffed correlation with pad = k-1-p.
 *   d_w     [kh, kw, cin, cout] float32 (effective kernel, after the RDFT reparameterisation)
 *   d_bias  [cout] float32 or NULL
 *   d_beta  [cout], d_gamma [cout(k), cout(c)] float32 (effective, after the nonnegative
 *           reparameterisation) — used when act is GDN/IGDN.
 * in/out dtypes are tfc_dtype.  `impl`: 0 = pick the fastest kernel that supports the shape,
 * 1 = force the CUDA-core tiled kernel, 2 = force the tcgen05 implicit-GEMM kernel, 3 = force
 * the tcgen05 scatter-GEMM kernel for thin-output transposed layers, 4 = force the tcgen05
 * kernel with a software im2col for thin-input layers (TFC_ERR_UNIMPLEMENTED if the shape is
 * not supported by the forced kernel).
 * d_workspace: tfc_signal_conv2d_workspace(...) bytes (weights repacked to bf16 etc.). */
typedef struct {
  int32_t n, h, w, cin;     /* input  [n, h, w, cin]  */
  int32_t ho, wo, cout;     /* output [n, ho, wo, cout] */
  int32_t kh, kw;
  int32_t sd, su;           /* strides_down, strides_up (same in both dimensions) */
  int32_t pad_top, pad_left;
  int32_t act;              /* tfc_activation */
  int32_t in_dtype, out_dtype;
  int32_t impl;
  int32_t flags;            /* TFC_CONV_* bits */
} tfc_conv_desc;

/* The workspace already holds the repacked weights (and gamma) written by an earlier call with
 * the same layer parameters and the same workspace: skip the repack kernels.  The weight area
 * sits at the start of the workspace and does not depend on n/h/w. */
#define TFC_CONV_WEIGHTS_STAGED 1

/* Which kernel tfc_signal_conv2d will run for this descriptor (it depends on in_dtype and on
 * h % sd / w % sd, not only on the layer): 0 CUDA-core tiled kernel, 1 tcgen05 implicit GEMM,
 * 2 space-to-depth + tcgen05, 3 scatter GEMM (thin output), 4 software im2col + tcgen05 (thin
 * input), 5 zero-extended copy + tcgen05 implicit GEMM, 6 sub-pixel correlation + depth-to-space (thin
 * output, transposed, more than 128 scatter columns); -1 for an invalid descriptor.  Each path
 * keeps its own weight layout at the start of the workspace, and path 0 uses the workspace as
 * plain scratch: TFC_CONV_WEIGHTS_STAGED is only valid when the previous call on the same
 * workspace took the same non-zero path with the same weights. */
int tfc_signal_conv2d_path(const tfc_conv_desc* desc);

size_t tfc_signal_conv2d_workspace(const tfc_conv_desc* desc);
int tfc_signal_conv2d(const tfc_conv_desc* desc, const void* d_in, const float* d_w,
                      const float* d_bias, const float* d_beta, const float* d_gamma,
                      void* d_out, void* d_workspace, size_t workspace_bytes, void* stream);

/* Channel-separable form of the same layer (signal_conv.py:466-485, :590-596; the layout of
 * tf.nn.depthwise_conv2d_native): desc->cin channels, desc->cout = cin * multiplier outputs,
 * output channel c * multiplier + q = input channel c with its filter q,
 *   d_w [kh, kw, cin, multiplier], d_bias [cout] or NULL, float32 in and out, no fused activation.
 * Not on the codecs' hot path: a CUDA-core kernel. */
int tfc_depthwise_conv2d(const tfc_conv_desc* desc, const float* d_in, const float* d_w,
                         const float* d_bias, float* d_out, void* stream);

/* Filter gradient of the same convolution (training step of examples/bls2017.py:156-185; in the
 * reference TF derives it from tf.nn.conv2d / conv2d_backprop_input):
 *   d_dw[a,b,c,o] = sum_{n,i,j} U[n, i*sd + a - pad_top, j*sd + b - pad_left, c] * dout[n,i,j,o]
 * d_x_bf16 [n,h,w,cin] and d_dy_bf16 [n,ho,wo,cout] are bfloat16, d_dw [kh,kw,cin,cout] float32
 * (overwritten).  The input gradient needs no entry point of its own: it is tfc_signal_conv2d
 * with sd and su swapped, the kernel flipped and transposed, and padding k-1-pad. */
int tfc_signal_conv2d_wgrad(const tfc_conv_desc* desc, const void* d_x_bf16,
                            const void* d_dy_bf16, float* d_dw, void* stream);

/* One Adam update over a flat float32 bucket with tf.train.AdamOptimizer's arithmetic
 * (bls2017.py:173-178): lr_t = lr*sqrt(1-beta2^t)/(1-beta1^t); p -= lr_t*m/(sqrt(v)+eps);
 * the gradient is multiplied by grad_scale first (1/world_size after a sum all-reduce). */
int tfc_adam_step(float* d_param, const float* d_grad, float* d_m, float* d_v, int64_t n,
                  float lr, float beta1, float beta2, float eps, int64_t step, float grad_scale,
                  void* stream);

/* Bias gradient of a layer (the reduction TF derives for tf.nn.bias_add, signal_conv.py:785-804):
 * d_out[c] = sum over pixels of d_x[pixel, c], float32. */
int tfc_column_sums(const float* d_x, int64_t pixels, int32_t channels, float* d_out,
                    void* stream);

/* Elementwise halves of GDN / IGDN in the training graph (gdn.py:171-202 and its gradient); the
 * C x C contraction x^2 . gamma and its two gradients are GEMMs issued by the caller.  float32,
 * channels_last [pixels, channels], channels a multiple of 4, buffers 16-byte aligned.
 *   forward:        d_norm = d_x2_gamma + beta ;  d_y = x * rsqrt(norm)      (inverse: x * sqrt(norm))
 *   backward_pre:   d_gx1 = gy * rsqrt(norm) ; d_gnorm = -gy * x * norm^-3/2 / 2
 *                   (inverse: gy * sqrt(norm) ; gy * x / sqrt(norm) / 2)
 *   backward_post:  d_gx += 2 * x * d_gx2          with d_gx2 = gnorm . gamma^T */
int tfc_gdn_train_forward(const float* d_x, const float* d_x2_gamma, const float* d_beta,
                          int64_t pixels, int32_t channels, int inverse, float* d_y,
                          float* d_norm, void* stream);
int tfc_gdn_train_backward_pre(const float* d_gy, const float* d_x, const float* d_norm,
                               int64_t pixels, int32_t channels, int inverse, float* d_gx1,
                               float* d_gnorm, void* stream);
int tfc_gdn_train_backward_post(float* d_gx, const float* d_x, const float* d_gx2, int64_t pixels,
                                int32_t channels, void* stream);
/* The same passes for a caller that runs the contraction and its two gradients as tcgen05 1x1
 * convolutions (tfc_signal_conv2d / tfc_signal_conv2d_wgrad on bfloat16 operands, the arithmetic
 * of the inference epilogue) instead of float32 library GEMMs:
 *   square:        d_x2_bf16 = bfloat16(x * x)
 *   forward_bf16:  as forward, y written as bfloat16 (what the next layer's convolution reads)
 *   backward_pre2: as backward_pre, gy float32 or bfloat16 (gy_dtype: tfc_dtype), plus a
 *                  bfloat16 copy of gnorm in d_gnorm_bf16 */
int tfc_gdn_train_square(const float* d_x, int64_t pixels, int32_t channels, void* d_x2_bf16,
                         void* stream);
int tfc_gdn_train_forward_bf16(const float* d_x, const float* d_x2_gamma, const float* d_beta,
                               int64_t pixels, int32_t channels, int inverse, void* d_y_bf16,
                               float* d_norm, void* stream);
int tfc_gdn_train_backward_pre2(const void* d_gy, int gy_dtype, const float* d_x,
                                const float* d_norm, int64_t pixels, int32_t channels, int inverse,
                                float* d_gx1, float* d_gnorm, void* d_gnorm_bf16, void* stream);

/* Stand-alone GDN / IGDN on a channels_last tensor of `pixels` x `channels` (gdn.py:171-206). */
int tfc_gdn(const void* d_x, int64_t pixels, int32_t channels, const float* d_beta,
            const float* d_gamma, int inverse, int in_dtype, int out_dtype, void* d_out,
            void* stream);

#ifdef __cplusplus
}
#endif
#endif /* TFC_B200_H_ */
