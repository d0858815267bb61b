// Elementwise kernels of the entropy models (HBM-bound; EB likelihood also MUFU-heavy).
//
// Replaces the TF graph pieces of /root/reference/tensorflow_compression/python/layers/
// entropy_models.py: EntropyBottleneck._logits_cumulative (:402-436), ._likelihood (:626-658),
// ._quantize/_dequantize (:600-624); SymmetricConditional index lookup (:869-881),
// ._quantize/_dequantize (:887-914), ._likelihood (:916-932); GaussianConditional
// ._standardized_cumulative (:965-969); lower_bound forward (ops/math_ops.py:132-184).
// The reference evaluates the density network with two full NHWC<->C-major transposes and
// batched matmuls of inner size <= 3; here each thread evaluates it in registers for one
// element, reading the input in its native channels_last order.
#include <cuda_bf16.h>

#include "common.cuh"

namespace {

constexpr int kThreads = 256;

inline int grid_for(int64_t work_items) {
  const int64_t blocks = (work_items + kThreads - 1) / kThreads;
  const int64_t cap = (int64_t)tfc::num_sms() * 16;
  return (int)std::max<int64_t>(1, std::min(blocks, cap));
}

// ---- EntropyBottleneck density network -------------------------------------------------
// Packed parameter stride per channel for L hidden layers of width F.
__host__ __device__ constexpr int eb_param_stride(int L, int F) {
  // layer 0: F*1 + F + F ; hidden->hidden (L-1 times): F*F + F + F ; last: 1*F + 1
  return L == 0 ? 2 : (3 * F) + (L - 1) * (F * F + 2 * F) + (F + 1);
}

template <int L, int F>
__device__ __forceinline__ float eb_logits(const float* __restrict__ p, float v) {
  if (L == 0) return p[0] * v + p[1];
  float h[F];
#pragma unroll
  for (int j = 0; j < F; ++j) {
    const float t = p[j] * v + p[F + j];
    h[j] = t + p[2 * F + j] * tanhf(t);
  }
  p += 3 * F;
#pragma unroll
  for (int l = 1; l < L; ++l) {
    float g[F];
#pragma unroll
    for (int j = 0; j < F; ++j) {
      float t = p[F * F + j];
#pragma unroll
      for (int k = 0; k < F; ++k) t = fmaf(p[j * F + k], h[k], t);
      // the reference computes matmul then adds the bias; the sum order inside a <=4-term
      // dot product is the only liberty taken here
      g[j] = t + p[F * F + F + j] * tanhf(t);
    }
#pragma unroll
    for (int j = 0; j < F; ++j) h[j] = g[j];
    p += F * F + 2 * F;
  }
  float out = p[F];
#pragma unroll
  for (int k = 0; k < F; ++k) out = fmaf(p[k], h[k], out);
  return out;
}

__device__ __forceinline__ float sigmoidf(float x) { return 1.0f / (1.0f + expf(-x)); }

template <int L, int F>
__device__ __forceinline__ float eb_likelihood(const float* __restrict__ p, float v, float bound) {
  const float lower = eb_logits<L, F>(p, v - 0.5f);
  const float upper = eb_logits<L, F>(p, v + 0.5f);
  const float t = lower + upper;
  const float sgn = t > 0.f ? -1.f : (t < 0.f ? 1.f : 0.f);  // -sign(lower + upper)
  const float lik = fabsf(sigmoidf(sgn * upper) - sigmoidf(sgn * lower));
  return fmaxf(lik, bound);
}

template <int L, int F>
__global__ void __launch_bounds__(kThreads)
eb_call_kernel(const float* __restrict__ x, int64_t count, int channels,
               const float* __restrict__ params, const float* __restrict__ medians,
               float bound, int mode, float* __restrict__ values, float* __restrict__ lik) {
  constexpr int stride = eb_param_stride(L, F);
  for (int64_t i = blockIdx.x * (int64_t)kThreads + threadIdx.x; i < count;
       i += (int64_t)gridDim.x * kThreads) {
    const int c = (int)(i % channels);
    float v = x[i];
    if (mode == 0) {
      const float m = __ldg(medians + c);
      v = floorf(v + (0.5f - m)) + m;
    }
    if (values) values[i] = v;
    lik[i] = eb_likelihood<L, F>(params + (int64_t)c * stride, v, bound);
  }
}

template <int L, int F>
__global__ void __launch_bounds__(kThreads)
eb_logits_kernel(const float* __restrict__ x, int channels, int samples,
                 const float* __restrict__ params, float* __restrict__ out) {
  constexpr int stride = eb_param_stride(L, F);
  const int64_t count = (int64_t)channels * samples;
  for (int64_t i = blockIdx.x * (int64_t)kThreads + threadIdx.x; i < count;
       i += (int64_t)gridDim.x * kThreads) {
    const int c = (int)(i / samples);
    out[i] = eb_logits<L, F>(params + (int64_t)c * stride, x[i]);
  }
}

__global__ void __launch_bounds__(kThreads)
eb_symbols_kernel(const float* __restrict__ x, int64_t count, int channels,
                  const float* __restrict__ medians, int32_t* __restrict__ sym) {
  for (int64_t i = blockIdx.x * (int64_t)kThreads + threadIdx.x; i < count;
       i += (int64_t)gridDim.x * kThreads) {
    const float m = __ldg(medians + (int)(i % channels));
    sym[i] = (int32_t)floorf(x[i] + (0.5f - m));
  }
}

template <typename OutT>
__device__ __forceinline__ OutT cast_out(float v);
template <> __device__ __forceinline__ float cast_out<float>(float v) { return v; }
template <> __device__ __forceinline__ __nv_bfloat16 cast_out<__nv_bfloat16>(float v) {
  return __float2bfloat16_rn(v);
}

template <typename OutT>
__global__ void __launch_bounds__(kThreads)
dequantize_kernel(const int32_t* __restrict__ sym, int64_t count, int channels,
                  const float* __restrict__ medians, OutT* __restrict__ out) {
  for (int64_t i = blockIdx.x * (int64_t)kThreads + threadIdx.x; i < count;
       i += (int64_t)gridDim.x * kThreads) {
    float v = (float)sym[i];
    if (medians) v += __ldg(medians + (int)(i % channels));
    out[i] = cast_out<OutT>(v);
  }
}

// ---- GaussianConditional -----------------------------------------------------------------
__device__ __forceinline__ float std_normal_cdf(float t) {
  return 0.5f * erfcf(-0.70710678118654752440f * t);
}

__global__ void __launch_bounds__(kThreads)
gaussian_call_kernel(const float* __restrict__ y, const float* __restrict__ sigma, int64_t count,
                     float scale_bound, float bound, int mode, float* __restrict__ values,
                     float* __restrict__ lik) {
  for (int64_t i = blockIdx.x * (int64_t)kThreads + threadIdx.x; i < count;
       i += (int64_t)gridDim.x * kThreads) {
    float v = y[i];
    if (mode == 0) v = floorf(v + 0.5f);
    if (values) values[i] = v;
    const float s = fmaxf(sigma[i], scale_bound);
    const float a = fabsf(v);
    const float up = std_normal_cdf((0.5f - a) / s);
    const float lo = std_normal_cdf((-0.5f - a) / s);
    lik[i] = fmaxf(up - lo, bound);
  }
}

constexpr int kMaxLevels = 1024;

__global__ void __launch_bounds__(kThreads)
gaussian_symbols_indexes_kernel(const float* __restrict__ y, const float* __restrict__ sigma,
                                int64_t count, const float* __restrict__ table, int levels,
                                float scale_bound, int32_t* __restrict__ sym,
                                int32_t* __restrict__ idx) {
  __shared__ float s_table[kMaxLevels];
  for (int k = threadIdx.x; k < levels; k += kThreads) s_table[k] = table[k];
  __syncthreads();
  for (int64_t i = blockIdx.x * (int64_t)kThreads + threadIdx.x; i < count;
       i += (int64_t)gridDim.x * kThreads) {
    if (sym) sym[i] = (int32_t)floorf(y[i] + 0.5f);
    if (idx) {
      const float s = fmaxf(sigma[i], scale_bound);
      // number of k in [0, levels-1) with table[k] < s  ==  (levels-1) - #{s <= table[k]}
      int lo = 0, hi = levels - 1;
      while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (s_table[mid] < s) lo = mid + 1; else hi = mid;
      }
      idx[i] = lo;
    }
  }
}

__global__ void __launch_bounds__(kThreads)
abs_bf16_kernel(const float* __restrict__ x, int64_t count, __nv_bfloat16* __restrict__ out) {
  for (int64_t i = blockIdx.x * (int64_t)kThreads + threadIdx.x; i < count;
       i += (int64_t)gridDim.x * kThreads)
    out[i] = __float2bfloat16_rn(fabsf(x[i]));
}

#define EB_DISPATCH(L_, F_, CALL)                                              \
  if (num_filters == L_ && (L_ == 0 || filter_width == F_)) { CALL(L_, F_); } else

int eb_unsupported(int num_filters, int width) {
  return tfc::fail(TFC_ERR_UNIMPLEMENTED,
                   "EntropyBottleneck filters: %d layers of width %d not built (supported: "
                   "() and uniform widths 1..4 with up to 4 layers)", num_filters, width);
}

}  // namespace

extern "C" int tfc_eb_call(const float* d_x, int64_t count, int32_t channels,
                           const float* d_params, int32_t num_filters, int32_t filter_width,
                           const float* d_medians, float likelihood_bound, int mode,
                           float* d_values, float* d_lik, void* stream) {
  if (count <= 0) return TFC_OK;
  if (channels <= 0 || count % channels)
    return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "count %lld is not a multiple of channels %d",
                     (long long)count, channels);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int grid = grid_for(count);
#define CALL(L_, F_)                                                                        \
  eb_call_kernel<L_, F_><<<grid, kThreads, 0, st>>>(d_x, count, channels, d_params,         \
                                                    d_medians, likelihood_bound, mode,      \
                                                    d_values, d_lik)
  EB_DISPATCH(3, 3, CALL) EB_DISPATCH(0, 1, CALL) EB_DISPATCH(1, 3, CALL) EB_DISPATCH(2, 3, CALL)
  EB_DISPATCH(4, 3, CALL) EB_DISPATCH(1, 1, CALL) EB_DISPATCH(2, 1, CALL) EB_DISPATCH(3, 1, CALL)
  EB_DISPATCH(1, 2, CALL) EB_DISPATCH(2, 2, CALL) EB_DISPATCH(3, 2, CALL) EB_DISPATCH(1, 4, CALL)
  EB_DISPATCH(2, 4, CALL) EB_DISPATCH(3, 4, CALL) EB_DISPATCH(4, 4, CALL)
  return eb_unsupported(num_filters, filter_width);
#undef CALL
  TFC_LAUNCHED("eb_call_kernel");
  return TFC_OK;
}

extern "C" int tfc_eb_logits_cumulative(const float* d_x, int32_t channels, int32_t samples,
                                        const float* d_params, int32_t num_filters,
                                        int32_t filter_width, float* d_logits, void* stream) {
  const int64_t count = (int64_t)channels * samples;
  if (count <= 0) return TFC_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int grid = grid_for(count);
#define CALL(L_, F_) \
  eb_logits_kernel<L_, F_><<<grid, kThreads, 0, st>>>(d_x, channels, samples, d_params, d_logits)
  EB_DISPATCH(3, 3, CALL) EB_DISPATCH(0, 1, CALL) EB_DISPATCH(1, 3, CALL) EB_DISPATCH(2, 3, CALL)
  EB_DISPATCH(4, 3, CALL) EB_DISPATCH(1, 1, CALL) EB_DISPATCH(2, 1, CALL) EB_DISPATCH(3, 1, CALL)
  EB_DISPATCH(1, 2, CALL) EB_DISPATCH(2, 2, CALL) EB_DISPATCH(3, 2, CALL) EB_DISPATCH(1, 4, CALL)
  EB_DISPATCH(2, 4, CALL) EB_DISPATCH(3, 4, CALL) EB_DISPATCH(4, 4, CALL)
  return eb_unsupported(num_filters, filter_width);
#undef CALL
  TFC_LAUNCHED("eb_logits_kernel");
  return TFC_OK;
}

extern "C" int tfc_eb_symbols(const float* d_x, int64_t count, int32_t channels,
                              const float* d_medians, int32_t* d_symbols, void* stream) {
  if (count <= 0) return TFC_OK;
  eb_symbols_kernel<<<grid_for(count), kThreads, 0, static_cast<cudaStream_t>(stream)>>>(
      d_x, count, channels, d_medians, d_symbols);
  TFC_LAUNCHED("eb_symbols_kernel");
  return TFC_OK;
}

extern "C" int tfc_eb_dequantize(const int32_t* d_symbols, int64_t count, int32_t channels,
                                 const float* d_medians, void* d_out, int out_bf16,
                                 void* stream) {
  if (count <= 0) return TFC_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (out_bf16)
    dequantize_kernel<__nv_bfloat16><<<grid_for(count), kThreads, 0, st>>>(
        d_symbols, count, channels, d_medians, static_cast<__nv_bfloat16*>(d_out));
  else
    dequantize_kernel<float><<<grid_for(count), kThreads, 0, st>>>(
        d_symbols, count, channels, d_medians, static_cast<float*>(d_out));
  TFC_LAUNCHED("dequantize_kernel");
  return TFC_OK;
}

extern "C" int tfc_symbols_to_float(const int32_t* d_symbols, int64_t count, void* d_out,
                                    int out_bf16, void* stream) {
  return tfc_eb_dequantize(d_symbols, count, 1, nullptr, d_out, out_bf16, stream);
}

extern "C" int tfc_gaussian_call(const float* d_y, const float* d_sigma, int64_t count,
                                 float scale_bound, float likelihood_bound, int mode,
                                 float* d_values, float* d_lik, void* stream) {
  if (count <= 0) return TFC_OK;
  gaussian_call_kernel<<<grid_for(count), kThreads, 0, static_cast<cudaStream_t>(stream)>>>(
      d_y, d_sigma, count, scale_bound, likelihood_bound, mode, d_values, d_lik);
  TFC_LAUNCHED("gaussian_call_kernel");
  return TFC_OK;
}

extern "C" int tfc_gaussian_symbols_indexes(const float* d_y, const float* d_sigma,
                                            int64_t count, const float* d_table, int32_t levels,
                                            float scale_bound, int32_t* d_symbols,
                                            int32_t* d_indexes, void* stream) {
  if (count <= 0) return TFC_OK;
  if (levels < 1 || levels > kMaxLevels)
    return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "scale table must have 1..%d levels, got %d",
                     kMaxLevels, levels);
  gaussian_symbols_indexes_kernel<<<grid_for(count), kThreads, 0,
                                    static_cast<cudaStream_t>(stream)>>>(
      d_y, d_sigma, count, d_table, levels, scale_bound, d_symbols, d_indexes);
  TFC_LAUNCHED("gaussian_symbols_indexes_kernel");
  return TFC_OK;
}

extern "C" int tfc_abs_to_bf16(const float* d_x, int64_t count, void* d_out_bf16, void* stream) {
  if (count <= 0) return TFC_OK;
  abs_bf16_kernel<<<grid_for(count), kThreads, 0, static_cast<cudaStream_t>(stream)>>>(
      d_x, count, static_cast<__nv_bfloat16*>(d_out_bf16));
  TFC_LAUNCHED("abs_bf16_kernel");
  return TFC_OK;
}
