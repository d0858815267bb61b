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


// ---- 128-bit versions ------------------------------------------------------------------------
// One float4 / int4 per thread and array, 2048 threads per SM: ~64 KB of loads in flight per SM,
// what a B200 needs to cover HBM latency at full bandwidth.  Used whenever the buffers are 16-byte
// aligned (the scalar kernels above take everything else and the count % 4 tail).
inline int grid_vec(int64_t nvec) {
  const int64_t blocks = (nvec + kThreads - 1) / kThreads;
  return (int)std::max<int64_t>(1, std::min(blocks, (int64_t)tfc::num_sms() * 8));
}
inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

__device__ __forceinline__ float gaussian_lik(float v, float sg, float scale_bound, float bound) {
  // one IEEE reciprocal instead of two divides (the kernel is FP32-pipe bound); the argument of
  // Phi moves by <= 1 ulp, far inside the 1e-3 bar on likelihoods
  const float r = 1.0f / fmaxf(sg, scale_bound);
  const float a = fabsf(v);
  return fmaxf(std_normal_cdf((0.5f - a) * r) - std_normal_cdf((-0.5f - a) * r), bound);
}

__global__ void __launch_bounds__(kThreads)
gaussian_call_vec_kernel(const float4* __restrict__ y, const float4* __restrict__ sigma,
                         int64_t nvec, float scale_bound, float bound, int mode,
                         float4* __restrict__ values, float4* __restrict__ lik) {
  for (int64_t i = blockIdx.x * (int64_t)kThreads + threadIdx.x; i < nvec;
       i += (int64_t)gridDim.x * kThreads) {
    float4 v = __ldg(y + i);
    const float4 s = __ldg(sigma + i);
    if (mode == 0) {
      v.x = floorf(v.x + 0.5f); v.y = floorf(v.y + 0.5f);
      v.z = floorf(v.z + 0.5f); v.w = floorf(v.w + 0.5f);
    }
    if (values) values[i] = v;
    float4 l;
    l.x = gaussian_lik(v.x, s.x, scale_bound, bound);
    l.y = gaussian_lik(v.y, s.y, scale_bound, bound);
    l.z = gaussian_lik(v.z, s.z, scale_bound, bound);
    l.w = gaussian_lik(v.w, s.w, scale_bound, bound);
    lik[i] = l;
  }
}

// Gradient of the Gaussian likelihood of the training graph (entropy_models.py:916-932 with
// :965-969; TF derives it through erfc, abs, the division by scale and the two lower_bound ops,
// math_ops.py:61-74).  With a = |v|, r = 1 / max(sigma, scale_bound), u = (.5 - a) r,
// l = (-.5 - a) r, phi(t) = exp(-t^2 / 2) / sqrt(2 pi):
//   d lik / d v     = sign(v) (phi(l) - phi(u)) r
//   d lik / d sigma = (l phi(l) - u phi(u)) r
// The likelihood bound passes the gradient where lik >= bound or the gradient is negative, the
// scale bound where sigma >= bound or the gradient (w.r.t. the bounded scale) is negative.
__device__ __forceinline__ void gaussian_lik_grad(float v, float sg, float g, float scale_bound,
                                                  float bound, float* gv, float* gs) {
  const float r = 1.0f / (scale_bound > 0.f ? fmaxf(sg, scale_bound) : sg);
  const float a = fabsf(v);
  const float u = (0.5f - a) * r, l = (-0.5f - a) * r;
  const float lik = std_normal_cdf(u) - std_normal_cdf(l);
  if (bound > 0.f && !(lik >= bound || g < 0.f)) g = 0.f;
  const float kInvSqrt2Pi = 0.3989422804014327f;
  const float pu = kInvSqrt2Pi * __expf(-0.5f * u * u), pl = kInvSqrt2Pi * __expf(-0.5f * l * l);
  const float sgn = v > 0.f ? 1.f : (v < 0.f ? -1.f : 0.f);
  *gv = g * sgn * (pl - pu) * r;
  float ds = g * (l * pl - u * pu) * r;
  if (scale_bound > 0.f && !(sg >= scale_bound || ds < 0.f)) ds = 0.f;
  *gs = ds;
}

__global__ void __launch_bounds__(kThreads)
gaussian_backward_kernel(const float* __restrict__ y, const float* __restrict__ sigma,
                         const float* __restrict__ grad, int64_t count, float scale_bound,
                         float bound, float* __restrict__ gy, float* __restrict__ gs) {
  for (int64_t i = blockIdx.x * (int64_t)kThreads + threadIdx.x; i < count;
       i += (int64_t)gridDim.x * kThreads) {
    float a, b;
    gaussian_lik_grad(y[i], sigma[i], grad[i], scale_bound, bound, &a, &b);
    gy[i] = a;
    gs[i] = b;
  }
}

__global__ void __launch_bounds__(kThreads)
gaussian_backward_vec_kernel(const float4* __restrict__ y, const float4* __restrict__ sigma,
                             const float4* __restrict__ grad, int64_t nvec, float scale_bound,
                             float bound, float4* __restrict__ gy, float4* __restrict__ gs) {
  for (int64_t i = blockIdx.x * (int64_t)kThreads + threadIdx.x; i < nvec;
       i += (int64_t)gridDim.x * kThreads) {
    const float4 v = __ldg(y + i), s = __ldg(sigma + i), g = __ldg(grad + i);
    float4 a, b;
    gaussian_lik_grad(v.x, s.x, g.x, scale_bound, bound, &a.x, &b.x);
    gaussian_lik_grad(v.y, s.y, g.y, scale_bound, bound, &a.y, &b.y);
    gaussian_lik_grad(v.z, s.z, g.z, scale_bound, bound, &a.z, &b.z);
    gaussian_lik_grad(v.w, s.w, g.w, scale_bound, bound, &a.w, &b.w);
    gy[i] = a;
    gs[i] = b;
  }
}

__device__ __forceinline__ int table_index(const float* s_table, int levels, float s) {
  int lo = 0, hi = levels - 1;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (s_table[mid] < s) lo = mid + 1; else hi = mid;
  }
  return lo;
}

__global__ void __launch_bounds__(kThreads)
gaussian_symbols_indexes_vec_kernel(const float4* __restrict__ y, const float4* __restrict__ sigma,
                                    int64_t nvec, const float* __restrict__ table, int levels,
                                    float scale_bound, int4* __restrict__ sym,
                                    int4* __restrict__ idx) {
  __shared__ float s_table[kMaxLevels];
  for (int k = threadIdx.x; k < levels; k += kThreads) s_table[k] = table[k];
  __syncthreads();
  for (int64_t i = blockIdx.x * (int64_t)kThreads + threadIdx.x; i < nvec;
       i += (int64_t)gridDim.x * kThreads) {
    if (sym) {
      const float4 v = __ldg(y + i);
      sym[i] = make_int4((int32_t)floorf(v.x + 0.5f), (int32_t)floorf(v.y + 0.5f),
                         (int32_t)floorf(v.z + 0.5f), (int32_t)floorf(v.w + 0.5f));
    }
    if (idx) {
      const float4 s = __ldg(sigma + i);
      idx[i] = make_int4(table_index(s_table, levels, fmaxf(s.x, scale_bound)),
                         table_index(s_table, levels, fmaxf(s.y, scale_bound)),
                         table_index(s_table, levels, fmaxf(s.z, scale_bound)),
                         table_index(s_table, levels, fmaxf(s.w, scale_bound)));
    }
  }
}

// channels % 4 == 0: the four elements of a vector are four consecutive channels
__global__ void __launch_bounds__(kThreads)
eb_symbols_vec_kernel(const float4* __restrict__ x, int64_t nvec, int channels,
                      const float4* __restrict__ medians, int4* __restrict__ sym) {
  const int cvec = channels >> 2;
  for (int64_t i = blockIdx.x * (int64_t)kThreads + threadIdx.x; i < nvec;
       i += (int64_t)gridDim.x * kThreads) {
    const float4 m = __ldg(medians + (int)(i % cvec));
    const float4 v = __ldg(x + i);
    sym[i] = make_int4((int32_t)floorf(v.x + (0.5f - m.x)), (int32_t)floorf(v.y + (0.5f - m.y)),
                       (int32_t)floorf(v.z + (0.5f - m.z)), (int32_t)floorf(v.w + (0.5f - m.w)));
  }
}

template <typename OutT>
__global__ void __launch_bounds__(kThreads)
dequantize_vec_kernel(const int4* __restrict__ sym, int64_t nvec, int channels,
                      const float4* __restrict__ medians, OutT* __restrict__ out) {
  const int cvec = channels >> 2;
  for (int64_t i = blockIdx.x * (int64_t)kThreads + threadIdx.x; i < nvec;
       i += (int64_t)gridDim.x * kThreads) {
    const int4 q = __ldg(sym + i);
    float4 v = make_float4((float)q.x, (float)q.y, (float)q.z, (float)q.w);
    if (medians) {
      const float4 m = __ldg(medians + (int)(i % cvec));
      v.x += m.x; v.y += m.y; v.z += m.z; v.w += m.w;
    }
    if (sizeof(OutT) == 4) {
      reinterpret_cast<float4*>(out)[i] = v;
    } else {
      __nv_bfloat162 a = __floats2bfloat162_rn(v.x, v.y), b = __floats2bfloat162_rn(v.z, v.w);
      uint2 w;
      w.x = *reinterpret_cast<uint32_t*>(&a);
      w.y = *reinterpret_cast<uint32_t*>(&b);
      reinterpret_cast<uint2*>(out)[i] = w;
    }
  }
}

__global__ void __launch_bounds__(kThreads)
abs_bf16_vec_kernel(const float4* __restrict__ x, int64_t nvec, uint2* __restrict__ out) {
  for (int64_t i = blockIdx.x * (int64_t)kThreads + threadIdx.x; i < nvec;
       i += (int64_t)gridDim.x * kThreads) {
    const float4 v = __ldg(x + i);
    __nv_bfloat162 a = __floats2bfloat162_rn(fabsf(v.x), fabsf(v.y));
    __nv_bfloat162 b = __floats2bfloat162_rn(fabsf(v.z), fabsf(v.w));
    uint2 w;
    w.x = *reinterpret_cast<uint32_t*>(&a);
    w.y = *reinterpret_cast<uint32_t*>(&b);
    out[i] = w;
  }
}

// EntropyBottleneck.__call__, four channels per thread.  The packed per-channel parameters are
// staged TRANSPOSED in shared memory ([parameter][channel]) so a warp's float4 reads of one
// parameter are contiguous (in global memory consecutive channels are `stride` floats apart:
// every lane would touch its own sector, ~230 of them per vector).  tanh and the sigmoid use
// ex2.approx / rcp.approx (absolute error ~1e-7; the bar on likelihoods is 1e-3): the
// likelihood needs 18 tanh + 2 sigmoid per element and is otherwise bound by the SFU pipe far
// below HBM speed.
__device__ __forceinline__ float fast_tanh(float t) {
  return 1.0f - __fdividef(2.0f, 1.0f + __expf(2.0f * t));
}
__device__ __forceinline__ float fast_sigmoid(float t) {
  return __fdividef(1.0f, 1.0f + __expf(-t));
}
__device__ __forceinline__ float4 f4(float a) { return make_float4(a, a, a, a); }
__device__ __forceinline__ float4 f4_fma(float4 a, float4 b, float4 c) {
  return make_float4(fmaf(a.x, b.x, c.x), fmaf(a.y, b.y, c.y), fmaf(a.z, b.z, c.z),
                     fmaf(a.w, b.w, c.w));
}
__device__ __forceinline__ float4 f4_gate(float4 t, float4 a) {   // t + a * tanh(t)
  return make_float4(fmaf(a.x, fast_tanh(t.x), t.x), fmaf(a.y, fast_tanh(t.y), t.y),
                     fmaf(a.z, fast_tanh(t.z), t.z), fmaf(a.w, fast_tanh(t.w), t.w));
}

template <int L, int F>
__device__ __forceinline__ float4 eb_logits4(const float* __restrict__ sp, int channels, int c0,
                                             float4 v) {
  auto P = [&](int j) { return *reinterpret_cast<const float4*>(sp + (size_t)j * channels + c0); };
  if (L == 0) return f4_fma(P(0), v, P(1));
  float4 h[F];
#pragma unroll
  for (int j = 0; j < F; ++j) h[j] = f4_gate(f4_fma(P(j), v, P(F + j)), P(2 * F + j));
  int base = 3 * F;
#pragma unroll
  for (int l = 1; l < L; ++l) {
    float4 g[F];
#pragma unroll
    for (int j = 0; j < F; ++j) {
      float4 t = P(base + F * F + j);
#pragma unroll
      for (int k = 0; k < F; ++k) t = f4_fma(P(base + j * F + k), h[k], t);
      g[j] = f4_gate(t, P(base + F * F + F + j));
    }
#pragma unroll
    for (int j = 0; j < F; ++j) h[j] = g[j];
    base += F * F + 2 * F;
  }
  float4 out = P(base + F);
#pragma unroll
  for (int k = 0; k < F; ++k) out = f4_fma(P(base + k), h[k], out);
  return out;
}

__device__ __forceinline__ float eb_lik_from_logits(float lower, float upper, float bound) {
  const float t = lower + upper;
  const float sgn = t > 0.f ? -1.f : (t < 0.f ? 1.f : 0.f);
  return fmaxf(fabsf(fast_sigmoid(sgn * upper) - fast_sigmoid(sgn * lower)), bound);
}

template <int L, int F>
__global__ void __launch_bounds__(kThreads)
eb_call_vec_kernel(const float4* __restrict__ x, int64_t nvec, int channels,
                   const float* __restrict__ params, const float4* __restrict__ medians,
                   float bound, int mode, float4* __restrict__ values, float4* __restrict__ lik) {
  extern __shared__ __align__(16) float sp[];
  constexpr int stride = eb_param_stride(L, F);
  for (int e = threadIdx.x; e < channels * stride; e += kThreads) {
    const int c = e / stride, j = e - c * stride;
    sp[(size_t)j * channels + c] = params[e];
  }
  __syncthreads();
  const int cvec = channels >> 2;
  for (int64_t i = blockIdx.x * (int64_t)kThreads + threadIdx.x; i < nvec;
       i += (int64_t)gridDim.x * kThreads) {
    const int cv = (int)(i % cvec);
    float4 v = __ldg(x + i);
    if (mode == 0) {
      const float4 m = __ldg(medians + cv);
      v.x = floorf(v.x + (0.5f - m.x)) + m.x; v.y = floorf(v.y + (0.5f - m.y)) + m.y;
      v.z = floorf(v.z + (0.5f - m.z)) + m.z; v.w = floorf(v.w + (0.5f - m.w)) + m.w;
    }
    if (values) values[i] = v;
    const float4 lo = eb_logits4<L, F>(sp, channels, 4 * cv,
                                       make_float4(v.x - .5f, v.y - .5f, v.z - .5f, v.w - .5f));
    const float4 up = eb_logits4<L, F>(sp, channels, 4 * cv,
                                       make_float4(v.x + .5f, v.y + .5f, v.z + .5f, v.w + .5f));
    lik[i] = make_float4(eb_lik_from_logits(lo.x, up.x, bound), eb_lik_from_logits(lo.y, up.y, bound),
                         eb_lik_from_logits(lo.z, up.z, bound), eb_lik_from_logits(lo.w, up.w, bound));
  }
}


// ---- EntropyBottleneck likelihood: backward (training step) ---------------------
// Gradient of lik = max(|sigmoid(s*upper) - sigmoid(s*lower)|, bound) (entropy_models.py:626-658,
// bound with the lower_bound gradient rule of ops/math_ops.py:61-74) with respect to the input
// and to the PACKED effective parameters (softplus / tanh already applied; autograd carries the
// gradient on to the stored variables).  The forward pass is recomputed per element; one thread
// owns one channel over a chunk of pixels, so the 43 parameter gradients of a channel accumulate
// in registers and reach memory as one atomic add per parameter and chunk.
template <int L, int F>
struct EbTape {          // activations of one evaluation of the density network
  float t0v;             // the input v
  float th[L][F];        // tanh(t) of every gated layer
  float h[L][F];         // layer outputs
};

template <int L, int F>
__device__ __forceinline__ float eb_forward_tape(const float* __restrict__ p, float v,
                                                 EbTape<L, F>& tp) {
  tp.t0v = v;
#pragma unroll
  for (int j = 0; j < F; ++j) {
    const float t = p[j] * v + p[F + j];
    tp.th[0][j] = tanhf(t);
    tp.h[0][j] = t + p[2 * F + j] * tp.th[0][j];
  }
  int base = 3 * F;
#pragma unroll
  for (int l = 1; l < L; ++l) {
#pragma unroll
    for (int j = 0; j < F; ++j) {
      float t = p[base + F * F + j];
#pragma unroll
      for (int k = 0; k < F; ++k) t = fmaf(p[base + j * F + k], tp.h[l - 1][k], t);
      tp.th[l][j] = tanhf(t);
      tp.h[l][j] = t + p[base + F * F + F + j] * tp.th[l][j];
    }
    base += F * F + 2 * F;
  }
  float out = p[base + F];
#pragma unroll
  for (int k = 0; k < F; ++k) out = fmaf(p[base + k], tp.h[L - 1][k], out);
  return out;
}

// adds dout * d(logit)/d(params) to g[], returns dout * d(logit)/dv
template <int L, int F>
__device__ __forceinline__ float eb_backward_tape(const float* __restrict__ p,
                                                  const EbTape<L, F>& tp, float dout, float* g) {
  constexpr int last = 3 * F + (L - 1) * (F * F + 2 * F);
  float dh[F];
  g[last + F] += dout;
#pragma unroll
  for (int k = 0; k < F; ++k) {
    g[last + k] += dout * tp.h[L - 1][k];
    dh[k] = dout * p[last + k];
  }
  int base = last;
#pragma unroll
  for (int l = L - 1; l >= 1; --l) {
    base -= F * F + 2 * F;
    float dprev[F];
#pragma unroll
    for (int k = 0; k < F; ++k) dprev[k] = 0.f;
#pragma unroll
    for (int j = 0; j < F; ++j) {
      const float th = tp.th[l][j];
      g[base + F * F + F + j] += dh[j] * th;
      const float dt = dh[j] * (1.f + p[base + F * F + F + j] * (1.f - th * th));
      g[base + F * F + j] += dt;
#pragma unroll
      for (int k = 0; k < F; ++k) {
        g[base + j * F + k] += dt * tp.h[l - 1][k];
        dprev[k] = fmaf(dt, p[base + j * F + k], dprev[k]);
      }
    }
#pragma unroll
    for (int k = 0; k < F; ++k) dh[k] = dprev[k];
  }
  float dv = 0.f;
#pragma unroll
  for (int j = 0; j < F; ++j) {
    const float th = tp.th[0][j];
    g[2 * F + j] += dh[j] * th;
    const float dt = dh[j] * (1.f + p[2 * F + j] * (1.f - th * th));
    g[F + j] += dt;
    g[j] += dt * tp.t0v;
    dv = fmaf(dt, p[j], dv);
  }
  return dv;
}

template <int L, int F>
__global__ void __launch_bounds__(128)
eb_likelihood_backward_kernel(const float* __restrict__ x, int64_t pixels, int channels,
                              int pixels_per_cta, const float* __restrict__ params, float bound,
                              const float* __restrict__ grad_lik, float* __restrict__ grad_x,
                              float* __restrict__ grad_params) {
  constexpr int stride = eb_param_stride(L, F);
  const int c = blockIdx.y * 128 + threadIdx.x;
  if (c >= channels) return;
  float p[stride], g[stride];
#pragma unroll
  for (int j = 0; j < stride; ++j) { p[j] = params[(int64_t)c * stride + j]; g[j] = 0.f; }
  const int64_t p0 = (int64_t)blockIdx.x * pixels_per_cta;
  const int64_t p1 = min(pixels, p0 + pixels_per_cta);
  for (int64_t q = p0; q < p1; ++q) {
    const int64_t i = q * channels + c;
    const float v = x[i];
    EbTape<L, F> lo, up;
    const float l = eb_forward_tape<L, F>(p, v - 0.5f, lo);
    const float u = eb_forward_tape<L, F>(p, v + 0.5f, up);
    const float t = l + u;
    const float s = t > 0.f ? -1.f : (t < 0.f ? 1.f : 0.f);
    const float su = 1.0f / (1.0f + expf(-s * u)), sl = 1.0f / (1.0f + expf(-s * l));
    const float diff = su - sl;
    const float sg = diff > 0.f ? 1.f : (diff < 0.f ? -1.f : 0.f);
    float gl = grad_lik[i];
    if (!(fabsf(diff) >= bound || gl < 0.f)) gl = 0.f;      // lower_bound: identity_if_towards
    const float du = gl * sg * s * su * (1.f - su);
    const float dl = -gl * sg * s * sl * (1.f - sl);
    grad_x[i] = eb_backward_tape<L, F>(p, up, du, g) + eb_backward_tape<L, F>(p, lo, dl, g);
  }
#pragma unroll
  for (int j = 0; j < stride; ++j) atomicAdd(grad_params + (int64_t)c * stride + j, g[j]);
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
  const size_t sp_bytes = (size_t)channels * eb_param_stride(num_filters, filter_width) * 4;
  const bool vec = (channels & 3) == 0 && aligned16(d_x) && aligned16(d_lik) &&
                   aligned16(d_values) && aligned16(d_medians) && sp_bytes <= 160 * 1024;
  const int64_t nvec = count / 4;
  const int vgrid = (int)std::max<int64_t>(1, std::min<int64_t>((nvec + kThreads - 1) / kThreads,
                                                                (int64_t)tfc::num_sms() * 4));
#define CALL(L_, F_)                                                                        \
  if (vec) {                                                                                \
    if (sp_bytes > 48 * 1024)                                                               \
      TFC_CUDA(cudaFuncSetAttribute(eb_call_vec_kernel<L_, F_>,                             \
                                    cudaFuncAttributeMaxDynamicSharedMemorySize,            \
                                    (int)sp_bytes));                                        \
    eb_call_vec_kernel<L_, F_><<<vgrid, kThreads, sp_bytes, st>>>(                          \
        reinterpret_cast<const float4*>(d_x), nvec, channels, d_params,                     \
        reinterpret_cast<const float4*>(d_medians), likelihood_bound, mode,                 \
        reinterpret_cast<float4*>(d_values), reinterpret_cast<float4*>(d_lik));             \
  } else                                                                                    \
    eb_call_kernel<L_, F_><<<grid, kThreads, 0, st>>>(d_x, count, channels, d_params,       \
                                                      d_medians, likelihood_bound, mode,    \
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

extern "C" int tfc_eb_likelihood_backward(const float* d_x, int64_t count, int32_t channels,
                                          const float* d_params, int32_t num_filters,
                                          int32_t filter_width, float likelihood_bound,
                                          const float* d_grad_lik, float* d_grad_x,
                                          float* d_grad_params, void* stream) {
  if (channels <= 0 || count < 0 || count % channels)
    return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "count %lld is not a multiple of channels %d",
                     (long long)count, channels);
  if (num_filters < 1)
    return eb_unsupported(num_filters, filter_width);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const size_t pbytes = (size_t)channels * eb_param_stride(num_filters, filter_width) * 4;
  TFC_CUDA(cudaMemsetAsync(d_grad_params, 0, pbytes, st));
  if (count == 0) return TFC_OK;
  const int64_t pixels = count / channels;
  const int ygrid = (channels + 127) / 128;
  // ~4 CTAs per SM; every CTA ends with one atomic per parameter and thread
  const int64_t want = std::max<int64_t>(1, (int64_t)tfc::num_sms() * 4 / ygrid);
  const int ppc = (int)std::max<int64_t>(1, (pixels + want - 1) / want);
  const dim3 grid((unsigned)((pixels + ppc - 1) / ppc), (unsigned)ygrid, 1);
#define CALL(L_, F_)                                                                          \
  eb_likelihood_backward_kernel<L_, F_><<<grid, 128, 0, st>>>(                                \
      d_x, pixels, channels, ppc, d_params, likelihood_bound, d_grad_lik, d_grad_x,           \
      d_grad_params)
  EB_DISPATCH(3, 3, CALL) EB_DISPATCH(1, 3, CALL) EB_DISPATCH(2, 3, CALL) EB_DISPATCH(4, 3, CALL)
  EB_DISPATCH(1, 1, CALL) EB_DISPATCH(2, 1, CALL) EB_DISPATCH(3, 1, CALL) EB_DISPATCH(1, 2, CALL)
  EB_DISPATCH(2, 2, CALL) EB_DISPATCH(3, 2, CALL) EB_DISPATCH(1, 4, CALL) EB_DISPATCH(2, 4, CALL)
  EB_DISPATCH(3, 4, CALL)
  return eb_unsupported(num_filters, filter_width);
#undef CALL
  TFC_LAUNCHED("eb_likelihood_backward_kernel");
  return TFC_OK;
}

extern "C" int tfc_eb_symbols(const float* d_x, int64_t count, int32_t channels,
                              const float* d_medians, int32_t* d_symbols, void* stream) {
  if (count <= 0) return TFC_OK;
  if (channels > 0 && (channels & 3) == 0 && count % channels == 0 && aligned16(d_x) &&
      aligned16(d_symbols) && aligned16(d_medians))
    eb_symbols_vec_kernel<<<grid_vec(count / 4), kThreads, 0, static_cast<cudaStream_t>(stream)>>>(
        reinterpret_cast<const float4*>(d_x), count / 4, channels,
        reinterpret_cast<const float4*>(d_medians), reinterpret_cast<int4*>(d_symbols));
  else
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
  const bool chan_ok = d_medians == nullptr ||
                       (channels > 0 && (channels & 3) == 0 && count % channels == 0 &&
                        aligned16(d_medians));
  if (chan_ok && (count & 3) == 0 && aligned16(d_symbols) && aligned16(d_out)) {
    const float4* med = reinterpret_cast<const float4*>(d_medians);
    const int4* q = reinterpret_cast<const int4*>(d_symbols);
    const int ch = d_medians ? channels : 4;
    if (out_bf16)
      dequantize_vec_kernel<__nv_bfloat16><<<grid_vec(count / 4), kThreads, 0, st>>>(
          q, count / 4, ch, med, static_cast<__nv_bfloat16*>(d_out));
    else
      dequantize_vec_kernel<float><<<grid_vec(count / 4), kThreads, 0, st>>>(
          q, count / 4, ch, med, static_cast<float*>(d_out));
    TFC_LAUNCHED("dequantize_kernel");
    return TFC_OK;
  }
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
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  int64_t done = 0;
  if (count >= 4 && aligned16(d_y) && aligned16(d_sigma) && aligned16(d_values) && aligned16(d_lik)) {
    done = count & ~(int64_t)3;
    gaussian_call_vec_kernel<<<grid_vec(done / 4), kThreads, 0, st>>>(
        reinterpret_cast<const float4*>(d_y), reinterpret_cast<const float4*>(d_sigma), done / 4,
        scale_bound, likelihood_bound, mode, reinterpret_cast<float4*>(d_values),
        reinterpret_cast<float4*>(d_lik));
  }
  if (done < count)
    gaussian_call_kernel<<<grid_for(count - done), kThreads, 0, st>>>(
        d_y + done, d_sigma + done, count - done, scale_bound, likelihood_bound, mode,
        d_values ? d_values + done : nullptr, d_lik + done);
  TFC_LAUNCHED("gaussian_call_kernel");
  return TFC_OK;
}

extern "C" int tfc_gaussian_likelihood_backward(const float* d_y, const float* d_sigma,
                                                const float* d_grad_lik, int64_t count,
                                                float scale_bound, float likelihood_bound,
                                                float* d_grad_y, float* d_grad_sigma,
                                                void* stream) {
  if (count <= 0) return TFC_OK;
  if (!d_y || !d_sigma || !d_grad_lik || !d_grad_y || !d_grad_sigma)
    return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "gaussian likelihood backward: null buffer");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  int64_t done = 0;
  if (count >= 4 && aligned16(d_y) && aligned16(d_sigma) && aligned16(d_grad_lik) &&
      aligned16(d_grad_y) && aligned16(d_grad_sigma)) {
    done = count & ~(int64_t)3;
    gaussian_backward_vec_kernel<<<grid_vec(done / 4), kThreads, 0, st>>>(
        reinterpret_cast<const float4*>(d_y), reinterpret_cast<const float4*>(d_sigma),
        reinterpret_cast<const float4*>(d_grad_lik), done / 4, scale_bound, likelihood_bound,
        reinterpret_cast<float4*>(d_grad_y), reinterpret_cast<float4*>(d_grad_sigma));
  }
  if (done < count)
    gaussian_backward_kernel<<<grid_for(count - done), kThreads, 0, st>>>(
        d_y + done, d_sigma + done, d_grad_lik + done, count - done, scale_bound,
        likelihood_bound, d_grad_y + done, d_grad_sigma + done);
  TFC_LAUNCHED("gaussian_backward_kernel");
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
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  int64_t done = 0;
  if (count >= 4 && aligned16(d_y) && aligned16(d_sigma) && aligned16(d_symbols) &&
      aligned16(d_indexes)) {
    done = count & ~(int64_t)3;
    gaussian_symbols_indexes_vec_kernel<<<grid_vec(done / 4), kThreads, 0, st>>>(
        reinterpret_cast<const float4*>(d_y), reinterpret_cast<const float4*>(d_sigma), done / 4,
        d_table, levels, scale_bound, reinterpret_cast<int4*>(d_symbols),
        reinterpret_cast<int4*>(d_indexes));
  }
  if (done < count)
    gaussian_symbols_indexes_kernel<<<grid_for(count - done), kThreads, 0, st>>>(
        d_y ? d_y + done : nullptr, d_sigma ? d_sigma + done : nullptr, count - done, d_table,
        levels, scale_bound, d_symbols ? d_symbols + done : nullptr,
        d_indexes ? d_indexes + done : nullptr);
  TFC_LAUNCHED("gaussian_symbols_indexes_kernel");
  return TFC_OK;
}

extern "C" int tfc_abs_to_bf16(const float* d_x, int64_t count, void* d_out_bf16, void* stream) {
  if (count <= 0) return TFC_OK;
  if ((count & 3) == 0 && aligned16(d_x) && (reinterpret_cast<uintptr_t>(d_out_bf16) & 7) == 0) {
    abs_bf16_vec_kernel<<<grid_vec(count / 4), kThreads, 0, static_cast<cudaStream_t>(stream)>>>(
        reinterpret_cast<const float4*>(d_x), count / 4, static_cast<uint2*>(d_out_bf16));
    TFC_LAUNCHED("abs_bf16_kernel");
    return TFC_OK;
  }
  abs_bf16_kernel<<<grid_for(count), kThreads, 0, static_cast<cudaStream_t>(stream)>>>(
      d_x, count, static_cast<__nv_bfloat16*>(d_out_bf16));
  TFC_LAUNCHED("abs_bf16_kernel");
  return TFC_OK;
}
