// Elementwise halves of GDN / IGDN for the training step (gdn.py:171-202 and the gradient TF
// derives for it).  The C x C contraction norm = x^2 . gamma and its two
// gradients are tensor-core GEMMs issued by the caller; what is left around them is HBM-bound
// elementwise work that autograd would run as ~12 launches per layer over the largest
// activations of the step.  Here it is three passes, one float4 per thread and array:
//   forward        norm = x^2.gamma + beta ;  y = x * rsqrt(norm)            (IGDN: x * sqrt(norm))
//   backward, pre  gx1 = gy * rsqrt(norm) ;  gnorm = -1/2 * gy * x * norm^-3/2
//                  (IGDN: gx1 = gy * sqrt(norm) ; gnorm = 1/2 * gy * x / sqrt(norm))
//   backward, post gx = gx1 + 2 * x * (gnorm . gamma^T)
// The bias-like gradient d beta = column sums of gnorm uses tfc_column_sums.
#include <cuda_bf16.h>

#include <algorithm>

#include "common.cuh"

namespace {

constexpr int kThreads = 256;

inline int grid_vec(int64_t nvec) {
  const int64_t blocks = (nvec + kThreads - 1) / kThreads;
  return (int)std::max<int64_t>(1, std::min(blocks, (int64_t)tfc::num_sms() * 8));
}
inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

template <bool INV>
__device__ __forceinline__ float fwd1(float x, float n) {
  return INV ? x * sqrtf(n) : x * rsqrtf(n);
}

__device__ __forceinline__ uint2 pack4_bf16(float a, float b, float c, float d) {
  const __nv_bfloat162 lo = __floats2bfloat162_rn(a, b), hi = __floats2bfloat162_rn(c, d);
  return make_uint2(*reinterpret_cast<const uint32_t*>(&lo), *reinterpret_cast<const uint32_t*>(&hi));
}
__device__ __forceinline__ float4 unpack4_bf16(uint2 v) {
  const float2 lo = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&v.x));
  const float2 hi = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&v.y));
  return make_float4(lo.x, lo.y, hi.x, hi.y);
}

// YB16: y leaves as bfloat16 (the next layer's convolution reads bfloat16: no cast pass)
template <bool INV, bool YB16>
__global__ void __launch_bounds__(kThreads)
gdn_train_forward_kernel(const float4* __restrict__ x, const float4* nraw,
                         const float4* __restrict__ beta, int64_t nvec, int cvec,
                         void* __restrict__ y, float4* norm) {   // norm may alias nraw
  for (int64_t i = blockIdx.x * (int64_t)kThreads + threadIdx.x; i < nvec;
       i += (int64_t)gridDim.x * kThreads) {
    const float4 b = __ldg(beta + (int)(i % cvec));
    const float4 v = __ldg(x + i);
    float4 n = nraw[i];
    n.x += b.x; n.y += b.y; n.z += b.z; n.w += b.w;
    norm[i] = n;
    const float4 r = make_float4(fwd1<INV>(v.x, n.x), fwd1<INV>(v.y, n.y), fwd1<INV>(v.z, n.z),
                                 fwd1<INV>(v.w, n.w));
    if (YB16) static_cast<uint2*>(y)[i] = pack4_bf16(r.x, r.y, r.z, r.w);
    else static_cast<float4*>(y)[i] = r;
  }
}

template <bool INV>
__device__ __forceinline__ void bwd1(float gy, float x, float n, float& gx1, float& gn) {
  if (INV) {
    const float r = sqrtf(n);
    gx1 = gy * r;
    gn = 0.5f * gy * x / r;
  } else {
    const float r = rsqrtf(n);
    gx1 = gy * r;
    gn = -0.5f * gy * x * r * r * r;
  }
}

template <bool INV>
__global__ void __launch_bounds__(kThreads)
gdn_train_backward_pre_kernel(const float4* __restrict__ gy, const float4* __restrict__ x,
                              const float4* __restrict__ norm, int64_t nvec,
                              float4* __restrict__ gx1, float4* __restrict__ gnorm) {
  for (int64_t i = blockIdx.x * (int64_t)kThreads + threadIdx.x; i < nvec;
       i += (int64_t)gridDim.x * kThreads) {
    const float4 g = __ldg(gy + i), v = __ldg(x + i), n = __ldg(norm + i);
    float4 a, b;
    bwd1<INV>(g.x, v.x, n.x, a.x, b.x);
    bwd1<INV>(g.y, v.y, n.y, a.y, b.y);
    bwd1<INV>(g.z, v.z, n.z, a.z, b.z);
    bwd1<INV>(g.w, v.w, n.w, a.w, b.w);
    gx1[i] = a;
    gnorm[i] = b;
  }
}

// x -> bf16(x^2): the A operand of the contraction when it runs on the tcgen05 1x1 convolution
__global__ void __launch_bounds__(kThreads)
gdn_train_square_kernel(const float4* __restrict__ x, int64_t nvec, uint2* __restrict__ x2) {
  for (int64_t i = blockIdx.x * (int64_t)kThreads + threadIdx.x; i < nvec;
       i += (int64_t)gridDim.x * kThreads) {
    const float4 v = __ldg(x + i);
    const __nv_bfloat162 lo = __floats2bfloat162_rn(v.x * v.x, v.y * v.y);
    const __nv_bfloat162 hi = __floats2bfloat162_rn(v.z * v.z, v.w * v.w);
    x2[i] = make_uint2(*reinterpret_cast<const uint32_t*>(&lo), *reinterpret_cast<const uint32_t*>(&hi));
  }
}

// backward, pre with a second, bfloat16 copy of gnorm (the operand of the two gradient GEMMs)
template <bool INV, bool GB16>
__global__ void __launch_bounds__(kThreads)
gdn_train_backward_pre2_kernel(const void* __restrict__ gy, const float4* __restrict__ x,
                               const float4* __restrict__ norm, int64_t nvec,
                               float4* __restrict__ gx1, float4* __restrict__ gnorm,
                               uint2* __restrict__ gnorm_bf16) {
  for (int64_t i = blockIdx.x * (int64_t)kThreads + threadIdx.x; i < nvec;
       i += (int64_t)gridDim.x * kThreads) {
    const float4 g = GB16 ? unpack4_bf16(__ldg(static_cast<const uint2*>(gy) + i))
                          : __ldg(static_cast<const float4*>(gy) + i);
    const float4 v = __ldg(x + i), n = __ldg(norm + i);
    float4 a, b;
    bwd1<INV>(g.x, v.x, n.x, a.x, b.x);
    bwd1<INV>(g.y, v.y, n.y, a.y, b.y);
    bwd1<INV>(g.z, v.z, n.z, a.z, b.z);
    bwd1<INV>(g.w, v.w, n.w, a.w, b.w);
    gx1[i] = a;
    gnorm[i] = b;
    gnorm_bf16[i] = pack4_bf16(b.x, b.y, b.z, b.w);
  }
}

__global__ void __launch_bounds__(kThreads)
gdn_train_backward_post_kernel(float4* __restrict__ gx, const float4* __restrict__ x,
                               const float4* __restrict__ gx2, int64_t nvec) {
  for (int64_t i = blockIdx.x * (int64_t)kThreads + threadIdx.x; i < nvec;
       i += (int64_t)gridDim.x * kThreads) {
    float4 a = gx[i];
    const float4 v = __ldg(x + i), g = __ldg(gx2 + i);
    a.x = fmaf(2.f * v.x, g.x, a.x); a.y = fmaf(2.f * v.y, g.y, a.y);
    a.z = fmaf(2.f * v.z, g.z, a.z); a.w = fmaf(2.f * v.w, g.w, a.w);
    gx[i] = a;
  }
}

int check_vec(int64_t pixels, int channels, std::initializer_list<const void*> ptrs) {
  if (pixels < 0 || channels <= 0 || (channels & 3))
    return tfc::fail(TFC_ERR_INVALID_ARGUMENT,
                     "GDN training kernels need a channel count that is a multiple of 4, got %d",
                     channels);
  for (const void* p : ptrs)
    if (!aligned16(p))
      return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "GDN training buffers must be 16-byte aligned");
  return TFC_OK;
}

}  // namespace

extern "C" int tfc_gdn_train_forward(const float* d_x, const float* d_x2_gamma,
                                     const float* d_beta, int64_t pixels, int32_t channels,
                                     int inverse, float* d_y, float* d_norm, void* stream) {
  int rc = check_vec(pixels, channels, {d_x, d_x2_gamma, d_beta, d_y, d_norm});
  if (rc) return rc;
  const int64_t nvec = pixels * channels / 4;
  if (nvec == 0) return TFC_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  auto a = reinterpret_cast<const float4*>(d_x);
  auto b = reinterpret_cast<const float4*>(d_x2_gamma);
  auto c = reinterpret_cast<const float4*>(d_beta);
  if (inverse)
    gdn_train_forward_kernel<true, false><<<grid_vec(nvec), kThreads, 0, st>>>(
        a, b, c, nvec, channels / 4, d_y, reinterpret_cast<float4*>(d_norm));
  else
    gdn_train_forward_kernel<false, false><<<grid_vec(nvec), kThreads, 0, st>>>(
        a, b, c, nvec, channels / 4, d_y, reinterpret_cast<float4*>(d_norm));
  TFC_LAUNCHED("gdn_train_forward_kernel");
  return TFC_OK;
}

extern "C" int tfc_gdn_train_forward_bf16(const float* d_x, const float* d_x2_gamma,
                                          const float* d_beta, int64_t pixels, int32_t channels,
                                          int inverse, void* d_y_bf16, float* d_norm,
                                          void* stream) {
  int rc = check_vec(pixels, channels, {d_x, d_x2_gamma, d_beta, d_norm});
  if (rc) return rc;
  if (reinterpret_cast<uintptr_t>(d_y_bf16) & 7)
    return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "GDN training buffers must be 16-byte aligned");
  const int64_t nvec = pixels * channels / 4;
  if (nvec == 0) return TFC_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  auto a = reinterpret_cast<const float4*>(d_x);
  auto b = reinterpret_cast<const float4*>(d_x2_gamma);
  auto c = reinterpret_cast<const float4*>(d_beta);
  if (inverse)
    gdn_train_forward_kernel<true, true><<<grid_vec(nvec), kThreads, 0, st>>>(
        a, b, c, nvec, channels / 4, d_y_bf16, reinterpret_cast<float4*>(d_norm));
  else
    gdn_train_forward_kernel<false, true><<<grid_vec(nvec), kThreads, 0, st>>>(
        a, b, c, nvec, channels / 4, d_y_bf16, reinterpret_cast<float4*>(d_norm));
  TFC_LAUNCHED("gdn_train_forward_kernel");
  return TFC_OK;
}

extern "C" int tfc_gdn_train_backward_pre(const float* d_gy, const float* d_x, const float* d_norm,
                                          int64_t pixels, int32_t channels, int inverse,
                                          float* d_gx1, float* d_gnorm, void* stream) {
  int rc = check_vec(pixels, channels, {d_gy, d_x, d_norm, d_gx1, d_gnorm});
  if (rc) return rc;
  const int64_t nvec = pixels * channels / 4;
  if (nvec == 0) return TFC_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  auto a = reinterpret_cast<const float4*>(d_gy);
  auto b = reinterpret_cast<const float4*>(d_x);
  auto c = reinterpret_cast<const float4*>(d_norm);
  if (inverse)
    gdn_train_backward_pre_kernel<true><<<grid_vec(nvec), kThreads, 0, st>>>(
        a, b, c, nvec, reinterpret_cast<float4*>(d_gx1), reinterpret_cast<float4*>(d_gnorm));
  else
    gdn_train_backward_pre_kernel<false><<<grid_vec(nvec), kThreads, 0, st>>>(
        a, b, c, nvec, reinterpret_cast<float4*>(d_gx1), reinterpret_cast<float4*>(d_gnorm));
  TFC_LAUNCHED("gdn_train_backward_pre_kernel");
  return TFC_OK;
}

extern "C" int tfc_gdn_train_backward_post(float* d_gx, const float* d_x, const float* d_gx2,
                                           int64_t pixels, int32_t channels, void* stream) {
  int rc = check_vec(pixels, channels, {d_gx, d_x, d_gx2});
  if (rc) return rc;
  const int64_t nvec = pixels * channels / 4;
  if (nvec == 0) return TFC_OK;
  gdn_train_backward_post_kernel<<<grid_vec(nvec), kThreads, 0,
                                   static_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<float4*>(d_gx), reinterpret_cast<const float4*>(d_x),
      reinterpret_cast<const float4*>(d_gx2), nvec);
  TFC_LAUNCHED("gdn_train_backward_post_kernel");
  return TFC_OK;
}

extern "C" int tfc_gdn_train_square(const float* d_x, int64_t pixels, int32_t channels,
                                    void* d_x2_bf16, void* stream) {
  int rc = check_vec(pixels, channels, {d_x});
  if (rc) return rc;
  if (reinterpret_cast<uintptr_t>(d_x2_bf16) & 7)
    return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "GDN training buffers must be 16-byte aligned");
  const int64_t nvec = pixels * channels / 4;
  if (nvec == 0) return TFC_OK;
  gdn_train_square_kernel<<<grid_vec(nvec), kThreads, 0, static_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const float4*>(d_x), nvec, static_cast<uint2*>(d_x2_bf16));
  TFC_LAUNCHED("gdn_train_square_kernel");
  return TFC_OK;
}

extern "C" int tfc_gdn_train_backward_pre2(const void* d_gy, int gy_dtype, const float* d_x,
                                           const float* d_norm, int64_t pixels, int32_t channels,
                                           int inverse, float* d_gx1, float* d_gnorm,
                                           void* d_gnorm_bf16, void* stream) {
  if (gy_dtype != TFC_F32 && gy_dtype != TFC_BF16)
    return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "gy must be float32 or bfloat16");
  int rc = check_vec(pixels, channels, {d_gy, d_x, d_norm, d_gx1, d_gnorm});
  if (rc) return rc;
  if (d_gnorm_bf16 == nullptr || (reinterpret_cast<uintptr_t>(d_gnorm_bf16) & 7))
    return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "GDN training buffers must be 16-byte aligned");
  const int64_t nvec = pixels * channels / 4;
  if (nvec == 0) return TFC_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  auto b = reinterpret_cast<const float4*>(d_x);
  auto c = reinterpret_cast<const float4*>(d_norm);
  auto o1 = reinterpret_cast<float4*>(d_gx1);
  auto o2 = reinterpret_cast<float4*>(d_gnorm);
  auto o3 = static_cast<uint2*>(d_gnorm_bf16);
  const int g = grid_vec(nvec);
  const bool b16 = gy_dtype == TFC_BF16;
  if (inverse && b16) gdn_train_backward_pre2_kernel<true, true><<<g, kThreads, 0, st>>>(d_gy, b, c, nvec, o1, o2, o3);
  else if (inverse) gdn_train_backward_pre2_kernel<true, false><<<g, kThreads, 0, st>>>(d_gy, b, c, nvec, o1, o2, o3);
  else if (b16) gdn_train_backward_pre2_kernel<false, true><<<g, kThreads, 0, st>>>(d_gy, b, c, nvec, o1, o2, o3);
  else gdn_train_backward_pre2_kernel<false, false><<<g, kThreads, 0, st>>>(d_gy, b, c, nvec, o1, o2, o3);
  TFC_LAUNCHED("gdn_train_backward_pre_kernel");
  return TFC_OK;
}
