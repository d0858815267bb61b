// CUDA-core tiled SignalConv2D / GDN kernels: the shape-generic path (any cin/cout, any
// stride/upsampling/padding in the normalised form of tfc_conv_desc).  The tcgen05 implicit
// GEMM in conv_umma.cu takes over for the shapes the codecs' heavy layers use.
//
// Replaces the stock TF kernels the reference dispatches to
// (/root/reference/tensorflow_compression/python/layers/signal_conv.py:516-518 tf.nn.conv2d,
// :693-695 tf.nn.conv2d_backprop_input, :804 bias_add; gdn.py:183-202 1x1 conv + rsqrt/sqrt).
#include <cuda_bf16.h>

#include "common.cuh"
#include "conv_common.cuh"

namespace {

constexpr int BM = 64, BN = 64, BK = 16, kThreads = 256;

template <typename T> __device__ __forceinline__ float ld_f(const T* p);
template <> __device__ __forceinline__ float ld_f<float>(const float* p) { return __ldg(p); }
template <> __device__ __forceinline__ float ld_f<__nv_bfloat16>(const __nv_bfloat16* p) {
  return __bfloat162float(*p);
}
template <typename T> __device__ __forceinline__ void st_f(T* p, float v);
template <> __device__ __forceinline__ void st_f<float>(float* p, float v) { *p = v; }
template <> __device__ __forceinline__ void st_f<__nv_bfloat16>(__nv_bfloat16* p, float v) {
  *p = __float2bfloat16_rn(v);
}

// epi_x: for GDN/IGDN the tensor that gets multiplied (pre-activation, fp32, [pixels, cout])
template <typename InT, typename OutT, bool SQUARE>
__global__ void __launch_bounds__(kThreads)
conv_tiled_kernel(const tfc_conv_desc d, const InT* __restrict__ in, const float* __restrict__ w,
                  const float* __restrict__ bias, const float* __restrict__ epi_x,
                  OutT* __restrict__ out, int act) {
  __shared__ float As[BK][BM + 4];
  __shared__ float Bs[BK][BN + 4];
  const int t = threadIdx.x;
  const int64_t pixels = (int64_t)d.n * d.ho * d.wo;
  const int64_t p0 = (int64_t)blockIdx.x * BM;
  const int o0 = blockIdx.y * BN;

  // this thread's A-load assignment: one pixel, 4 consecutive channels of the K chunk
  const int lp = t % BM, lc = (t / BM) * 4;
  const int64_t gp = p0 + lp;
  const bool pvalid = gp < pixels;
  int n = 0, oi = 0, oj = 0;
  if (pvalid) {
    n = (int)(gp / ((int64_t)d.ho * d.wo));
    const int r = (int)(gp - (int64_t)n * d.ho * d.wo);
    oi = r / d.wo;
    oj = r - oi * d.wo;
  }
  // B-load assignment: row k of the chunk, 4 consecutive output channels
  const int bk = t / 16, bo = (t % 16) * 4;
  const int ty = t / 16, tx = t % 16;

  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  const int hu = d.h * d.su, wu = d.w * d.su;
  for (int a = 0; a < d.kh; ++a) {
    const int yu = oi * d.sd + a - d.pad_top;
    const bool yok = pvalid && yu >= 0 && yu < hu && (yu % d.su) == 0;
    const int yi = yu / d.su;
    for (int b = 0; b < d.kw; ++b) {
      const int xu = oj * d.sd + b - d.pad_left;
      const bool ok = yok && xu >= 0 && xu < wu && (xu % d.su) == 0;
      const int xi = xu / d.su;
      // a whole tap is skipped by the block only if no pixel of the tile needs it
      if (!__syncthreads_or(ok)) continue;
      const InT* src = in + (((int64_t)n * d.h + yi) * d.w + xi) * d.cin;
      const float* wt = w + (int64_t)(a * d.kw + b) * d.cin * d.cout;
      for (int c0 = 0; c0 < d.cin; c0 += BK) {
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const int c = c0 + lc + q;
          float v = 0.f;
          if (ok && c < d.cin) {
            v = ld_f<InT>(src + c);
            if (SQUARE) v = v * v;
          }
          As[lc + q][lp] = v;
        }
        {
          const int c = c0 + bk;
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const int o = o0 + bo + q;
            Bs[bk][bo + q] = (c < d.cin && o < d.cout) ? __ldg(wt + (int64_t)c * d.cout + o) : 0.f;
          }
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < BK; ++k) {
          const float4 av = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
          const float4 bv = *reinterpret_cast<const float4*>(&Bs[k][tx * 4]);
          const float aa[4] = {av.x, av.y, av.z, av.w};
          const float bb[4] = {bv.x, bv.y, bv.z, bv.w};
#pragma unroll
          for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(aa[i], bb[j], acc[i][j]);
        }
        __syncthreads();
      }
    }
  }

#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int64_t p = p0 + ty * 4 + i;
    if (p >= pixels) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int o = o0 + tx * 4 + j;
      if (o >= d.cout) continue;
      float v = acc[i][j] + (bias ? __ldg(bias + o) : 0.f);
      if (act == TFC_ACT_RELU) v = fmaxf(v, 0.f);
      else if (act == TFC_ACT_GDN) v = epi_x[p * d.cout + o] * rsqrtf(v);
      else if (act == TFC_ACT_IGDN) v = epi_x[p * d.cout + o] * sqrtf(v);
      st_f<OutT>(out + p * d.cout + o, v);
    }
  }
}

template <typename InT, typename OutT, bool SQUARE>
int launch_tiled(const tfc_conv_desc& d, const void* in, const float* w, const float* bias,
                 const float* epi_x, void* out, int act, cudaStream_t st) {
  const int64_t pixels = (int64_t)d.n * d.ho * d.wo;
  if (pixels == 0) return TFC_OK;
  dim3 grid((unsigned)((pixels + BM - 1) / BM), (unsigned)((d.cout + BN - 1) / BN));
  conv_tiled_kernel<InT, OutT, SQUARE><<<grid, kThreads, 0, st>>>(
      d, static_cast<const InT*>(in), w, bias, epi_x, static_cast<OutT*>(out), act);
  TFC_LAUNCHED("conv_tiled_kernel");
  return TFC_OK;
}

template <bool SQUARE>
int dispatch_tiled(const tfc_conv_desc& d, int in_dtype, int out_dtype, const void* in,
                   const float* w, const float* bias, const float* epi_x, void* out, int act,
                   cudaStream_t st) {
  if (in_dtype == TFC_F32 && out_dtype == TFC_F32)
    return launch_tiled<float, float, SQUARE>(d, in, w, bias, epi_x, out, act, st);
  if (in_dtype == TFC_F32 && out_dtype == TFC_BF16)
    return launch_tiled<float, __nv_bfloat16, SQUARE>(d, in, w, bias, epi_x, out, act, st);
  if (in_dtype == TFC_BF16 && out_dtype == TFC_F32)
    return launch_tiled<__nv_bfloat16, float, SQUARE>(d, in, w, bias, epi_x, out, act, st);
  if (in_dtype == TFC_BF16 && out_dtype == TFC_BF16)
    return launch_tiled<__nv_bfloat16, __nv_bfloat16, SQUARE>(d, in, w, bias, epi_x, out, act, st);
  return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "bad dtype %d/%d", in_dtype, out_dtype);
}

}  // namespace

namespace tfc {

size_t conv_direct_workspace(const tfc_conv_desc& d) {
  if (d.act == TFC_ACT_GDN || d.act == TFC_ACT_IGDN)
    return (size_t)d.n * d.ho * d.wo * d.cout * sizeof(float);
  return 0;
}

int conv_direct(const tfc_conv_desc& d, const void* in, const float* w, const float* bias,
                const float* beta, const float* gamma, void* out, void* ws, size_t ws_bytes,
                cudaStream_t st) {
  const bool gdn = d.act == TFC_ACT_GDN || d.act == TFC_ACT_IGDN;
  if (!gdn)
    return dispatch_tiled<false>(d, d.in_dtype, d.out_dtype, in, w, bias, nullptr, out, d.act, st);
  if (ws_bytes < conv_direct_workspace(d) || ws == nullptr)
    return fail(TFC_ERR_INVALID_ARGUMENT, "conv workspace too small: need %zu bytes",
                conv_direct_workspace(d));
  float* pre = static_cast<float*>(ws);
  int rc = dispatch_tiled<false>(d, d.in_dtype, TFC_F32, in, w, bias, nullptr, pre, TFC_ACT_NONE, st);
  if (rc) return rc;
  return gdn_direct(pre, (int64_t)d.n * d.ho * d.wo, d.cout, beta, gamma,
                    d.act == TFC_ACT_IGDN, TFC_F32, d.out_dtype, out, st);
}

// GDN as a 1x1 convolution of x^2 with gamma[k, c], bias beta, epilogue x * (r)sqrt(norm)
int gdn_direct(const void* x, int64_t pixels, int channels, const float* beta,
               const float* gamma, bool inverse, int in_dtype, int out_dtype, void* out,
               cudaStream_t st) {
  if (in_dtype != TFC_F32)
    return fail(TFC_ERR_UNIMPLEMENTED, "stand-alone GDN takes float32 input");
  if (pixels >= (1ll << 31))
    return fail(TFC_ERR_INVALID_ARGUMENT, "too many pixels for one GDN call");
  tfc_conv_desc g{};
  g.n = 1; g.h = 1; g.w = (int32_t)pixels; g.cin = channels;
  g.ho = 1; g.wo = (int32_t)pixels; g.cout = channels;
  g.kh = g.kw = 1; g.sd = g.su = 1; g.pad_top = g.pad_left = 0;
  return dispatch_tiled<true>(g, TFC_F32, out_dtype, x, gamma, beta,
                              static_cast<const float*>(x), out,
                              inverse ? TFC_ACT_IGDN : TFC_ACT_GDN, st);
}

}  // namespace tfc
