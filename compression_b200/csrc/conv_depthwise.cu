// Channel-separable SignalConv2D (signal_conv.py:466-485 downsampled, :590-596 upsampled): every
// input channel is convolved with its own `mult` filters, output channel = c * mult + q — the
// layout of tf.nn.depthwise_conv2d_native.  Not on the codecs' hot path (bls2017 / bmshj2018 use
// dense layers only): a CUDA-core kernel in the normalised form of tfc_b200.h,
//   out[n,i,j,c*mult+q] = bias + sum_{a,b} U[n, i*sd + a - pt, j*sd + b - pl, c] * w[a,b,c,q]
// with U the input zero-stuffed by su.  One thread per output element; consecutive threads are
// consecutive output channels, so the loads of a tap are coalesced over c.
#include <algorithm>

#include "common.cuh"
#include "conv_common.cuh"

namespace {

struct DwParams {
  int32_t n, h, w, c, mult, ho, wo, kh, kw, sd, su, pt, pl;
  const float* in;
  const float* wgt;
  const float* bias;
  float* out;
};

__global__ void __launch_bounds__(256) depthwise_kernel(const DwParams p) {
  const int cout = p.c * p.mult;
  const int64_t total = (int64_t)p.n * p.ho * p.wo * cout;
  for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total;
       t += (int64_t)gridDim.x * blockDim.x) {
    const int co = (int)(t % cout);
    int64_t r = t / cout;
    const int j = (int)(r % p.wo);
    r /= p.wo;
    const int i = (int)(r % p.ho);
    const int n = (int)(r / p.ho);
    const int c = co / p.mult, q = co - c * p.mult;
    float acc = p.bias ? p.bias[co] : 0.f;
    for (int a = 0; a < p.kh; ++a) {
      const int y = i * p.sd + a - p.pt;
      if (y < 0 || y % p.su) continue;
      const int iy = y / p.su;
      if (iy >= p.h) continue;
      for (int b = 0; b < p.kw; ++b) {
        const int x = j * p.sd + b - p.pl;
        if (x < 0 || x % p.su) continue;
        const int ix = x / p.su;
        if (ix >= p.w) continue;
        acc = fmaf(__ldg(p.in + (((int64_t)n * p.h + iy) * p.w + ix) * p.c + c),
                   __ldg(p.wgt + (((int64_t)a * p.kw + b) * p.c + c) * p.mult + q), acc);
      }
    }
    p.out[t] = acc;
  }
}

}  // namespace

extern "C" int tfc_depthwise_conv2d(const tfc_conv_desc* d, const float* d_in, const float* d_w,
                                    const float* d_bias, float* d_out, void* stream) {
  if (!d || d->n < 0 || d->h <= 0 || d->w <= 0 || d->cin <= 0 || d->ho <= 0 || d->wo <= 0 ||
      d->cout <= 0 || d->kh <= 0 || d->kw <= 0 || d->sd <= 0 || d->su <= 0)
    return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "bad conv geometry");
  if (d->cout % d->cin)
    return tfc::fail(TFC_ERR_INVALID_ARGUMENT,
                     "channel-separable: cout (%d) must be a multiple of cin (%d)", d->cout, d->cin);
  if (d->in_dtype != TFC_F32 || d->out_dtype != TFC_F32 || d->act != TFC_ACT_NONE)
    return tfc::fail(TFC_ERR_UNIMPLEMENTED,
                     "channel-separable convolution is float32 without a fused activation");
  if (d->n == 0) return TFC_OK;
  DwParams p{};
  p.n = d->n; p.h = d->h; p.w = d->w; p.c = d->cin; p.mult = d->cout / d->cin;
  p.ho = d->ho; p.wo = d->wo; p.kh = d->kh; p.kw = d->kw; p.sd = d->sd; p.su = d->su;
  p.pt = d->pad_top; p.pl = d->pad_left;
  p.in = d_in; p.wgt = d_w; p.bias = d_bias; p.out = d_out;
  const int64_t total = (int64_t)d->n * d->ho * d->wo * d->cout;
  const int grid = (int)std::min<int64_t>((total + 255) / 256, (int64_t)tfc::num_sms() * 32);
  depthwise_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(p);
  TFC_LAUNCHED("depthwise_kernel");
  return TFC_OK;
}
