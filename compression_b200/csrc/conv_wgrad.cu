// Weight gradient of SignalConv2D (bls2017 training step): in the reference this
// is the filter gradient TF derives for tf.nn.conv2d / conv2d_backprop_input
// (signal_conv.py:516-518, :693-695).  In the normalised form of tfc_b200.h,
//     out[n,i,j,o] = sum_{a,b,c} U[n, i*sd + a - pt, j*sd + b - pl, c] * w[a,b,c,o]
// (U = input zero-stuffed by su), the gradient is, per tap (a,b), a [cin x P] . [P x cout] GEMM
// over the P pixels where the tap lands inside the image:
//     dw[a,b,c,o] = sum_{n,i,j} U[n, i*sd + a - pt, j*sd + b - pl, c] * dout[n,i,j,o].
// NHWC makes both operands pixel-major with the channels contiguous — exactly the col-major A /
// row-major B fragments of the warp-level MMA (mma.sync via nvcuda::wmma, bf16 in, fp32
// accumulate), so tiles go global -> shared -> fragments without a transpose.  Pixels are split
// over `kslices` CTAs per (tap, channel tile); partial sums are added with fp32 atomics.
//
// (The forward transforms are tcgen05; the filter gradient is ~1/3 of a training step's FLOPs
// and is the next candidate for a tcgen05 kernel with MN-major operand descriptors.)
#include <cuda_bf16.h>
#include <mma.h>

#include <algorithm>
#include <cmath>
#include <cstdlib>

#include "common.cuh"
#include "conv_common.cuh"

namespace {

using namespace nvcuda;

constexpr int kBM = 64, kBN = 64, kBK = 32;
constexpr int kLd = 72;                  // padded leading dimension (multiple of 8)
constexpr int kThreads = 128;

struct WgradParams {
  int32_t n, h, w, cin, ho, wo, cout, kh, kw, sd, su, pt, pl;
  int32_t ctiles, otiles, kslices;
  // thin layers (3 image channels): the kw taps of one kernel row are contiguous in memory on
  // the thin side ((column, channel) of NHWC), so one "tap" is a whole kernel row and the thin
  // dimension becomes kw * channels wide (27 of a 64-wide tile instead of 3).
  //   fuse 1: su == 1, thin cin:  x element e = b * cin + c of row a
  //   fuse 2: su  > 1, thin cout: dy element e = (kw - 1 - b) * cout + o of row a
  int32_t fuse;
  int64_t pixels;                        // enumerated pixels: outputs (su == 1) or inputs (su > 1)
  const __nv_bfloat16* x;
  const __nv_bfloat16* dy;
  float* dw;
};

__global__ void __launch_bounds__(kThreads, 8)
conv_wgrad_kernel(const WgradParams p) {
  __shared__ __align__(32) __nv_bfloat16 xs[kBK * kLd];
  __shared__ __align__(32) __nv_bfloat16 ys[kBK * kLd];
  __shared__ int64_t x_off[kBK], y_off[kBK];
  __shared__ int32_t e_lo[kBK], e_hi[kBK];   // fused side: valid element range of the row
  __shared__ __align__(32) float acc_s[kBM * kBN];

  const int tile = blockIdx.x;
  const int ct = tile / p.otiles, ot = tile - ct * p.otiles;
  const int tap = blockIdx.y;
  const int a = p.fuse ? tap : tap / p.kw, b = p.fuse ? 0 : tap - a * p.kw;
  const int cin_t = p.fuse == 1 ? p.kw * p.cin : p.cin;      // tile extents of the two sides
  const int cout_t = p.fuse == 2 ? p.kw * p.cout : p.cout;
  const int c0 = ct * kBM, o0 = ot * kBN;
  const int warp = threadIdx.x >> 5;
  const int wm = (warp >> 1) * 32, wn = (warp & 1) * 32;

  wmma::fragment<wmma::accumulator, 16, 16, 16, float> acc[2][2];
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int j = 0; j < 2; ++j) wmma::fill_fragment(acc[i][j], 0.f);

  const int64_t per = (p.pixels + p.kslices - 1) / p.kslices;
  const int64_t p_begin = (int64_t)blockIdx.z * per;
  const int64_t p_end = min(p.pixels, p_begin + per);
  const bool vec_x = (p.cin & 7) == 0 && p.fuse != 1, vec_y = (p.cout & 7) == 0 && p.fuse != 2;
  const int gw = p.su > 1 ? p.w : p.wo, gh = p.su > 1 ? p.h : p.ho;   // enumeration grid

  for (int64_t base = p_begin; base < p_end; base += kBK) {
    if (threadIdx.x < kBK) {
      const int64_t q = base + threadIdx.x;
      int64_t xo = -1, yo = -1;
      int lo = 0, hi = 0x7FFFFFFF;
      if (q < p_end && p.fuse == 1) {        // whole kernel row a of output pixel (i, j)
        // 32-bit division (pixels < 2^31 is checked on the host): this decomposition sits on the
        // serial part of every iteration
        const uint32_t qq = (uint32_t)q, row = qq / (uint32_t)gw;
        const int j = (int)(qq - row * (uint32_t)gw);
        const int nn = (int)(row / (uint32_t)gh);
        const int i = (int)(row - (uint32_t)nn * (uint32_t)gh);
        const int yi = i * p.sd + a - p.pt, xj0 = j * p.sd - p.pl;
        if (yi >= 0 && yi < p.h) {
          xo = (((int64_t)nn * p.h + yi) * p.w + xj0) * p.cin + ((int64_t)1 << 40);  // may be < 0
          yo = (((int64_t)nn * p.ho + i) * p.wo + j) * p.cout;
          lo = max(0, -xj0) * p.cin;
          hi = min(p.kw, p.w - xj0) * p.cin;
        }
      } else if (q < p_end && p.fuse == 2) { // input pixel (i, j) against kernel row a of dy
        // 32-bit division (pixels < 2^31 is checked on the host): this decomposition sits on the
        // serial part of every iteration
        const uint32_t qq = (uint32_t)q, row = qq / (uint32_t)gw;
        const int j = (int)(qq - row * (uint32_t)gw);
        const int nn = (int)(row / (uint32_t)gh);
        const int i = (int)(row - (uint32_t)nn * (uint32_t)gh);
        const int oi = p.su * i - a + p.pt, oj0 = p.su * j + p.pl - (p.kw - 1);
        if (oi >= 0 && oi < p.ho) {
          xo = (((int64_t)nn * p.h + i) * p.w + j) * p.cin;
          yo = (((int64_t)nn * p.ho + oi) * p.wo + oj0) * p.cout + ((int64_t)1 << 40);
          lo = max(0, -oj0) * p.cout;
          hi = min(p.kw, p.wo - oj0) * p.cout;
        }
      } else if (q < p_end) {
        // 32-bit division (pixels < 2^31 is checked on the host): this decomposition sits on the
        // serial part of every iteration
        const uint32_t qq = (uint32_t)q, row = qq / (uint32_t)gw;
        const int j = (int)(qq - row * (uint32_t)gw);
        const int nn = (int)(row / (uint32_t)gh);
        const int i = (int)(row - (uint32_t)nn * (uint32_t)gh);
        if (p.su > 1) {           // (i, j) input pixel; the tap sends it to one output pixel
          const int oi = p.su * i - a + p.pt, oj = p.su * j - b + p.pl;
          if (oi >= 0 && oi < p.ho && oj >= 0 && oj < p.wo) {
            xo = (((int64_t)nn * p.h + i) * p.w + j) * p.cin;
            yo = (((int64_t)nn * p.ho + oi) * p.wo + oj) * p.cout;
          }
        } else {                  // (i, j) output pixel; the tap reads one input pixel
          const int yi = i * p.sd + a - p.pt, xj = j * p.sd + b - p.pl;
          if (yi >= 0 && yi < p.h && xj >= 0 && xj < p.w) {
            xo = (((int64_t)nn * p.h + yi) * p.w + xj) * p.cin;
            yo = (((int64_t)nn * p.ho + i) * p.wo + j) * p.cout;
          }
        }
      }
      x_off[threadIdx.x] = xo;
      y_off[threadIdx.x] = yo;
      e_lo[threadIdx.x] = lo;
      e_hi[threadIdx.x] = hi;
    }
    __syncthreads();
    // tiles: 32 pixel rows x 64 channels, zero where the tap is outside / channels run out
    for (int e = threadIdx.x; e < kBK * 8; e += kThreads) {
      const int r = e >> 3, ch = (e & 7) * 8;
      uint4 vx = make_uint4(0, 0, 0, 0), vy = make_uint4(0, 0, 0, 0);
      int64_t xo = x_off[r], yo = y_off[r];
      if (xo >= 0) {
        const int lo = e_lo[r], hi = e_hi[r];
        if (p.fuse == 1) xo -= (int64_t)1 << 40;
        if (p.fuse == 2) yo -= (int64_t)1 << 40;
        if (vec_x) {
          if (c0 + ch < p.cin) vx = *reinterpret_cast<const uint4*>(p.x + xo + c0 + ch);
        } else {
          const int xl = p.fuse == 1 ? lo : 0, xh = p.fuse == 1 ? hi : p.cin;
          __nv_bfloat16 t[8];
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            const int e = c0 + ch + k;
            t[k] = (e >= xl && e < xh) ? p.x[xo + e] : __float2bfloat16(0.f);
          }
          vx = *reinterpret_cast<const uint4*>(t);
        }
        if (vec_y) {
          if (o0 + ch < p.cout) vy = *reinterpret_cast<const uint4*>(p.dy + yo + o0 + ch);
        } else {
          const int yl = p.fuse == 2 ? lo : 0, yh = p.fuse == 2 ? hi : p.cout;
          __nv_bfloat16 t[8];
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            const int e = o0 + ch + k;
            t[k] = (e >= yl && e < yh) ? p.dy[yo + e] : __float2bfloat16(0.f);
          }
          vy = *reinterpret_cast<const uint4*>(t);
        }
      }
      *reinterpret_cast<uint4*>(xs + r * kLd + ch) = vx;
      *reinterpret_cast<uint4*>(ys + r * kLd + ch) = vy;
    }
    __syncthreads();
#pragma unroll
    for (int k0 = 0; k0 < kBK; k0 += 16) {
      wmma::fragment<wmma::matrix_a, 16, 16, 16, __nv_bfloat16, wmma::col_major> fa[2];
      wmma::fragment<wmma::matrix_b, 16, 16, 16, __nv_bfloat16, wmma::row_major> fb[2];
#pragma unroll
      for (int i = 0; i < 2; ++i) wmma::load_matrix_sync(fa[i], xs + k0 * kLd + wm + 16 * i, kLd);
#pragma unroll
      for (int j = 0; j < 2; ++j) wmma::load_matrix_sync(fb[j], ys + k0 * kLd + wn + 16 * j, kLd);
#pragma unroll
      for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int j = 0; j < 2; ++j) wmma::mma_sync(acc[i][j], fa[i], fb[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int j = 0; j < 2; ++j)
      wmma::store_matrix_sync(acc_s + (wm + 16 * i) * kBN + wn + 16 * j, acc[i][j], kBN,
                              wmma::mem_row_major);
  __syncthreads();
  float* dst = p.dw + (int64_t)tap * (p.fuse ? p.kw : 1) * p.cin * p.cout;   // row a / tap (a, b)
  for (int e = threadIdx.x; e < kBM * kBN; e += kThreads) {
    const int c = c0 + e / kBN, o = o0 + e % kBN;
    if (c < cin_t && o < cout_t) {
      const float v = acc_s[e];
      // fuse 1: c = b * cin + c' is already the offset inside row a;  fuse 2: o = (kw-1-b)*cout + o'
      int64_t at = (int64_t)c * p.cout + o;
      if (p.fuse == 2) {
        const int bb = p.kw - 1 - o / p.cout, oo = o - (o / p.cout) * p.cout;
        at = ((int64_t)bb * p.cin + c) * p.cout + oo;
      }
      if (p.kslices == 1) dst[at] = v;
      else atomicAdd(dst + at, v);
    }
  }
}

// bias gradient: column sums of a [pixels, channels] float32 tensor (channels % 4 == 0: one float4
// per thread, 32 pixel rows per CTA pass and four passes in flight per thread, shared-memory tree
// over the rows, one atomic per channel quad and CTA).  All CTAs add into the same `channels`
// floats — a handful of cache lines, where L2 retires about one atomic per clock — so the grid is
// two CTAs per SM and the add is the 16-byte vector atomic when `out` is aligned for it: with
// 8 CTAs per SM and scalar atomics the 150 K adds took ~28 us whatever the input size
// (profiles/r1l_train_elementwise_full_summary.md).
template <bool VEC_ATOMIC>
__global__ void __launch_bounds__(256)
column_sums_kernel(const float4* __restrict__ x, int64_t pixels, int cvec,
                   float* __restrict__ out) {
  extern __shared__ float4 part[];             // [rows_per_pass][cvec]
  const int rows = 256 / min(cvec, 256) > 0 ? 256 / min(cvec, 256) : 1;
  const int col = threadIdx.x % cvec, row = threadIdx.x / cvec;
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  if (row < rows) {
    const int64_t step = (int64_t)gridDim.x * rows;
    int64_t q = (int64_t)blockIdx.x * rows + row;
    float4 a1 = acc, a2 = acc, a3 = acc;
    for (; q + 3 * step < pixels; q += 4 * step) {
      const float4 v0 = __ldg(x + q * cvec + col);
      const float4 v1 = __ldg(x + (q + step) * cvec + col);
      const float4 v2 = __ldg(x + (q + 2 * step) * cvec + col);
      const float4 v3 = __ldg(x + (q + 3 * step) * cvec + col);
      acc.x += v0.x; acc.y += v0.y; acc.z += v0.z; acc.w += v0.w;
      a1.x += v1.x; a1.y += v1.y; a1.z += v1.z; a1.w += v1.w;
      a2.x += v2.x; a2.y += v2.y; a2.z += v2.z; a2.w += v2.w;
      a3.x += v3.x; a3.y += v3.y; a3.z += v3.z; a3.w += v3.w;
    }
    for (; q < pixels; q += step) {
      const float4 v = __ldg(x + q * cvec + col);
      acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
    }
    acc.x += a1.x + (a2.x + a3.x); acc.y += a1.y + (a2.y + a3.y);
    acc.z += a1.z + (a2.z + a3.z); acc.w += a1.w + (a2.w + a3.w);
    part[row * cvec + col] = acc;
  }
  __syncthreads();
  if (row == 0) {
    for (int r = 1; r < rows; ++r) {
      const float4 v = part[r * cvec + col];
      acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
    }
    if (VEC_ATOMIC) {
      atomicAdd(reinterpret_cast<float4*>(out) + col, acc);
    } else {
      atomicAdd(out + 4 * col + 0, acc.x);
      atomicAdd(out + 4 * col + 1, acc.y);
      atomicAdd(out + 4 * col + 2, acc.z);
      atomicAdd(out + 4 * col + 3, acc.w);
    }
  }
}

// any channel count (<= 256 here): every thread walks elements with a fixed residue modulo
// `channels` — the stride is the thread count rounded DOWN to a multiple of channels (the launch
// has at least `channels` threads) — then the CTA folds its threads' sums per channel in shared
// memory, so memory sees one atomic per channel and CTA instead of one per thread
__global__ void __launch_bounds__(256)
column_sums_scalar_kernel(const float* __restrict__ x, int64_t total, int channels,
                          float* __restrict__ out) {
  __shared__ float part[256];
  for (int c = threadIdx.x; c < channels; c += 256) part[c] = 0.f;
  __syncthreads();
  const int64_t stride = (int64_t)gridDim.x * 256 / channels * channels;
  const int64_t i0 = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (i0 < stride) {
    float acc = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;      // four loads in flight per thread
    int64_t i = i0;
    for (; i + 3 * stride < total; i += 4 * stride) {
      const float v0 = __ldg(x + i), v1 = __ldg(x + i + stride), v2 = __ldg(x + i + 2 * stride),
                  v3 = __ldg(x + i + 3 * stride);
      acc += v0; a1 += v1; a2 += v2; a3 += v3;
    }
    for (; i < total; i += stride) acc += __ldg(x + i);
    acc += a1 + (a2 + a3);
    atomicAdd(part + (int)(i0 % channels), acc);
  }
  __syncthreads();
  for (int c = threadIdx.x; c < channels; c += 256) atomicAdd(out + c, part[c]);
}

// fused Adam over a flat parameter bucket (tf.train.AdamOptimizer semantics: the bias
// correction is folded into the step size, epsilon is NOT scaled)
__global__ void adam_kernel(float* __restrict__ param, const float* __restrict__ grad,
                            float* __restrict__ m, float* __restrict__ v, int64_t n, float lr_t,
                            float beta1, float beta2, float eps, float grad_scale) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x) {
    const float g = grad[i] * grad_scale;
    const float mi = beta1 * m[i] + (1.f - beta1) * g;
    const float vi = beta2 * v[i] + (1.f - beta2) * g * g;
    m[i] = mi;
    v[i] = vi;
    param[i] -= lr_t * mi / (sqrtf(vi) + eps);
  }
}

}  // namespace

extern "C" int tfc_signal_conv2d_wgrad(const tfc_conv_desc* d, const void* d_x_bf16,
                                       const void* d_dy_bf16, float* d_dw, void* stream) {
  if (!d || d->n < 0 || d->h <= 0 || d->w <= 0 || d->cin <= 0 || d->ho <= 0 || d->wo <= 0 ||
      d->cout <= 0 || d->kh <= 0 || d->kw <= 0 || d->sd <= 0 || d->su <= 0)
    return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "bad conv geometry");
  if (d->sd > 1 && d->su > 1)
    return tfc::fail(TFC_ERR_UNIMPLEMENTED, "strides_down and strides_up together");
  if ((reinterpret_cast<uintptr_t>(d_x_bf16) & 15) || (reinterpret_cast<uintptr_t>(d_dy_bf16) & 15))
    return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "conv buffers must be 16-byte aligned");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int taps = d->kh * d->kw;
  TFC_CUDA(cudaMemsetAsync(d_dw, 0, (size_t)taps * d->cin * d->cout * sizeof(float), st));
  if (d->n == 0) return TFC_OK;
  {
    static int force_wmma = -1;
    if (force_wmma < 0) {
      const char* e = getenv("TFC_WGRAD_WMMA");
      force_wmma = (e && e[0] == '1') ? 1 : 0;
    }
    if (!force_wmma && tfc::conv_wgrad_umma_supported(*d))
      return tfc::conv_wgrad_umma(*d, d_x_bf16, d_dy_bf16, d_dw, st);
    if (!force_wmma && tfc::conv_thin_wgrad_supported(*d, d_x_bf16, d_dy_bf16))
      return tfc::conv_thin_wgrad(*d, d_x_bf16, d_dy_bf16, d_dw, st);
  }
  WgradParams p{};
  p.n = d->n; p.h = d->h; p.w = d->w; p.cin = d->cin; p.ho = d->ho; p.wo = d->wo;
  p.cout = d->cout; p.kh = d->kh; p.kw = d->kw; p.sd = d->sd; p.su = d->su;
  p.pt = d->pad_top; p.pl = d->pad_left;
  p.fuse = 0;
  if (d->su == 1 && d->cin < 16 && d->kw * d->cin <= kBM) p.fuse = 1;
  if (d->su > 1 && d->cout < 16 && d->kw * d->cout <= kBN) p.fuse = 2;
  p.ctiles = ((p.fuse == 1 ? d->kw * d->cin : d->cin) + kBM - 1) / kBM;
  p.otiles = ((p.fuse == 2 ? d->kw * d->cout : d->cout) + kBN - 1) / kBN;
  p.pixels = d->su > 1 ? (int64_t)d->n * d->h * d->w : (int64_t)d->n * d->ho * d->wo;
  if (p.pixels >= ((int64_t)1 << 31))
    return tfc::fail(TFC_ERR_UNIMPLEMENTED, "filter gradient over 2^31 or more pixels");
  const int grid_taps = p.fuse ? d->kh : taps;
  const int64_t ctas = (int64_t)grid_taps * p.ctiles * p.otiles;
  // CTAs per SM: the kernel is latency bound per CTA (index decomposition -> gather -> MMA with
  // a barrier between each), so time falls with the CTAs resident per SM until the partial sums'
  // atomics take over; measured on the 3-channel layers of bls2017: 1.71 / 0.93 / 0.66 / 0.53 /
  // 0.46 / 0.53 ms at 1 / 2 / 3 / 4 / 6 / 8 per SM (TFC_WGRAD_WANT overrides, for experiments)
  static int want_mult = -1;
  if (want_mult < 0) {
    const char* e = getenv("TFC_WGRAD_WANT");
    want_mult = e ? std::max(1, atoi(e)) : 0;
  }
  const int64_t want = (want_mult ? want_mult : (p.fuse ? 6 : 4)) * (int64_t)tfc::num_sms();
  int64_t ks = std::max<int64_t>(1, (want + ctas - 1) / ctas);
  ks = std::min<int64_t>(ks, std::max<int64_t>(1, p.pixels / 1024));
  p.kslices = (int)std::min<int64_t>(ks, 65535);
  p.x = static_cast<const __nv_bfloat16*>(d_x_bf16);
  p.dy = static_cast<const __nv_bfloat16*>(d_dy_bf16);
  p.dw = d_dw;
  dim3 grid(p.ctiles * p.otiles, grid_taps, p.kslices);
  conv_wgrad_kernel<<<grid, kThreads, 0, st>>>(p);
  TFC_LAUNCHED("conv_wgrad_kernel");
  return TFC_OK;
}

extern "C" int tfc_column_sums(const float* d_x, int64_t pixels, int32_t channels, float* d_out,
                               void* stream) {
  if (pixels < 0 || channels <= 0)
    return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "bad column-sum shape");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  TFC_CUDA(cudaMemsetAsync(d_out, 0, (size_t)channels * sizeof(float), st));
  if (pixels == 0) return TFC_OK;
  const int sms = tfc::num_sms();
  if ((channels & 3) == 0 && channels <= 1024 && (reinterpret_cast<uintptr_t>(d_x) & 15) == 0) {
    const int cvec = channels / 4;
    const int rows = std::max(1, 256 / cvec);
    const int grid = (int)std::min<int64_t>((pixels + rows - 1) / rows, (int64_t)sms * 2);
    const size_t smem = (size_t)rows * cvec * sizeof(float4);
    if ((reinterpret_cast<uintptr_t>(d_out) & 15) == 0)
      column_sums_kernel<true><<<grid, 256, smem, st>>>(reinterpret_cast<const float4*>(d_x),
                                                        pixels, cvec, d_out);
    else
      column_sums_kernel<false><<<grid, 256, smem, st>>>(reinterpret_cast<const float4*>(d_x),
                                                         pixels, cvec, d_out);
  } else {
    if (channels > 256)
      return tfc::fail(TFC_ERR_UNIMPLEMENTED,
                       "column sums over %d channels need a 16-byte aligned float4 layout", channels);
    const int64_t total = pixels * channels;
    const int grid = (int)std::max<int64_t>((channels + 255) / 256,
                                            std::min<int64_t>((total + 255) / 256, (int64_t)sms * 8));
    column_sums_scalar_kernel<<<grid, 256, 0, st>>>(d_x, total, channels, d_out);
  }
  TFC_LAUNCHED("column_sums_kernel");
  return TFC_OK;
}

extern "C" int tfc_adam_step(float* d_param, const float* d_grad, float* d_m, float* d_v,
                             int64_t n, float lr, float beta1, float beta2, float eps,
                             int64_t step, float grad_scale, void* stream) {
  if (n < 0 || step < 1) return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "bad Adam arguments");
  if (n == 0) return TFC_OK;
  // tf.train.AdamOptimizer: lr_t = lr * sqrt(1 - beta2^t) / (1 - beta1^t)
  const double lr_t = (double)lr * std::sqrt(1.0 - std::pow((double)beta2, (double)step)) /
                      (1.0 - std::pow((double)beta1, (double)step));
  const int grid = (int)std::min<int64_t>((n + 255) / 256, (int64_t)tfc::num_sms() * 8);
  adam_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      d_param, d_grad, d_m, d_v, n, (float)lr_t, beta1, beta2, eps, grad_scale);
  TFC_LAUNCHED("adam_kernel");
  return TFC_OK;
}
