// C-ABI entry points of the transforms: validation + dispatch between the tcgen05 path and
// the CUDA-core path.
#include <algorithm>

#include "common.cuh"
#include <cuda_bf16.h>

#include "conv_common.cuh"

namespace {

// ---- strided correlation over an image whose size is not a multiple of the stride -----------
// The implicit-GEMM kernel views the input as (s*C, W/s, s, H/s, N) for its TMA boxes, which
// needs H and W to be multiples of s (1080p: 135 -> 68 rows in the fourth analysis layer).  The
// missing rows / columns are exactly the zero padding of same_zeros, so the image is copied
// once into the workspace with its size rounded up (one pass over the input, ~0.1 ms against
// the 28 ms of the CUDA-core kernel) and convolved there.
__global__ void __launch_bounds__(256)
pad_image_kernel(const uint4* __restrict__ in, int n, int h, int w, int h2, int w2, int vec_per_px,
                 uint4* __restrict__ out) {
  const int64_t total = (int64_t)n * h2 * w2 * vec_per_px;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int v = (int)(i % vec_per_px);
    int64_t px = i / vec_per_px;
    const int x = (int)(px % w2);
    px /= w2;
    const int y = (int)(px % h2);
    const int b = (int)(px / h2);
    uint4 val = make_uint4(0, 0, 0, 0);
    if (y < h && x < w) val = in[(((int64_t)b * h + y) * w + x) * vec_per_px + v];
    out[i] = val;
  }
}

bool needs_pad(const tfc_conv_desc& d) {
  return d.sd > 1 && d.su == 1 && (d.h % d.sd || d.w % d.sd) && d.in_dtype == TFC_BF16 &&
         d.cin % 8 == 0;
}

tfc_conv_desc padded_desc(const tfc_conv_desc& d) {
  tfc_conv_desc e = d;
  e.h = (d.h + d.sd - 1) / d.sd * d.sd;
  e.w = (d.w + d.sd - 1) / d.sd * d.sd;
  return e;
}

size_t align256(size_t x) { return (x + 255) & ~size_t(255); }

// ---- thin-output transposed convolution as a sub-pixel convolution ---------------------------
// A transposed convolution with stride s writes output pixel (s*i + py, s*j + px) from the input
// pixels around (i, j) through the kernel taps a = (pt - py) mod s, + s, + 2s, ... only: all
// s*s output phases of one input position together are an ordinary stride-1 correlation with
// ceil-ish(k / s) taps per dimension and s*s*cout output channels, followed by a depth-to-space
// permutation.  For the last synthesis layer of bls2017 (9x9, stride 4, 128 -> 3; SURVEY a21:
// examples/bls2017.py:122-127) that is a 3x3 correlation 128 -> 48 on the 64 x 64 grid: one
// implicit GEMM with N = 48 instead of 16 phase GEMMs with N = 3 padded to 16 (0.70 -> see
// profiles/ for the training step), and kh*kw*cout = 243 columns is more than the scatter kernel
// (conv_upthin.cu, <= 128) takes.
struct SubPixel {
  int kh2, kw2;             // taps of the equivalent correlation
  int dmin_h, dmin_w;       // first input offset (the correlation's padding is -dmin)
  int cout2;                // su * su * cout
};

inline int floor_div_i(int a, int b) { return (a >= 0) ? a / b : -((-a + b - 1) / b); }
inline int ceil_div_i(int a, int b) { return -floor_div_i(-a, b); }

SubPixel sub_pixel_geometry(const tfc_conv_desc& d) {
  SubPixel g;
  g.dmin_h = ceil_div_i(-d.pad_top, d.su);
  g.dmin_w = ceil_div_i(-d.pad_left, d.su);
  g.kh2 = floor_div_i(d.su - 1 + d.kh - 1 - d.pad_top, d.su) - g.dmin_h + 1;
  g.kw2 = floor_div_i(d.su - 1 + d.kw - 1 - d.pad_left, d.su) - g.dmin_w + 1;
  g.cout2 = d.su * d.su * d.cout;
  return g;
}

tfc_conv_desc sub_pixel_desc(const tfc_conv_desc& d) {
  const SubPixel g = sub_pixel_geometry(d);
  tfc_conv_desc e = d;
  e.ho = d.h; e.wo = d.w; e.cout = g.cout2;
  e.kh = g.kh2; e.kw = g.kw2; e.sd = 1; e.su = 1;
  e.pad_top = -g.dmin_h; e.pad_left = -g.dmin_w;
  e.act = TFC_ACT_NONE; e.out_dtype = TFC_F32; e.impl = 2; e.flags = d.flags;
  return e;
}

bool sub_pixel_supported(const tfc_conv_desc& d) {
  if (d.su < 2 || d.sd != 1 || d.in_dtype != TFC_BF16) return false;
  if (d.act != TFC_ACT_NONE && d.act != TFC_ACT_RELU) return false;
  if (d.ho != d.h * d.su || d.wo != d.w * d.su) return false;
  if (d.cout * d.su * d.su > 64 || d.cout > 4) return false;
  const tfc_conv_desc e = sub_pixel_desc(d);
  if (e.kh < 1 || e.kw < 1 || e.pad_top < 0 || e.pad_left < 0) return false;
  return tfc::conv_umma_supported(e, nullptr);
}

// w [kh, kw, cin, cout] -> w2 [kh2, kw2, cin, (py*su + px)*cout + o]
__global__ void __launch_bounds__(256)
sub_pixel_weights_kernel(const float* __restrict__ w, int kh, int kw, int cin, int cout, int su,
                         int pt, int pl, int kh2, int kw2, int dmin_h, int dmin_w,
                         float* __restrict__ w2) {
  const int cout2 = su * su * cout;
  const int64_t total = (int64_t)kh2 * kw2 * cin * cout2;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int oc = (int)(i % cout2);
    int64_t r = i / cout2;
    const int c = (int)(r % cin);
    r /= cin;
    const int tb = (int)(r % kw2), ta = (int)(r / kw2);
    const int o = oc % cout, ph = oc / cout, px = ph % su, py = ph / su;
    const int a = su * (ta + dmin_h) - py + pt, b = su * (tb + dmin_w) - px + pl;
    float v = 0.f;
    if (a >= 0 && a < kh && b >= 0 && b < kw) v = w[(((int64_t)a * kw + b) * cin + c) * cout + o];
    w2[i] = v;
  }
}

// depth-to-space + bias + ReLU + output dtype: one thread per output row segment of su pixels
// (su * cout contiguous values in y2 and in the output); consecutive threads walk an output row
__global__ void __launch_bounds__(256)
sub_pixel_shuffle_kernel(const float* __restrict__ y2, int64_t rows_in, int w, int su, int cout,
                         const float* __restrict__ bias, int relu, int out_dtype,
                         void* __restrict__ out) {
  const int row_elems = su * cout;
  const int64_t total = rows_in * su * w;               // rows_in = n * h
  for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total;
       t += (int64_t)gridDim.x * blockDim.x) {
    const int64_t j = t % w, r = t / w;
    const int py = (int)(r % su);
    const int64_t ni = r / su;
    const float* src = y2 + ((ni * w + j) * su + py) * row_elems;
    const int64_t dst = (r * w + j) * row_elems;         // r = ni * su + py is the output row
    for (int e = 0; e < row_elems; ++e) {
      float v = src[e] + (bias ? bias[e % cout] : 0.f);
      if (relu) v = fmaxf(v, 0.f);
      if (out_dtype == TFC_F32) static_cast<float*>(out)[dst + e] = v;
      else if (out_dtype == TFC_BF16) static_cast<__nv_bfloat16*>(out)[dst + e] = __float2bfloat16_rn(v);
      else static_cast<uint8_t*>(out)[dst + e] = (uint8_t)fminf(fmaxf(rintf(v * 255.f), 0.f), 255.f);
    }
  }
}

size_t sub_pixel_workspace(const tfc_conv_desc& d) {
  const tfc_conv_desc e = sub_pixel_desc(d);
  return align256(tfc::conv_umma_workspace(e)) +
         align256((size_t)e.kh * e.kw * e.cin * e.cout * 4) +
         align256((size_t)e.n * e.ho * e.wo * e.cout * 4) + 256;
}

int validate(const tfc_conv_desc* d) {
  if (!d) return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "null conv descriptor");
  if (d->n < 0 || d->h <= 0 || d->w <= 0 || d->cin <= 0 || d->ho <= 0 || d->wo <= 0 ||
      d->cout <= 0 || d->kh <= 0 || d->kw <= 0 || d->sd <= 0 || d->su <= 0)
    return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "bad conv geometry");
  if (d->act < TFC_ACT_NONE || d->act > TFC_ACT_IGDN)
    return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "bad activation %d", d->act);
  if (d->in_dtype < TFC_F32 || d->in_dtype > TFC_U8 || d->out_dtype < TFC_F32 ||
      d->out_dtype > TFC_U8)
    return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "bad dtype");
  return TFC_OK;
}

// 0: CUDA-core path, 1: tcgen05 implicit GEMM, 2: space-to-depth + tcgen05, 3: scatter GEMM,
// 4: thin-input im2col + tcgen05, 5: zero-extended copy + tcgen05 implicit GEMM,
// 6: thin-output transposed layer as a sub-pixel correlation + depth-to-space
int pick_path(const tfc_conv_desc& d, const char** why) {
  if (d.impl == 1) return 0;
  if (d.impl == 3) return tfc::conv_upthin_supported(d, why) ? 3 : 0;
  if (d.impl == 4) return tfc::conv_thin_supported(d, why) ? 4 : 0;
  if (d.impl == 0 && tfc::conv_upthin_supported(d, nullptr)) return 3;
  if (d.impl == 0 && tfc::conv_thin_supported(d, nullptr)) return 4;
  if (d.impl == 0 && sub_pixel_supported(d)) return 6;
  if (tfc::conv_umma_supported(d, why)) return 1;
  if (needs_pad(d) && tfc::conv_umma_supported(padded_desc(d), nullptr)) return 5;
  const char* why2 = nullptr;
  if (tfc::conv_s2d_supported(d, &why2)) return 2;
  return 0;
}

}  // namespace

extern "C" size_t tfc_signal_conv2d_workspace(const tfc_conv_desc* desc) {
  if (validate(desc)) return 0;
  size_t need = 256;
  const char* why = nullptr;
  const int path = pick_path(*desc, &why);
  if (path == 0) need = std::max(need, tfc::conv_direct_workspace(*desc));
  if (path == 1) need = std::max(need, tfc::conv_umma_workspace(*desc));
  if (path == 2) need = std::max(need, tfc::conv_s2d_workspace(*desc));
  if (path == 3) need = std::max(need, tfc::conv_upthin_workspace(*desc));
  if (path == 4) need = std::max(need, tfc::conv_thin_workspace(*desc));
  if (path == 5) {
    const tfc_conv_desc e = padded_desc(*desc);
    need = std::max(need, align256(tfc::conv_umma_workspace(e)) +
                              (size_t)e.n * e.h * e.w * e.cin * 2 + 256);
  }
  if (path == 6) need = std::max(need, sub_pixel_workspace(*desc));
  return need;
}

extern "C" int tfc_signal_conv2d_path(const tfc_conv_desc* desc) {
  if (validate(desc)) return -1;
  const char* why = nullptr;
  return pick_path(*desc, &why);
}

extern "C" int tfc_signal_conv2d(const tfc_conv_desc* desc, const void* d_in, const float* d_w,
                                 const float* d_bias, const float* d_beta, const float* d_gamma,
                                 void* d_out, void* d_workspace, size_t workspace_bytes,
                                 void* stream) {
  int rc = validate(desc);
  if (rc) return rc;
  const bool gdn = desc->act == TFC_ACT_GDN || desc->act == TFC_ACT_IGDN;
  if (gdn && (!d_beta || !d_gamma))
    return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "GDN activation needs beta and gamma");
  if (desc->n == 0) return TFC_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const char* why = "unknown";
  const int path = pick_path(*desc, &why);
  if (desc->impl >= 2 && path == 0)
    return tfc::fail(TFC_ERR_UNIMPLEMENTED, "tcgen05 conv path does not support this shape: %s",
                     why);
  if ((desc->in_dtype == TFC_U8 && path != 4) || (desc->out_dtype == TFC_U8 && path != 3 && path != 6))
    return tfc::fail(TFC_ERR_UNIMPLEMENTED,
                     "uint8 images are read by the thin-input kernel and written by the "
                     "thin-output kernel only (this shape takes kernel path %d)", path);
  if (path == 1)
    return tfc::conv_umma(*desc, d_in, d_w, d_bias, d_beta, d_gamma, d_out, d_workspace,
                          workspace_bytes, st);
  if (path == 3)
    return tfc::conv_upthin(*desc, d_in, d_w, d_bias, d_out, d_workspace, workspace_bytes, st);
  if (path == 4)
    return tfc::conv_thin(*desc, d_in, d_w, d_bias, d_beta, d_gamma, d_out, d_workspace,
                          workspace_bytes, st);
  if (path == 5) {
    // workspace: [implicit-GEMM workspace (weights first: offsets independent of the image)]
    //            [zero-extended image]
    const tfc_conv_desc e = padded_desc(*desc);
    const size_t inner = align256(tfc::conv_umma_workspace(e));
    const size_t img = (size_t)e.n * e.h * e.w * e.cin * 2;
    if (d_workspace == nullptr || workspace_bytes < inner + img)
      return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "conv workspace too small: need %zu bytes",
                       inner + img + 256);
    if (reinterpret_cast<uintptr_t>(d_in) & 15)
      return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "conv buffers must be 16-byte aligned");
    uint4* padded = reinterpret_cast<uint4*>(static_cast<uint8_t*>(d_workspace) + inner);
    const int vec = desc->cin / 8;
    const int64_t total = (int64_t)e.n * e.h * e.w * vec;
    const int grid = (int)std::min<int64_t>((total + 255) / 256, (int64_t)tfc::num_sms() * 16);
    pad_image_kernel<<<grid, 256, 0, st>>>(static_cast<const uint4*>(d_in), desc->n, desc->h,
                                           desc->w, e.h, e.w, vec, padded);
    TFC_LAUNCHED("pad_image_kernel");
    return tfc::conv_umma(e, padded, d_w, d_bias, d_beta, d_gamma, d_out, d_workspace, inner, st);
  }
  if (path == 6) {
    // workspace: [implicit-GEMM workspace (repacked weights first)] [w2 float32] [y2 float32]
    const tfc_conv_desc e = sub_pixel_desc(*desc);
    const SubPixel g = sub_pixel_geometry(*desc);
    const size_t inner = align256(tfc::conv_umma_workspace(e));
    const size_t wbytes = align256((size_t)e.kh * e.kw * e.cin * e.cout * 4);
    if (d_workspace == nullptr || workspace_bytes < sub_pixel_workspace(*desc) - 256)
      return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "conv workspace too small: need %zu bytes",
                       sub_pixel_workspace(*desc));
    float* w2 = reinterpret_cast<float*>(static_cast<uint8_t*>(d_workspace) + inner);
    float* y2 = reinterpret_cast<float*>(static_cast<uint8_t*>(d_workspace) + inner + wbytes);
    if (!(desc->flags & TFC_CONV_WEIGHTS_STAGED)) {
      const int64_t total = (int64_t)e.kh * e.kw * e.cin * e.cout;
      sub_pixel_weights_kernel<<<(int)std::min<int64_t>((total + 255) / 256, 4096), 256, 0, st>>>(
          d_w, desc->kh, desc->kw, desc->cin, desc->cout, desc->su, desc->pad_top, desc->pad_left,
          g.kh2, g.kw2, g.dmin_h, g.dmin_w, w2);
      TFC_LAUNCHED("sub_pixel_weights_kernel");
    }
    rc = tfc::conv_umma(e, d_in, w2, nullptr, nullptr, nullptr, y2, d_workspace, inner, st);
    if (rc) return rc;
    const int64_t rows_in = (int64_t)desc->n * desc->h;
    const int64_t threads = rows_in * desc->su * desc->w;
    sub_pixel_shuffle_kernel<<<(int)std::min<int64_t>((threads + 255) / 256,
                                                      (int64_t)tfc::num_sms() * 16),
                               256, 0, st>>>(y2, rows_in, desc->w, desc->su, desc->cout, d_bias,
                                             desc->act == TFC_ACT_RELU, desc->out_dtype, d_out);
    TFC_LAUNCHED("sub_pixel_shuffle_kernel");
    return TFC_OK;
  }
  if (path == 2)
    return tfc::conv_s2d(*desc, d_in, d_w, d_bias, d_beta, d_gamma, d_out, d_workspace,
                         workspace_bytes, st);
  return tfc::conv_direct(*desc, d_in, d_w, d_bias, d_beta, d_gamma, d_out, d_workspace,
                          workspace_bytes, st);
}

extern "C" int tfc_gdn(const void* d_x, int64_t pixels, int32_t channels, const float* d_beta,
                       const float* d_gamma, int inverse, int in_dtype, int out_dtype,
                       void* d_out, void* stream) {
  if (pixels <= 0) return TFC_OK;
  return tfc::gdn_direct(d_x, pixels, channels, d_beta, d_gamma, inverse != 0, in_dtype,
                         out_dtype, d_out, static_cast<cudaStream_t>(stream));
}
