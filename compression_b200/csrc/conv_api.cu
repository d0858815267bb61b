// C-ABI entry points of the transforms: validation + dispatch between the tcgen05 path and
// the CUDA-core path.
#include <algorithm>

#include "common.cuh"
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
// 4: thin-input im2col + tcgen05, 5: zero-extended copy + tcgen05 implicit GEMM
int pick_path(const tfc_conv_desc& d, const char** why) {
  if (d.impl == 1) return 0;
  if (d.impl == 3) return tfc::conv_upthin_supported(d, why) ? 3 : 0;
  if (d.impl == 4) return tfc::conv_thin_supported(d, why) ? 4 : 0;
  if (d.impl == 0 && tfc::conv_upthin_supported(d, nullptr)) return 3;
  if (d.impl == 0 && tfc::conv_thin_supported(d, nullptr)) return 4;
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
  if ((desc->in_dtype == TFC_U8 && path != 4) || (desc->out_dtype == TFC_U8 && path != 3))
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
