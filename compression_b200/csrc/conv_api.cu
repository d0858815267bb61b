// C-ABI entry points of the transforms: validation + dispatch between the tcgen05 path and
// the CUDA-core path.
#include <algorithm>

#include "common.cuh"
#include "conv_common.cuh"

namespace {

int validate(const tfc_conv_desc* d) {
  if (!d) return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "null conv descriptor");
  if (d->n < 0 || d->h <= 0 || d->w <= 0 || d->cin <= 0 || d->ho <= 0 || d->wo <= 0 ||
      d->cout <= 0 || d->kh <= 0 || d->kw <= 0 || d->sd <= 0 || d->su <= 0)
    return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "bad conv geometry");
  if (d->act < TFC_ACT_NONE || d->act > TFC_ACT_IGDN)
    return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "bad activation %d", d->act);
  if ((d->in_dtype != TFC_F32 && d->in_dtype != TFC_BF16) ||
      (d->out_dtype != TFC_F32 && d->out_dtype != TFC_BF16))
    return tfc::fail(TFC_ERR_INVALID_ARGUMENT, "bad dtype");
  return TFC_OK;
}

// 0: CUDA-core path, 1: tcgen05 implicit GEMM, 2: space-to-depth + tcgen05, 3: scatter GEMM,
// 4: thin-input im2col + tcgen05
int pick_path(const tfc_conv_desc& d, const char** why) {
  if (d.impl == 1) return 0;
  if (d.impl == 3) return tfc::conv_upthin_supported(d, why) ? 3 : 0;
  if (d.impl == 4) return tfc::conv_thin_supported(d, why) ? 4 : 0;
  if (d.impl == 0 && tfc::conv_upthin_supported(d, nullptr)) return 3;
  if (d.impl == 0 && tfc::conv_thin_supported(d, nullptr)) return 4;
  if (tfc::conv_umma_supported(d, why)) return 1;
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
  return need;
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
  if (path == 1)
    return tfc::conv_umma(*desc, d_in, d_w, d_bias, d_beta, d_gamma, d_out, d_workspace,
                          workspace_bytes, st);
  if (path == 3)
    return tfc::conv_upthin(*desc, d_in, d_w, d_bias, d_out, d_workspace, workspace_bytes, st);
  if (path == 4)
    return tfc::conv_thin(*desc, d_in, d_w, d_bias, d_beta, d_gamma, d_out, d_workspace,
                          workspace_bytes, st);
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
