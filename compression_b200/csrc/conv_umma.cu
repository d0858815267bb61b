// tcgen05 / TMEM / TMA implicit-GEMM SignalConv2D with fused bias + GDN/IGDN epilogue.
// (stub until the kernel lands: reports every shape as unsupported so the dispatcher uses the
// CUDA-core path)
#include "common.cuh"
#include "conv_common.cuh"

namespace tfc {
bool conv_umma_supported(const tfc_conv_desc&, const char** why) {
  if (why) *why = "tcgen05 kernel not built yet";
  return false;
}
size_t conv_umma_workspace(const tfc_conv_desc&) { return 0; }
int conv_umma(const tfc_conv_desc&, const void*, const float*, const float*, const float*,
              const float*, void*, void*, size_t, cudaStream_t) {
  return fail(TFC_ERR_UNIMPLEMENTED, "tcgen05 conv path not built yet");
}
}  // namespace tfc
