// Internal interfaces between the conv dispatcher (conv_api.cu) and its implementations.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

#include "tfc_b200.h"

namespace tfc {

// CUDA-core tiled path (conv_direct.cu): any shape
size_t conv_direct_workspace(const tfc_conv_desc& d);
int conv_direct(const tfc_conv_desc& d, const void* in, const float* w, const float* bias,
                const float* beta, const float* gamma, void* out, void* ws, size_t ws_bytes,
                cudaStream_t st);
int gdn_direct(const void* x, int64_t pixels, int channels, const float* beta,
               const float* gamma, bool inverse, int in_dtype, int out_dtype, void* out,
               cudaStream_t st);

// tcgen05 implicit-GEMM path (conv_umma.cu): bf16 operands, fp32 accumulation in TMEM
bool conv_umma_supported(const tfc_conv_desc& d, const char** why);
size_t conv_umma_workspace(const tfc_conv_desc& d);
int conv_umma(const tfc_conv_desc& d, const void* in, const float* w, const float* bias,
              const float* beta, const float* gamma, void* out, void* ws, size_t ws_bytes,
              cudaStream_t st);

// SWIZZLE_128B bf16 tiled tensor map (cuTensorMapEncodeTiled through the runtime's driver
// entry point); defined in conv_umma.cu
int make_tensor_map_bf16(CUtensorMap* m, void* base, int rank, const cuuint64_t* dims,
                         const cuuint64_t* strides_bytes, const cuuint32_t* box);

// un-swizzled float32 / bf16 / uint8 (tfc_dtype) tiled tensor map (image patches of conv_thin.cu)
int make_tensor_map_plain(CUtensorMap* m, int dtype, void* base, int rank, const cuuint64_t* dims,
                          const cuuint64_t* strides_bytes, const cuuint32_t* box);

// 128-pixel tile shape (th x tw) that wastes the fewest pixels on an hs x ws output grid
void conv_pick_tile(int hs, int ws, int* th, int* tw);

// thin-input layers (3 image channels) with a software im2col feeding tcgen05 (conv_thin.cu)
bool conv_thin_supported(const tfc_conv_desc& d, const char** why);
size_t conv_thin_workspace(const tfc_conv_desc& d);
int conv_thin(const tfc_conv_desc& d, const void* in, const float* w, const float* bias,
              const float* beta, const float* gamma, void* out, void* ws, size_t ws_bytes,
              cudaStream_t st);

// thin-output transposed layers as a scatter GEMM (conv_upthin.cu)
bool conv_upthin_supported(const tfc_conv_desc& d, const char** why);
size_t conv_upthin_workspace(const tfc_conv_desc& d);
int conv_upthin(const tfc_conv_desc& d, const void* in, const float* w, const float* bias,
                void* out, void* ws, size_t ws_bytes, cudaStream_t st);

// filter gradient on tcgen05 with MN-major operands (conv_wgrad_umma.cu): both channel counts
// multiples of 64, the M side exactly 128; `dw` must be zeroed (slices add atomically)
bool conv_wgrad_umma_supported(const tfc_conv_desc& d);
// the same for the layers with 3 channels on one side (conv_thin.cu: software im2col as A operand)
bool conv_thin_wgrad_supported(const tfc_conv_desc& d, const void* x_bf16, const void* dy_bf16);
int conv_thin_wgrad(const tfc_conv_desc& d, const void* x_bf16, const void* dy_bf16, float* dw,
                    cudaStream_t st);
int conv_wgrad_umma(const tfc_conv_desc& d, const void* x_bf16, const void* dy_bf16, float* dw,
                    cudaStream_t st);

// thin-input layers via space-to-depth + conv_umma (conv_s2d.cu)
bool conv_s2d_supported(const tfc_conv_desc& d, const char** why);
size_t conv_s2d_workspace(const tfc_conv_desc& d);
int conv_s2d(const tfc_conv_desc& d, const void* in, const float* w, const float* bias,
             const float* beta, const float* gamma, void* out, void* ws, size_t ws_bytes,
             cudaStream_t st);

}  // namespace tfc
