// Thin-input SignalConv2D (the first analysis layer: 3 input channels, 5x5 stride 2 or 9x9
// stride 4) on the tensor cores.
//
// A stride-s correlation over a cin-channel image is rewritten as a stride-1 correlation over
// the space-to-depth image xs[n, h/s, w/s, (hp, wp, c)] — s*s*cin channels, zero-padded to 64 so
// that one kernel tap is exactly one 128-byte K-block of the tcgen05 implicit GEMM in
// conv_umma.cu — with the kernel taps regrouped accordingly:
//     out[i, j] = sum_{a,b,c} in[s*i + a - pt, s*j + b - pl, c] * w[a, b, c]
//              = sum_{qa,qb} sum_{hp,wp,c} xs[i + qa, j + qb, (hp, wp, c)] * w[s*qa + hp + pt, ...]
// (a - pt = s*qa + hp).  A 5x5/s2 kernel becomes 3x3, a 9x9/s4 kernel becomes 3x3 as well.
// Same reference semantics as conv_umma.cu (signal_conv.py:487-524); the bias / GDN epilogue is
// that kernel's.
#include <cuda_bf16.h>

#include "common.cuh"
#include "conv_common.cuh"

namespace {

constexpr int kPadC = 64;

inline int floor_div(int a, int b) { return (a >= 0) ? a / b : -((-a + b - 1) / b); }

struct Regroup {
  int qmin_h, qmin_w, kh2, kw2;   // new kernel support and its (negative) first offset
};

Regroup regroup(const tfc_conv_desc& d) {
  Regroup r;
  r.qmin_h = floor_div(0 - d.pad_top, d.sd);
  r.qmin_w = floor_div(0 - d.pad_left, d.sd);
  r.kh2 = floor_div(d.kh - 1 - d.pad_top, d.sd) - r.qmin_h + 1;
  r.kw2 = floor_div(d.kw - 1 - d.pad_left, d.sd) - r.qmin_w + 1;
  return r;
}

tfc_conv_desc inner_desc(const tfc_conv_desc& d) {
  const Regroup r = regroup(d);
  tfc_conv_desc e = d;
  e.h = d.h / d.sd; e.w = d.w / d.sd; e.cin = kPadC;
  e.kh = r.kh2; e.kw = r.kw2; e.sd = 1; e.su = 1;
  e.pad_top = -r.qmin_h; e.pad_left = -r.qmin_w;
  e.in_dtype = TFC_BF16;
  e.impl = 2;
  return e;
}

template <typename InT>
__global__ void __launch_bounds__(256)
space_to_depth_kernel(const InT* __restrict__ in, int n, int h, int w, int cin, int s,
                      __nv_bfloat16* __restrict__ xs) {
  const int h2 = h / s, w2 = w / s;
  const int64_t pixels = (int64_t)n * h2 * w2;
  const int real = s * s * cin;
  for (int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; p < pixels;
       p += (int64_t)gridDim.x * blockDim.x) {
    const int j2 = (int)(p % w2);
    const int i2 = (int)((p / w2) % h2);
    const int b = (int)(p / ((int64_t)w2 * h2));
    const InT* src = in + (((int64_t)b * h + (int64_t)i2 * s) * w + (int64_t)j2 * s) * cin;
    uint4* dst = reinterpret_cast<uint4*>(xs + p * kPadC);
#pragma unroll
    for (int g = 0; g < kPadC / 8; ++g) {
      __nv_bfloat16 v[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const int ch = g * 8 + e;
        float f = 0.f;
        if (ch < real) {
          const int hp = ch / (s * cin);
          const int rem = ch - hp * s * cin;      // wp * cin + c: contiguous in the source row
          f = (float)src[(int64_t)hp * w * cin + rem];
        }
        v[e] = __float2bfloat16_rn(f);
      }
      dst[g] = *reinterpret_cast<const uint4*>(v);
    }
  }
}

// w [kh, kw, cin, cout] -> w2 [kh2, kw2, 64, cout]
__global__ void regroup_weights_kernel(const float* __restrict__ w, int kh, int kw, int cin,
                                       int cout, int s, int pt, int pl, int qmin_h, int qmin_w,
                                       int kh2, int kw2, float* __restrict__ w2) {
  const int64_t total = (int64_t)kh2 * kw2 * kPadC * cout;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int o = (int)(i % cout);
    const int ch = (int)((i / cout) % kPadC);
    const int qb = (int)((i / ((int64_t)cout * kPadC)) % kw2);
    const int qa = (int)(i / ((int64_t)cout * kPadC * kw2));
    float v = 0.f;
    if (ch < s * s * cin) {
      const int hp = ch / (s * cin);
      const int rem = ch - hp * s * cin;
      const int wp = rem / cin, c = rem - wp * cin;
      const int a = s * (qa + qmin_h) + hp + pt;
      const int b = s * (qb + qmin_w) + wp + pl;
      if (a >= 0 && a < kh && b >= 0 && b < kw) v = w[(((int64_t)a * kw + b) * cin + c) * cout + o];
    }
    w2[i] = v;
  }
}

size_t align256(size_t x) { return (x + 255) & ~size_t(255); }

}  // namespace

namespace tfc {

bool conv_s2d_supported(const tfc_conv_desc& d, const char** why) {
  const char* dummy;
  if (!why) why = &dummy;
  if (d.su != 1) { *why = "space-to-depth path needs strides_up == 1"; return false; }
  if (d.cin % 64 == 0) { *why = "cin is already a multiple of 64"; return false; }
  if (d.sd * d.sd * d.cin > kPadC) { *why = "strides_down^2 * cin must be <= 64"; return false; }
  if (d.h % d.sd || d.w % d.sd) { *why = "input size must be a multiple of strides_down"; return false; }
  const tfc_conv_desc e = inner_desc(d);
  return conv_umma_supported(e, why);
}

size_t conv_s2d_workspace(const tfc_conv_desc& d) {
  const tfc_conv_desc e = inner_desc(d);
  const size_t xs = align256((size_t)e.n * e.h * e.w * kPadC * 2);
  const size_t w2 = align256((size_t)e.kh * e.kw * kPadC * d.cout * 4);
  return xs + w2 + align256(conv_umma_workspace(e)) + 256;
}

int conv_s2d(const tfc_conv_desc& d, const void* in, const float* w, const float* bias,
             const float* beta, const float* gamma, void* out, void* ws, size_t ws_bytes,
             cudaStream_t st) {
  if (ws == nullptr || ws_bytes < conv_s2d_workspace(d))
    return fail(TFC_ERR_INVALID_ARGUMENT, "conv workspace too small: need %zu bytes",
                conv_s2d_workspace(d));
  const tfc_conv_desc e = inner_desc(d);
  const Regroup r = regroup(d);
  // layout: [regrouped fp32 weights][inner conv workspace (bf16 weights, gamma)][xs] — the
  // weight areas come first so that their offsets do not depend on the image size
  uint8_t* base = static_cast<uint8_t*>(ws);
  const size_t xs_bytes = align256((size_t)e.n * e.h * e.w * kPadC * 2);
  const size_t w2_bytes = align256((size_t)e.kh * e.kw * kPadC * d.cout * 4);
  const size_t inner_bytes = align256(conv_umma_workspace(e));
  float* w2 = reinterpret_cast<float*>(base);
  void* inner_ws = base + w2_bytes;
  __nv_bfloat16* xs = reinterpret_cast<__nv_bfloat16*>(base + w2_bytes + inner_bytes);
  (void)xs_bytes;
  const int64_t pixels = (int64_t)e.n * e.h * e.w;
  const int grid = (int)std::min<int64_t>((pixels + 255) / 256, (int64_t)num_sms() * 16);
  if (d.in_dtype == TFC_F32)
    space_to_depth_kernel<float><<<grid, 256, 0, st>>>(static_cast<const float*>(in), d.n, d.h,
                                                       d.w, d.cin, d.sd, xs);
  else
    space_to_depth_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>(
        static_cast<const __nv_bfloat16*>(in), d.n, d.h, d.w, d.cin, d.sd, xs);
  TFC_LAUNCHED("space_to_depth_kernel");
  if (!(d.flags & TFC_CONV_WEIGHTS_STAGED)) {
    const int64_t wtotal = (int64_t)e.kh * e.kw * kPadC * d.cout;
    regroup_weights_kernel<<<(int)std::min<int64_t>((wtotal + 255) / 256, 1024), 256, 0, st>>>(
        w, d.kh, d.kw, d.cin, d.cout, d.sd, d.pad_top, d.pad_left, r.qmin_h, r.qmin_w, e.kh, e.kw,
        w2);
    TFC_LAUNCHED("regroup_weights_kernel");
  }
  return conv_umma(e, xs, w2, bias, beta, gamma, out, inner_ws, inner_bytes, st);
}

}  // namespace tfc
