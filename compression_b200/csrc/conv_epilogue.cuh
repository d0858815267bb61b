// Epilogues of the tcgen05 convolution kernels (conv_umma.cu, conv_thin.cu): one thread per
// TMEM lane / tile pixel.  bias (+ReLU), or the fused GDN / IGDN of gdn.py:183-202:
//   phase 1: x = D1 + bias kept as packed bf16 in registers, x^2 -> K-major SW128 bf16 tile(s) in
//            shared memory (operand A of the second GEMM, which overwrites D1 in place)
//   phase 2: y = x * rsqrt(beta + D2)   (IGDN: * sqrt)
#pragma once
#include <cuda_bf16.h>

#include "umma_ptx.cuh"

namespace tfc {
namespace epi {

using namespace tfc::ptx;

constexpr int kABytes = 128 * 64 * 2;   // one 128-row x 64-channel bf16 operand tile

__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}

__device__ __forceinline__ float rsqrt_approx(float x) {
  float y;
  asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float sqrt_approx(float x) {
  float y;
  asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// 256-bit global store (STG.256): one full 32-byte sector per lane, so that a warp whose lanes
// write different pixels still moves whole sectors.  `p` must be 32-byte aligned.
__device__ __forceinline__ void st_global_v8(void* p, uint32_t a0, uint32_t a1, uint32_t a2,
                                             uint32_t a3, uint32_t a4, uint32_t a5, uint32_t a6,
                                             uint32_t a7) {
  asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(p), "r"(a0),
               "r"(a1), "r"(a2), "r"(a3), "r"(a4), "r"(a5), "r"(a6), "r"(a7)
               : "memory");
}

// 32 consecutive channels of one pixel.  `wide`: the destination is 32-byte aligned.
__device__ __forceinline__ void store_chunk(void* out, bool out_f32, bool wide, int64_t elem,
                                            const float (&y)[32]) {
  if (out_f32) {
    float* dst = static_cast<float*>(out) + elem;
    if (wide) {
#pragma unroll
      for (int q = 0; q < 4; ++q)
        st_global_v8(dst + 8 * q, __float_as_uint(y[8 * q]), __float_as_uint(y[8 * q + 1]),
                     __float_as_uint(y[8 * q + 2]), __float_as_uint(y[8 * q + 3]),
                     __float_as_uint(y[8 * q + 4]), __float_as_uint(y[8 * q + 5]),
                     __float_as_uint(y[8 * q + 6]), __float_as_uint(y[8 * q + 7]));
    } else {
#pragma unroll
      for (int q = 0; q < 8; ++q)
        reinterpret_cast<float4*>(dst)[q] = make_float4(y[4 * q], y[4 * q + 1], y[4 * q + 2], y[4 * q + 3]);
    }
  } else {
    __nv_bfloat16* dst = static_cast<__nv_bfloat16*>(out) + elem;
    uint32_t w[16];
#pragma unroll
    for (int q = 0; q < 16; ++q) w[q] = pack_bf16(y[2 * q], y[2 * q + 1]);
    if (wide) {
      st_global_v8(dst, w[0], w[1], w[2], w[3], w[4], w[5], w[6], w[7]);
      st_global_v8(dst + 16, w[8], w[9], w[10], w[11], w[12], w[13], w[14], w[15]);
    } else {
#pragma unroll
      for (int q = 0; q < 4; ++q)
        reinterpret_cast<uint4*>(dst)[q] = make_uint4(w[4 * q], w[4 * q + 1], w[4 * q + 2], w[4 * q + 3]);
    }
  }
}

// Eight epilogue warps: warp pair (w, w + 4) shares a TMEM lane quadrant (a warp may only read
// lanes 32*(warp%4)...), the first taking the lower half of the channels and the second the
// upper half — two warps per scheduler to hide the tcgen05.ld / MUFU latencies, and half the
// register stash per thread.
constexpr int kEpilogueWarps = 8;
constexpr int kEpilogueThreads = kEpilogueWarps * 32;

// MODE 1 / 2: GDN / IGDN.  This thread handles channels [32*cbase, 32*(cbase + NCH)) of pixel
// `r`.  `dacc`: TMEM address of this warp's lanes of the accumulator, `a2`: the x^2 tile(s),
// `elem0`: pixel * cout.
//
// Output: with `map_out` (bf16, cout a multiple of 64) the finished tile is written back into
// the x^2 buffer — free once the second GEMM has completed, same swizzled layout — and leaves
// as one TMA tile store per 64 channels, issued by the `leader` thread; ragged tiles are clipped
// by the TMA unit.  The leader waits for the store to have read the buffer before the next
// tile's x^2 goes in.  Without `map_out` every thread stores its own channels (store_chunk).
struct TileOrigin { int j, i, n; };   // tensor-map coordinates of the tile's first pixel (n < 0: no store)

// KO: knock-out bits of the bottleneck hunt (profiles/r3f_*): results are WRONG when non-zero
template <int MODE, int NCH, int KO = 0>
__device__ __forceinline__ void gdn_tile(uint32_t dacc, uint8_t* a2, int r, int cbase,
                                         const float* s_bias, const float* s_beta,
                                         uint64_t* a2_full, uint64_t* d2_full, uint32_t it,
                                         bool valid, void* out, bool out_f32, bool wide,
                                         int64_t elem0, const CUtensorMap* map_out,
                                         TileOrigin org, bool leader, int nblocks,
                                         uint32_t a2_full_pair = 0) {
  // a2_full_pair != 0 (CTA-pair MMAs): the "x^2 is in shared memory" arrival goes to the pair
  // leader's barrier at that shared::cluster address instead of this CTA's a2_full
  uint32_t xs[NCH * 16];                           // x = D1 + bias as packed bf16 pairs
  if (map_out != nullptr && it > 0) {
    if (leader) bulk_wait_read();
    named_barrier(1, kEpilogueThreads);
  }
  // phase 1: x -> registers (bf16), x^2 -> bf16 K-major SW128 tile(s) in shared memory
#pragma unroll
  for (int c = 0; c < NCH; ++c) {
    const int cc = cbase + c;
    uint32_t v[32];
    tmem_ld_32x32(dacc + cc * 32, v);
    tmem_ld_wait();
    uint8_t* row = a2 + (cc >> 1) * kABytes + r * 128;
    const float4* bs = reinterpret_cast<const float4*>(s_bias + cc * 32);
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float4 kz = make_float4(0.f, 0.f, 0.f, 0.f);
      const float4 b0 = (KO & 1) ? kz : bs[2 * q], b1 = (KO & 1) ? kz : bs[2 * q + 1];
      const float x0 = __uint_as_float(v[8 * q + 0]) + b0.x;
      const float x1 = __uint_as_float(v[8 * q + 1]) + b0.y;
      const float x2 = __uint_as_float(v[8 * q + 2]) + b0.z;
      const float x3 = __uint_as_float(v[8 * q + 3]) + b0.w;
      const float x4 = __uint_as_float(v[8 * q + 4]) + b1.x;
      const float x5 = __uint_as_float(v[8 * q + 5]) + b1.y;
      const float x6 = __uint_as_float(v[8 * q + 6]) + b1.z;
      const float x7 = __uint_as_float(v[8 * q + 7]) + b1.w;
      xs[c * 16 + 4 * q + 0] = pack_bf16(x0, x1);
      xs[c * 16 + 4 * q + 1] = pack_bf16(x2, x3);
      xs[c * 16 + 4 * q + 2] = pack_bf16(x4, x5);
      xs[c * 16 + 4 * q + 3] = pack_bf16(x6, x7);
      const int physc = ((cc & 1) * 4 + q) ^ (r & 7);
      if (!(KO & 8) || x0 == 12345.f)
      *reinterpret_cast<uint4*>(row + physc * 16) =
          make_uint4(pack_bf16(x0 * x0, x1 * x1), pack_bf16(x2 * x2, x3 * x3),
                     pack_bf16(x4 * x4, x5 * x5), pack_bf16(x6 * x6, x7 * x7));
    }
  }
  tc_fence_before();
  fence_proxy_async_smem();
  if (a2_full_pair) mbar_arrive_cluster(a2_full_pair); else mbar_arrive(a2_full);
  mbar_wait(d2_full, it & 1);
  tc_fence_after();
  // phase 2: y = x * rsqrt(beta + D2)  (IGDN: x * sqrt)
#pragma unroll
  for (int c = 0; c < NCH; ++c) {
    const int cc = cbase + c;
    uint32_t g[32];
    if (KO & 32) {
#pragma unroll
      for (int e = 0; e < 32; ++e) g[e] = xs[c * 16 + (e >> 1)];
    } else {
      tmem_ld_32x32(dacc + cc * 32, g);
      tmem_ld_wait();
    }
    const float4* be = reinterpret_cast<const float4*>(s_beta + cc * 32);
    float y[32];
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      const float4 b4 = (KO & 1) ? make_float4(1.f, 1.f, 1.f, 1.f) : be[q];
      const float bb[4] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const uint32_t pr = xs[c * 16 + 2 * q + (e >> 1)];
        const float x = __uint_as_float((e & 1) ? (pr & 0xFFFF0000u) : (pr << 16));
        const float nrm = __uint_as_float(g[4 * q + e]) + bb[e];
        y[4 * q + e] = (KO & 4) ? x * nrm : x * (MODE == 2 ? sqrt_approx(nrm) : rsqrt_approx(nrm));
      }
    }
    if (map_out != nullptr) {
      uint8_t* row = a2 + (cc >> 1) * kABytes + r * 128;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int physc = ((cc & 1) * 4 + q) ^ (r & 7);
        if (!(KO & 16) || y[8 * q] == 12345.f)
        *reinterpret_cast<uint4*>(row + physc * 16) =
            make_uint4(pack_bf16(y[8 * q], y[8 * q + 1]), pack_bf16(y[8 * q + 2], y[8 * q + 3]),
                       pack_bf16(y[8 * q + 4], y[8 * q + 5]), pack_bf16(y[8 * q + 6], y[8 * q + 7]));
      }
    } else if (valid) {
      store_chunk(out, out_f32, wide, elem0 + cc * 32, y);
    }
  }
  if (map_out != nullptr) {
    fence_proxy_async_smem();
    named_barrier(1, kEpilogueThreads);
    if (leader && !(KO & 16)) {
      for (int kc = 0; kc < nblocks && org.n >= 0; ++kc)
        tma_store_4d(map_out, a2 + kc * kABytes, kc * 64, org.j, org.i, org.n);
      bulk_commit();
    }
  }
}

// bias (+ReLU) on the 32-channel chunks half, half + 2, half + 4, ... (half = 0 / 1)
__device__ __forceinline__ void bias_tile(uint32_t dacc, int half, int npad, int cout,
                                          const float* s_bias, bool relu, bool valid, void* out,
                                          bool out_f32, bool wide, int64_t elem0) {
  for (int c0 = half * 32; c0 < npad; c0 += 64) {
    uint32_t v[32];
    tmem_ld_32x32(dacc + c0, v);
    tmem_ld_wait();
    const float4* bs = reinterpret_cast<const float4*>(s_bias + c0);
    float y[32];
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      const float4 b4 = bs[q];
      y[4 * q + 0] = __uint_as_float(v[4 * q + 0]) + b4.x;
      y[4 * q + 1] = __uint_as_float(v[4 * q + 1]) + b4.y;
      y[4 * q + 2] = __uint_as_float(v[4 * q + 2]) + b4.z;
      y[4 * q + 3] = __uint_as_float(v[4 * q + 3]) + b4.w;
    }
    if (relu) {
#pragma unroll
      for (int e = 0; e < 32; ++e) y[e] = fmaxf(y[e], 0.f);
    }
    if (valid) {
      const int nvalid = min(32, cout - c0);
      if (nvalid == 32 && (cout & 7) == 0) {
        store_chunk(out, out_f32, wide, elem0 + c0, y);
      } else {
#pragma unroll
        for (int e = 0; e < 32; ++e) {
          if (e < nvalid) {
            if (out_f32) static_cast<float*>(out)[elem0 + c0 + e] = y[e];
            else static_cast<__nv_bfloat16*>(out)[elem0 + c0 + e] = __float2bfloat16_rn(y[e]);
          }
        }
      }
    }
  }
}

// the destination admits 256-bit stores: 32-byte aligned base and pixel pitch
inline bool wide_stores_ok(const void* out, int cout, bool out_f32) {
  return (reinterpret_cast<uintptr_t>(out) & 31) == 0 && (cout * (out_f32 ? 4 : 2)) % 32 == 0;
}

}  // namespace epi
}  // namespace tfc
