// Filter gradient of SignalConv2D on the tensor cores (tcgen05 / TMEM / TMA) for the layers whose
// two channel counts are multiples of 64 (the 128 -> 128 layers of bls2017).
//
// In the normalised form of tfc_b200.h the gradient of one kernel tap is a GEMM over pixels,
//     G[tap][cb][cs] = sum_{n,i,j} big[n, s*i + a - p, s*j + b - q, cb] * small[n, i, j, cs],
// where for a strided correlation (su == 1) big = the layer input, small = the output gradient,
// s = strides_down, and dw[a,b,c,o] = G[a,b][c][o]; for a transposed convolution (su > 1) the roles
// swap — big = the output gradient, small = the layer input, s = strides_up, pad k-1-pad — and
// dw[a,b,c,o] = G[kh-1-a, kw-1-b][o][c].  The reduction dimension is the PIXEL index, and NHWC
// keeps the channels contiguous: both operands are "MN-major" for the MMA, which tcgen05 takes
// directly through the major bits of the instruction descriptor — no transpose anywhere.
//
// Per CTA: one tap and a slice of the 128-pixel tiles.  A TMA producer thread loads, per tile,
// the tap's box of `big` (the stride-s view of conv_umma.cu: one dense box per tap, borders
// zero-filled by the TMA unit) and the tile of `small`, one 16 KB 128-byte-swizzled slab per 64
// channels; the MMA thread issues 8 MMAs (K = 16 pixels each) per tile into ONE TMEM accumulator
// D[cb x cs] that lives for the whole CTA; four epilogue warps read it once at the end and add
// it to dw with fp32 atomics (the slices of a tap meet there).
#include <cuda_bf16.h>

#include <algorithm>

#include "common.cuh"
#include "conv_common.cuh"
#include "umma_ptx.cuh"

namespace {

using namespace tfc::ptx;

constexpr int kSlab = 16 * 1024;      // 128 pixels x 64 channels of bf16
constexpr int kThreads = 64 + 128;    // producer warp, MMA warp, four epilogue warps

struct WTap {
  int16_t dh, dw;       // box shift in units of the stride (rows / columns of the small grid)
  int16_t ph, pw;       // parity inside the stride
  int32_t out_tap;      // a * kw + b of the filter tap this lands in
};

struct WParams {
  int32_t n, cb, cs;              // channels of big / small
  int32_t th, tw, tiles_h, tiles_w, num_tiles;
  int32_t stages, kslices;
  int32_t transposed;             // 0: dw[tap][cb][cs]   1: dw[tap][cs][cb]
  int32_t cin, cout;
  float* dw;
  WTap taps[81];
};

struct __align__(8) Barriers {
  uint64_t full[4], empty[4];
  uint64_t d_full;
  uint32_t tmem_base;
  uint32_t pad;
};

// MN-major operand: rows (= K = pixels) of 128 bytes (64 channels), 128-byte swizzle; groups of
// 8 rows are 1024 B apart (stride byte offset), 64-channel slabs `lbo` bytes apart (leading byte
// offset) — the canonical ((8,n),(8,k)):((1,LBO),(8,SBO)) layout in 16-byte units.
__device__ __forceinline__ uint64_t make_desc_mn_sw128(uint32_t smem_addr, uint32_t lbo) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFFu) >> 4);
  d |= (uint64_t)((lbo >> 4) & 0x3FFFu) << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}

__global__ void __launch_bounds__(kThreads, 1)
conv_wgrad_umma_kernel(const __grid_constant__ CUtensorMap map_big,
                       const __grid_constant__ CUtensorMap map_small, const WParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // aligned with pointer arithmetic on the __shared__ array (not through an integer): the
  // compiler then knows every pointer derived from `smem` is a shared-memory address and emits
  // LDS / STS instead of generic LD / ST (twice the wavefronts for a broadcast 128-bit load)
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  const int nb = p.cb / 64, ns = p.cs / 64;
  const int stage_bytes = (nb + ns) * kSlab;
  Barriers* bars = reinterpret_cast<Barriers*>(smem + p.stages * stage_bytes);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t tmem_cols = p.cs <= 32 ? 32 : (p.cs <= 64 ? 64 : (p.cs <= 128 ? 128 : 256));

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&map_big);
    prefetch_tmap(&map_small);
    for (int s = 0; s < 4; ++s) { mbar_init(&bars->full[s], 1); mbar_init(&bars->empty[s], 1); }
    mbar_init(&bars->d_full, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(&bars->tmem_base, tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = bars->tmem_base;

  const WTap tap = p.taps[blockIdx.x];
  const int per = (p.num_tiles + p.kslices - 1) / p.kslices;
  const int tile0 = (int)blockIdx.y * per;
  const int tile1 = min(tile0 + per, p.num_tiles);

  if (warp == 0) {
    if (lane == 0) {
      int stage = 0;
      uint32_t ph = 0;
      const int per_img = p.tiles_h * p.tiles_w;
      for (int tile = tile0; tile < tile1; ++tile) {
        const int n = tile / per_img;
        const int r = tile - n * per_img;
        const int ti = r / p.tiles_w;
        const int i0 = ti * p.th, j0 = (r - ti * p.tiles_w) * p.tw;
        mbar_wait(&bars->empty[stage], ph ^ 1);
        uint8_t* sa = smem + stage * stage_bytes;
        mbar_expect_tx(&bars->full[stage], (uint32_t)stage_bytes);
        for (int c = 0; c < nb; ++c)
          tma_load_5d(sa + c * kSlab, &map_big, &bars->full[stage], tap.pw * p.cb + c * 64,
                      j0 + tap.dw, tap.ph, i0 + tap.dh, n);
        for (int c = 0; c < ns; ++c) {
          asm volatile(
              "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes"
              " [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(smem_u32(sa + (nb + c) * kSlab)),
              "l"(reinterpret_cast<uint64_t>(&map_small)), "r"(smem_u32(&bars->full[stage])),
              "r"(c * 64), "r"(j0), "r"(i0), "r"(n)
              : "memory");
        }
        if (++stage == p.stages) { stage = 0; ph ^= 1; }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      // kind::f16, bf16 x bf16 -> f32, A and B MN-major (bits 15, 16)
      const uint32_t idesc = make_idesc_bf16(p.cb, p.cs) | (1u << 15) | (1u << 16);
      int stage = 0;
      uint32_t ph = 0;
      bool first = true;
      for (int tile = tile0; tile < tile1; ++tile) {
        mbar_wait(&bars->full[stage], ph);
        tc_fence_after();
        const uint32_t sa = smem_u32(smem + stage * stage_bytes);
        const uint64_t da = make_desc_mn_sw128(sa, kSlab);
        const uint64_t db = make_desc_mn_sw128(sa + nb * kSlab, kSlab);
#pragma unroll
        for (int k = 0; k < 8; ++k) {      // 16 pixel rows = 2048 bytes per step
          umma_bf16(tmem, da + (uint64_t)(k * 128), db + (uint64_t)(k * 128), idesc, !first);
          first = false;
        }
        umma_commit(&bars->empty[stage]);
        if (++stage == p.stages) { stage = 0; ph ^= 1; }
      }
      umma_commit(&bars->d_full);
    }
  } else {
    // epilogue: warp w reads TMEM lanes 32*(w%4) .. +31 (= rows cb), all cs columns
    if (tile1 > tile0) {
      const int quad = warp & 3;
      const int row = quad * 32 + lane;
      mbar_wait(&bars->d_full, 0);
      tc_fence_after();
      float* dst = p.dw + (int64_t)tap.out_tap * p.cin * p.cout;
      for (int c0 = 0; c0 < p.cs; c0 += 32) {
        uint32_t v[32];
        tmem_ld_32x32(tmem + ((uint32_t)(quad * 32) << 16) + (uint32_t)c0, v);
        tmem_ld_wait();
        if (row < p.cb) {
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const int col = c0 + j;
            if (col < p.cs) {
              const int64_t at = p.transposed ? (int64_t)col * p.cout + row
                                              : (int64_t)row * p.cout + col;
              atomicAdd(dst + at, __uint_as_float(v[j]));
            }
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, tmem_cols);
}

inline int floor_div(int a, int b) { return (a >= 0) ? a / b : -((-a + b - 1) / b); }

}  // namespace

namespace tfc {

bool conv_wgrad_umma_supported(const tfc_conv_desc& d) {
  if (d.sd > 1 && d.su > 1) return false;
  const int cb = d.su > 1 ? d.cout : d.cin, cs = d.su > 1 ? d.cin : d.cout;
  if (cb != 128 && cb != 64) return false;            // UMMA M
  if (cb == 64) return false;                         // M = 64 uses another TMEM layout: not built
  if (cs % 64 || cs > 256 || cs < 64) return false;   // UMMA N, whole 64-channel slabs
  if (d.kh * d.kw > 81) return false;
  const int s = d.su > 1 ? d.su : d.sd;
  const int hb = d.su > 1 ? d.ho : d.h, wb = d.su > 1 ? d.wo : d.w;
  const int hs = d.su > 1 ? d.h : d.ho, ws = d.su > 1 ? d.w : d.wo;
  if (hb % s || wb % s || s > 8) return false;
  if (d.su > 1 && (hb != hs * s || wb != ws * s)) return false;
  if (d.n <= 0 || (int64_t)d.n * hs * ws < 128) return false;
  return true;
}

int conv_wgrad_umma(const tfc_conv_desc& d, const void* x, const void* dy, float* dw,
                    cudaStream_t st) {
  const bool up = d.su > 1;
  WParams p{};
  p.n = d.n; p.cin = d.cin; p.cout = d.cout;
  p.cb = up ? d.cout : d.cin;
  p.cs = up ? d.cin : d.cout;
  p.transposed = up ? 1 : 0;
  p.dw = dw;
  const int s = up ? d.su : d.sd;
  const int hb = up ? d.ho : d.h, wb = up ? d.wo : d.w;
  const int hs = up ? d.h : d.ho, ws = up ? d.w : d.wo;
  const void* big = up ? dy : x;
  const void* small = up ? x : dy;
  const int pt = up ? d.kh - 1 - d.pad_top : d.pad_top;
  const int pl = up ? d.kw - 1 - d.pad_left : d.pad_left;
  int count = 0;
  for (int a = 0; a < d.kh; ++a) {
    for (int b = 0; b < d.kw; ++b) {
      WTap t{};
      const int ya = a - pt, xb = b - pl;
      const int qh = floor_div(ya, s), qw = floor_div(xb, s);
      t.dh = (int16_t)qh; t.dw = (int16_t)qw;
      t.ph = (int16_t)(ya - qh * s); t.pw = (int16_t)(xb - qw * s);
      t.out_tap = up ? (d.kh - 1 - a) * d.kw + (d.kw - 1 - b) : a * d.kw + b;
      p.taps[count++] = t;
    }
  }
  conv_pick_tile(hs, ws, &p.th, &p.tw);
  p.tiles_h = (hs + p.th - 1) / p.th;
  p.tiles_w = (ws + p.tw - 1) / p.tw;
  p.num_tiles = d.n * p.tiles_h * p.tiles_w;
  const int stage_bytes = (p.cb / 64 + p.cs / 64) * kSlab;
  p.stages = std::max(2, std::min(4, (int)((227 * 1024 - 2048) / stage_bytes)));
  if ((size_t)p.stages * stage_bytes + 2048 > 227 * 1024)
    return fail(TFC_ERR_UNIMPLEMENTED, "filter gradient: tile does not fit in shared memory");
  const int sms = num_sms();
  // one CTA per SM (the stages fill shared memory) and uniform work per CTA: a single wave of
  // taps x slices CTAs, two waves when one would leave more than 30 % of the SMs idle
  p.kslices = std::max(1, sms / count);
  if (count * p.kslices * 10 < sms * 7) p.kslices = std::max(1, 2 * sms / count);
  p.kslices = std::min(p.kslices, p.num_tiles);
  const size_t smem = 1024 + (size_t)p.stages * stage_bytes + sizeof(Barriers) + 64;

  CUtensorMap map_big, map_small;
  {
    const cuuint64_t C = p.cb, S = s;
    cuuint64_t dims[5] = {S * C, (cuuint64_t)wb / S, S, (cuuint64_t)hb / S, (cuuint64_t)d.n};
    cuuint64_t str[4] = {S * C * 2, (cuuint64_t)wb * C * 2, S * wb * C * 2,
                         (cuuint64_t)hb * wb * C * 2};
    cuuint32_t box[5] = {64, (cuuint32_t)p.tw, 1, (cuuint32_t)p.th, 1};
    int rc = make_tensor_map_bf16(&map_big, const_cast<void*>(big), 5, dims, str, box);
    if (rc) return rc;
  }
  {
    const cuuint64_t C = p.cs;
    cuuint64_t dims[4] = {C, (cuuint64_t)ws, (cuuint64_t)hs, (cuuint64_t)d.n};
    cuuint64_t str[3] = {C * 2, (cuuint64_t)ws * C * 2, (cuuint64_t)hs * ws * C * 2};
    cuuint32_t box[4] = {64, (cuuint32_t)p.tw, (cuuint32_t)p.th, 1};
    int rc = make_tensor_map_bf16(&map_small, const_cast<void*>(small), 4, dims, str, box);
    if (rc) return rc;
  }
  static tfc::PerDeviceOnce attr;
  if (attr.first()) {
    TFC_CUDA(cudaFuncSetAttribute(conv_wgrad_umma_kernel,
                                  cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
  }
  dim3 grid(count, p.kslices, 1);
  conv_wgrad_umma_kernel<<<grid, kThreads, smem, st>>>(map_big, map_small, p);
  TFC_LAUNCHED("conv_wgrad_umma_kernel");
  return TFC_OK;
}

}  // namespace tfc
