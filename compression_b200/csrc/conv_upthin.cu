// Thin-output transposed SignalConv2D (the last synthesis layer: 192 -> 3 channels, 5x5 up 2)
// as a "scatter GEMM" on the tensor cores.
//
// Reference semantics: signal_conv.py:629-707 (tf.nn.conv2d_backprop_input with explicit
// padding) + bias_add (:785-804), in the normalised form of tfc_b200.h:
//     out[n,i,j,o] = bias[o] + sum_{a,b,c} U[n, i+a-pt, j+b-pl, c] * w[a,b,c,o],
//     U[su*i', su*j'] = in[i',j'], zero elsewhere.
//
// With cout tiny, the implicit GEMM of conv_umma.cu re-reads the (large) input once per tap and
// issues N=16 MMAs — L2-bound at 3 % of the tensor peak.  Here the roles are swapped:
//     D[pixel, (a,b,o)] = sum_c in[pixel, c] * w[a,b,c,o]          (N = kh*kw*cout <= 128)
// is ONE dense GEMM that reads every input pixel once (TMA -> smem -> tcgen05.mma, fp32
// accumulators in TMEM), and the epilogue does the col2im: the accumulator tile goes to shared
// memory and every output pixel of the tile's interior gathers its (kh/su)*(kw/su) taps from
// there — no atomics, fixed summation order.  Input tiles overlap by the tap halo (1 pixel
// each side for 5x5 up 2); the halo is recomputed, not exchanged.
//
// Algorithmic traffic: cin*2 bytes read per input pixel, su^2*cout*4 bytes written -> HBM bound.
#include <cuda.h>
#include <cuda_bf16.h>

#include <algorithm>

#include "common.cuh"
#include "conv_common.cuh"
#include "umma_ptx.cuh"

namespace {

using namespace tfc::ptx;

constexpr int kTileM = 128;
constexpr int kTH = 8, kTW = 16;       // input tile (with halo): kTH * kTW == kTileM
constexpr int kBlockK = 64;
constexpr int kABytes = kTileM * kBlockK * 2;
constexpr int kMaxNpad = 128;
constexpr int kThreads = 192;          // producer warp, MMA warp, 4 epilogue warps
constexpr int kMaxKChunks = 4;         // cin <= 256

struct UpParams {
  int32_t n, h, w, cin;             // input
  int32_t ho, wo, cout;             // output
  int32_t kh, kw, su, pt, pl;
  int32_t lo_h, lo_w;               // halo before the interior
  int32_t ih, iw;                   // interior (input pixels) per tile
  int32_t tiles_h, tiles_w, num_tiles;
  int32_t chunk;                    // tiles per CTA (0: persistent grid)
  int32_t npad, ld;                 // GEMM N, row stride (floats) of the staged accumulator tile
  int32_t kchunks, stages;
  int32_t relu, out_dtype;          // out_dtype: tfc_dtype (uint8: quantize_image)
  const float* bias;
  void* out;
};

struct __align__(8) UpBarriers {
  uint64_t full[8], empty[8];
  uint64_t d_full[2], d_empty[2];
  uint64_t w_full;
  uint32_t tmem_base;
  uint32_t pad;
};

__device__ __forceinline__ void named_bar_sync(int id, int count) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(count) : "memory");
}

// quantize_image (examples/bmshj2018.py:60-63): saturate_cast<uint8>(round(x * 255)), half to even
__device__ __forceinline__ uint32_t quantize_u8(float x) {
  return (uint32_t)fminf(fmaxf(rintf(x * 255.0f), 0.f), 255.f);
}

__device__ __forceinline__ void tmem_ld_32x16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]),
        "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]),
        "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}

// FAST: the bmshj2018 shape (5x5, up 2, 3 output channels, pad 2) with every gather offset a
// compile-time constant; otherwise the generic gather with run-time loops.
constexpr int kFastLd = 84;   // 25 taps * 3 channels = 75 -> 80 accumulator columns + 4 padding

template <bool FAST>
__global__ void __launch_bounds__(kThreads, 1)
conv_upthin_kernel(const __grid_constant__ CUtensorMap map_a,
                   const __grid_constant__ CUtensorMap map_w, const UpParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // aligned with pointer arithmetic on the __shared__ array (not through an integer): the
  // compiler then knows every pointer derived from `smem` is a shared-memory address and emits
  // LDS / STS instead of generic LD / ST (twice the wavefronts for a broadcast 128-bit load)
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  // carve: [W: kchunks x npad*128 B (1 KB aligned)] [A ring: stages x 16 KB] [D: 2 x 128 x ld
  // floats] [bias] [barriers]
  const int w_chunk = (p.npad * kBlockK * 2 + 1023) & ~1023;
  uint8_t* sw = smem;
  uint8_t* sa = sw + p.kchunks * w_chunk;
  float* sd = reinterpret_cast<float*>(sa + p.stages * kABytes);
  float* s_bias = sd + 2 * kTileM * p.ld;
  UpBarriers* bars = reinterpret_cast<UpBarriers*>(s_bias + 16);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x < 16) s_bias[threadIdx.x] = (p.bias && threadIdx.x < p.cout) ? p.bias[threadIdx.x] : 0.f;
  if (warp == 0 && lane == 0) {
    prefetch_tmap(&map_a);
    prefetch_tmap(&map_w);
    for (int s = 0; s < p.stages; ++s) { mbar_init(&bars->full[s], 1); mbar_init(&bars->empty[s], 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(&bars->d_full[s], 1); mbar_init(&bars->d_empty[s], 128); }
    mbar_init(&bars->w_full, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(&bars->tmem_base, 256);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = bars->tmem_base;
  // persistent (chunk == 0): tiles b, b + grid, ...; chunked: `chunk` consecutive tiles per CTA,
  // many more CTAs than SMs, so the hardware places them on whatever SMs are free (used when
  // the range coder of another batch occupies part of the GPU)
  const int tile0 = p.chunk ? (int)blockIdx.x * p.chunk : (int)blockIdx.x;
  const int tile1 = p.chunk ? min(tile0 + p.chunk, p.num_tiles) : p.num_tiles;
  const int tile_step = p.chunk ? 1 : (int)gridDim.x;

  auto tile_coords = [&](int tile, int& n, int& i0, int& j0) {
    const int per_img = p.tiles_h * p.tiles_w;
    n = tile / per_img;
    const int r = tile - n * per_img;
    const int ti = r / p.tiles_w;
    i0 = ti * p.ih;                       // first interior input row / col of the tile
    j0 = (r - ti * p.tiles_w) * p.iw;
  };

  if (warp == 0) {
    if (lane == 0) {
      // weights: resident for the whole kernel
      mbar_expect_tx(&bars->w_full, p.kchunks * p.npad * kBlockK * 2);
      for (int kc = 0; kc < p.kchunks; ++kc)
        tma_load_2d(sw + kc * w_chunk, &map_w, &bars->w_full, kc * kBlockK, 0);
      int stage = 0;
      uint32_t ph = 0;
      for (int tile = tile0; tile < tile1; tile += tile_step) {
        int n, i0, j0;
        tile_coords(tile, n, i0, j0);
        for (int kc = 0; kc < p.kchunks; ++kc) {
          mbar_wait(&bars->empty[stage], ph ^ 1);
          mbar_expect_tx(&bars->full[stage], kABytes);
          tma_load_5d(sa + stage * kABytes, &map_a, &bars->full[stage], kc * kBlockK,
                      j0 - p.lo_w, 0, i0 - p.lo_h, n);
          if (++stage == p.stages) { stage = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      const uint32_t idesc = make_idesc_bf16(kTileM, p.npad);
      mbar_wait(&bars->w_full, 0);
      int stage = 0;
      uint32_t ph = 0, it = 0;
      for (int tile = tile0; tile < tile1; tile += tile_step, ++it) {
        const int acc = it & 1;
        mbar_wait(&bars->d_empty[acc], ((it >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t d = tmem + acc * 128;
        for (int kc = 0; kc < p.kchunks; ++kc) {
          mbar_wait(&bars->full[stage], ph);
          tc_fence_after();
          const uint64_t da = make_desc_sw128(smem_u32(sa + stage * kABytes));
          const uint64_t db = make_desc_sw128(smem_u32(sw + kc * w_chunk));
#pragma unroll
          for (int k = 0; k < kBlockK / 16; ++k)
            umma_bf16(d, da + (k * 2), db + (k * 2), idesc, (kc | k) != 0);
          umma_commit(&bars->empty[stage]);
          if (++stage == p.stages) { stage = 0; ph ^= 1; }
        }
        umma_commit(&bars->d_full[acc]);
      }
    }
  } else {
    const int quad = warp & 3;
    const int r = quad * 32 + lane;                  // TMEM lane == input pixel of the tile
    const int t = threadIdx.x - 64;                  // 0..127 within the epilogue group
    const uint32_t lane_addr = (uint32_t)(quad * 32) << 16;
    const int OH = p.ih * p.su, OW = p.iw * p.su;    // output pixels of the tile's interior
    const int nout = OH * OW;
    const int taps_n = p.kw * p.cout;                // accumulator columns per kernel row
    uint32_t it = 0;
    for (int tile = tile0; tile < tile1; tile += tile_step, ++it) {
      int n, i0, j0;
      tile_coords(tile, n, i0, j0);
      const int acc = it & 1;
      float* dst = sd + acc * kTileM * p.ld + r * p.ld;
      mbar_wait(&bars->d_full[acc], (it >> 1) & 1);
      tc_fence_after();
      const uint32_t d = tmem + acc * 128 + lane_addr;
      if constexpr (FAST) {
        uint32_t v[80];
#pragma unroll
        for (int c = 0; c < 5; ++c)
          tmem_ld_32x16(d + c * 16, reinterpret_cast<uint32_t(&)[16]>(v[c * 16]));
        tmem_ld_wait();
#pragma unroll
        for (int q = 0; q < 20; ++q)
          *reinterpret_cast<uint4*>(dst + 4 * q) = make_uint4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
      } else {
        for (int c0 = 0; c0 < p.npad; c0 += 16) {
          uint32_t v[16];
          tmem_ld_32x16(d + c0, v);
          tmem_ld_wait();
#pragma unroll
          for (int q = 0; q < 4; ++q)
            *reinterpret_cast<uint4*>(dst + c0 + 4 * q) = make_uint4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
        }
      }
      tc_fence_before();
      mbar_arrive(&bars->d_empty[acc]);              // TMEM is free again; the tile lives in smem
      named_bar_sync(1, 128);
      const float* tilebuf = sd + acc * kTileM * p.ld;
      if constexpr (FAST) {
        // thread <-> interior input pixel (li, lj); it owns the 2x2 output pixels
        // (2i'+pi, 2j'+pj) and sums, for each, the taps a = pi (mod 2), b = pj (mod 2):
        // source pixel offset (pi + a - 2) / 2, accumulator column (a*5 + b)*3 + o
        const int li = r >> 4, lj = r & 15;
        const int ii = i0 + li - 1, jj = j0 + lj - 1;         // input pixel
        if (li >= 1 && li <= 6 && lj >= 1 && lj <= 14 && ii < p.h && jj < p.w) {
          const float* base = tilebuf + r * kFastLd;
          const float b0 = s_bias[0], b1 = s_bias[1], b2 = s_bias[2];
#pragma unroll
          for (int pi = 0; pi < 2; ++pi) {
            float y[6];
#pragma unroll
            for (int pj = 0; pj < 2; ++pj) {
              float s0 = b0, s1 = b1, s2 = b2;
#pragma unroll
              for (int a = pi; a < 5; a += 2) {
#pragma unroll
                for (int b = pj; b < 5; b += 2) {
                  const int dl = (pi + a - 2) / 2, dj = (pj + b - 2) / 2;
                  const float* src = base + (dl * kTW + dj) * kFastLd + (a * 5 + b) * 3;
                  s0 += src[0]; s1 += src[1]; s2 += src[2];
                }
              }
              if (p.relu) { s0 = fmaxf(s0, 0.f); s1 = fmaxf(s1, 0.f); s2 = fmaxf(s2, 0.f); }
              y[pj * 3 + 0] = s0; y[pj * 3 + 1] = s1; y[pj * 3 + 2] = s2;
            }
            const int64_t pix = ((int64_t)n * p.ho + (2 * ii + pi)) * p.wo + 2 * jj;
            if (p.out_dtype == TFC_F32) {
              float2* o2 = reinterpret_cast<float2*>(static_cast<float*>(p.out) + pix * 3);
              o2[0] = make_float2(y[0], y[1]); o2[1] = make_float2(y[2], y[3]); o2[2] = make_float2(y[4], y[5]);
            } else if (p.out_dtype == TFC_U8) {
              // two pixels = 6 bytes at an even offset (wo is even): three 16-bit stores
              uint16_t* o3 = reinterpret_cast<uint16_t*>(static_cast<uint8_t*>(p.out) + pix * 3);
              o3[0] = (uint16_t)(quantize_u8(y[0]) | (quantize_u8(y[1]) << 8));
              o3[1] = (uint16_t)(quantize_u8(y[2]) | (quantize_u8(y[3]) << 8));
              o3[2] = (uint16_t)(quantize_u8(y[4]) | (quantize_u8(y[5]) << 8));
            } else {
              uint32_t* o1 = reinterpret_cast<uint32_t*>(static_cast<__nv_bfloat16*>(p.out) + pix * 3);
              __nv_bfloat162 h0 = __floats2bfloat162_rn(y[0], y[1]), h1 = __floats2bfloat162_rn(y[2], y[3]),
                             h2 = __floats2bfloat162_rn(y[4], y[5]);
              o1[0] = *reinterpret_cast<uint32_t*>(&h0);
              o1[1] = *reinterpret_cast<uint32_t*>(&h1);
              o1[2] = *reinterpret_cast<uint32_t*>(&h2);
            }
          }
        }
        continue;
      }
      // generic col2im gather: output pixel (oi, oj) of the interior sums its taps
      for (int q = t; q < nout; q += 128) {
        const int oi = q / OW, oj = q - oi * OW;
        const int i = i0 * p.su + oi, j = j0 * p.su + oj;
        if (i >= p.ho || j >= p.wo) continue;
        // taps a with (oi + a - pt) % su == 0; input row li = lo + (oi + a - pt) / su
        int a0 = (p.pt - oi) % p.su; if (a0 < 0) a0 += p.su;
        int b0 = (p.pl - oj) % p.su; if (b0 < 0) b0 += p.su;
        float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;   // up to 4 output channels in registers
        for (int a = a0; a < p.kh; a += p.su) {
          const int li = p.lo_h + (oi + a - p.pt) / p.su;   // exact division (may be negative numerator)
          for (int b = b0; b < p.kw; b += p.su) {
            const int lj = p.lo_w + (oj + b - p.pl) / p.su;
            const float* src = tilebuf + (li * kTW + lj) * p.ld + a * taps_n + b * p.cout;
            s0 += src[0];
            if (p.cout > 1) s1 += src[1];
            if (p.cout > 2) s2 += src[2];
            if (p.cout > 3) s3 += src[3];
          }
        }
        const float y[4] = {s0 + s_bias[0], s1 + s_bias[1], s2 + s_bias[2], s3 + s_bias[3]};
        const int64_t pix = ((int64_t)n * p.ho + i) * p.wo + j;
#pragma unroll
        for (int o = 0; o < 4; ++o) {
          if (o < p.cout) {
            const float v = p.relu ? fmaxf(y[o], 0.f) : y[o];
            if (p.out_dtype == TFC_F32) static_cast<float*>(p.out)[pix * p.cout + o] = v;
            else if (p.out_dtype == TFC_U8) static_cast<uint8_t*>(p.out)[pix * p.cout + o] = (uint8_t)quantize_u8(v);
            else static_cast<__nv_bfloat16*>(p.out)[pix * p.cout + o] = __float2bfloat16_rn(v);
          }
        }
      }
      // the buffer `acc` is rewritten two tiles from now; the barrier at that tile's staging
      // step (after its own d_full wait) orders this gather before it
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 256);
}

// w fp32 [kh*kw][cin][cout] -> bf16 [npad rows = (tap, o)][cin]
__global__ void repack_scatter_weights_kernel(const float* __restrict__ w, int taps, int cin,
                                              int cout, int npad, __nv_bfloat16* __restrict__ out) {
  const int total = npad * cin;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int c = i % cin, row = i / cin;
    float v = 0.f;
    if (row < taps * cout) {
      const int tap = row / cout, o = row - tap * cout;
      v = w[((int64_t)tap * cin + c) * cout + o];
    }
    out[i] = __float2bfloat16_rn(v);
  }
}

inline int floor_div(int a, int b) { return (a >= 0) ? a / b : -((-a + b - 1) / b); }

struct Plan {
  int npad, ld, lo_h, hi_h, lo_w, hi_w, ih, iw, stages;
  size_t smem;
};

bool make_plan(const tfc_conv_desc& d, Plan* pl, const char** why) {
  const int n_real = d.kh * d.kw * d.cout;
  pl->npad = (n_real + 15) / 16 * 16;
  if (pl->npad > kMaxNpad) { *why = "kh*kw*cout must be <= 128"; return false; }
  pl->ld = pl->npad + 4;
  pl->lo_h = floor_div(d.pad_top, d.su);
  pl->lo_w = floor_div(d.pad_left, d.su);
  pl->hi_h = std::max(0, floor_div(d.kh - 2 - d.pad_top, d.su) + 1);
  pl->hi_w = std::max(0, floor_div(d.kw - 2 - d.pad_left, d.su) + 1);
  if (pl->lo_h < 0 || pl->lo_w < 0) { *why = "negative halo"; return false; }
  pl->ih = kTH - pl->lo_h - pl->hi_h;
  pl->iw = kTW - pl->lo_w - pl->hi_w;
  if (pl->ih < 1 || pl->iw < 1) { *why = "kernel too large for the tile"; return false; }
  const int kchunks = d.cin / kBlockK;
  const size_t fixed = 1024 + (size_t)kchunks * ((pl->npad * kBlockK * 2 + 1023) & ~1023) +
                       (size_t)2 * kTileM * pl->ld * 4 + 64 + sizeof(UpBarriers) + 64;
  int stages = (int)((227 * 1024 - fixed) / kABytes);
  stages = std::min(stages, 8);
  // tfc_configure(TFC_CFG_CONV_STAGES): a caller that wants range-coder CTAs of another batch to
  // share the SM limits the operand rings; this kernel is HBM bound and its 16 KB stages only
  // cover latency, so one more than the implicit-GEMM kernel's count is kept
  const int cfg = tfc::g_conv_stages.load(std::memory_order_relaxed);
  if (cfg >= 2) stages = std::min(stages, cfg + 1);
  if (stages < 2) { *why = "shared memory"; return false; }
  pl->stages = stages;
  pl->smem = fixed + (size_t)stages * kABytes;
  return true;
}

}  // namespace

namespace tfc {

bool conv_upthin_supported(const tfc_conv_desc& d, const char** why) {
  const char* dummy;
  if (!why) why = &dummy;
  if (d.in_dtype != TFC_BF16) { *why = "input must be bfloat16"; return false; }
  if (d.su < 2 || d.sd != 1) { *why = "scatter path needs strides_up >= 2, strides_down == 1"; return false; }
  if (d.cin % kBlockK || d.cin / kBlockK > kMaxKChunks) { *why = "cin must be 64, 128, 192 or 256"; return false; }
  if (d.cout > 4) { *why = "cout must be <= 4"; return false; }
  if (d.act != TFC_ACT_NONE && d.act != TFC_ACT_RELU) { *why = "GDN is not fused on the scatter path"; return false; }
  if (d.ho != d.h * d.su || d.wo != d.w * d.su) { *why = "output must be strides_up * input"; return false; }
  if (d.pad_top < 0 || d.pad_left < 0 || d.pad_top >= d.kh || d.pad_left >= d.kw) { *why = "padding"; return false; }
  Plan pl;
  return make_plan(d, &pl, why);
}

size_t conv_upthin_workspace(const tfc_conv_desc& d) {
  return (size_t)kMaxNpad * d.cin * 2 + 256;
}

int conv_upthin(const tfc_conv_desc& d, const void* in, const float* w, const float* bias,
                void* out, void* ws, size_t ws_bytes, cudaStream_t st) {
  const char* why = "";
  Plan pl;
  if (!conv_upthin_supported(d, &why) || !make_plan(d, &pl, &why))
    return fail(TFC_ERR_UNIMPLEMENTED, "scatter conv path does not support this shape: %s", why);
  if (ws == nullptr || ws_bytes < conv_upthin_workspace(d))
    return fail(TFC_ERR_INVALID_ARGUMENT, "conv workspace too small: need %zu bytes",
                conv_upthin_workspace(d));
  if ((reinterpret_cast<uintptr_t>(in) & 15) || (reinterpret_cast<uintptr_t>(ws) & 255))
    return fail(TFC_ERR_INVALID_ARGUMENT, "conv buffers must be 16-byte (workspace 256-byte) aligned");
  UpParams p{};
  p.n = d.n; p.h = d.h; p.w = d.w; p.cin = d.cin; p.ho = d.ho; p.wo = d.wo; p.cout = d.cout;
  p.kh = d.kh; p.kw = d.kw; p.su = d.su; p.pt = d.pad_top; p.pl = d.pad_left;
  p.lo_h = pl.lo_h; p.lo_w = pl.lo_w; p.ih = pl.ih; p.iw = pl.iw;
  p.tiles_h = (d.h + pl.ih - 1) / pl.ih;
  p.tiles_w = (d.w + pl.iw - 1) / pl.iw;
  p.num_tiles = d.n * p.tiles_h * p.tiles_w;
  p.npad = pl.npad; p.ld = pl.ld; p.kchunks = d.cin / kBlockK; p.stages = pl.stages;
  p.relu = d.act == TFC_ACT_RELU; p.out_dtype = d.out_dtype;
  p.bias = bias; p.out = out;

  __nv_bfloat16* wb = static_cast<__nv_bfloat16*>(ws);
  if (!(d.flags & TFC_CONV_WEIGHTS_STAGED)) {
    repack_scatter_weights_kernel<<<(pl.npad * d.cin + 255) / 256, 256, 0, st>>>(
        w, d.kh * d.kw, d.cin, d.cout, pl.npad, wb);
    TFC_LAUNCHED("repack_scatter_weights_kernel");
  }

  CUtensorMap map_a, map_w;
  {
    const cuuint64_t C = d.cin;
    cuuint64_t dims[5] = {C, (cuuint64_t)d.w, 1, (cuuint64_t)d.h, (cuuint64_t)d.n};
    cuuint64_t str[4] = {C * 2, (cuuint64_t)d.w * C * 2, (cuuint64_t)d.w * C * 2,
                         (cuuint64_t)d.h * d.w * C * 2};
    cuuint32_t box[5] = {kBlockK, kTW, 1, kTH, 1};
    int rc = make_tensor_map_bf16(&map_a, const_cast<void*>(in), 5, dims, str, box);
    if (rc) return rc;
  }
  {
    cuuint64_t dims[2] = {(cuuint64_t)d.cin, (cuuint64_t)pl.npad};
    cuuint64_t str[1] = {(cuuint64_t)d.cin * 2};
    cuuint32_t box[2] = {kBlockK, (cuuint32_t)pl.npad};
    int rc = make_tensor_map_bf16(&map_w, wb, 2, dims, str, box);
    if (rc) return rc;
  }
  const bool fast = d.su == 2 && d.kh == 5 && d.kw == 5 && d.cout == 3 && d.pad_top == 2 &&
                    d.pad_left == 2 && pl.ld == kFastLd && pl.lo_h == 1 && pl.lo_w == 1 &&
                    pl.ih == 6 && pl.iw == 14;
  auto kern = fast ? conv_upthin_kernel<true> : conv_upthin_kernel<false>;
  TFC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem));
  p.chunk = tiles_per_cta() * 4;    // small tiles: four times the chunk of the big-tile kernel
  const int grid = p.chunk ? (p.num_tiles + p.chunk - 1) / p.chunk
                           : std::min(p.num_tiles, persistent_grid());
  kern<<<grid, kThreads, pl.smem, st>>>(map_a, map_w, p);
  TFC_LAUNCHED("conv_upthin_kernel");
  return TFC_OK;
}

}  // namespace tfc
