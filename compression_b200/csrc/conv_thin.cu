// Thin-input SignalConv2D (the first analysis layer: 3 image channels -> 128/192 filters, 5x5
// stride 2 in bmshj2018, 9x9 stride 4 in bls2017) as ONE tcgen05 kernel fed by a software
// im2col, with the bias / GDN epilogue of conv_epilogue.cuh.
//
// Reference semantics: signal_conv.py:487-524 (tf.nn.conv2d with explicit same_zeros padding),
// :804 bias_add, gdn.py:183-202.
//
// The whole receptive field of an output pixel is only K = kh*kw*cin values (75 / 243), so the
// GEMM per 128-pixel tile is D[128, cout] = A[128, K] * W[K, cout] with K padded to 16:
//   * W (K-major bf16, ceil(K/64) blocks) and gamma stay RESIDENT in shared memory for the life
//     of the CTA — one TMA load each; the tile loop itself reads only the image.
//   * A is built by two "builder" warps straight from the fp32 / bf16 NHWC image into the
//     128-byte-swizzled K-major layout tcgen05.mma expects (row = pixel, 16-byte chunk c of
//     row r stored at chunk c ^ (r & 7)), through a ring of 16 KB stages.  The k -> (a, b, c)
//     decomposition is resolved at compile time (kernel templated on kh, kw, cin).
//   * warp 0: TMA (weights once, an image patch per tile, and — TS form — the stores of the
//     staged output tiles); warp 1: MMA issuer; warps 2-3: builders; warps 4-11 (4-19 in the TS
//     form: GDN with a bfloat16 output, see the kernel comment): epilogue.
// Compared with the space-to-depth formulation (conv_s2d.cu, 9 k-blocks of 64 mostly-zero
// channels per tile, weights streamed per tile) this issues 5 instead of 36 MMAs per tile and
// takes the shared-memory fill traffic per tile from 360 KB to the 8 KB image patch.
#include <cuda.h>
#include <cuda_bf16.h>

#include <algorithm>
#include <type_traits>

#include "common.cuh"
#include "conv_common.cuh"
#include "conv_epilogue.cuh"
#include "umma_ptx.cuh"

namespace {

using namespace tfc::ptx;
namespace epi = tfc::epi;

constexpr int kEpiThreads = tfc::epi::kEpilogueThreads;
constexpr int kThreads = 64 + kEpiThreads + 64;
// the tensor-memory (TS) GDN epilogue: sixteen warps, four per lane quadrant
constexpr int kEpiThreadsTs = 512;
constexpr int kThreadsTs = 64 + 64 + kEpiThreadsTs;
constexpr int kBuilders = 64;         // two warps, two tile rows per thread
constexpr int kABytes = epi::kABytes;
constexpr int kMaxStages = 4;

// Bias and beta folded into the GEMMs (TS form).  The last k-block of W uses `tail` of its 64
// K columns; the three bfloat16 pieces of the bias sit in the next three (inside the padding of
// the last 16-wide K step, where the im2col matrix carries 1.0), and the pieces of beta at
// positions 13..15 of the 16-wide slice starting at `beta_slice` — columns no main-GEMM step
// multiplies by anything but zero.
// Phase stamps (tools/a0_trace.py): clock64() of CTA 0's phase boundaries over tiles 8..15 of the
// tensor-memory GDN epilogue, 64 predicated 8-byte stores per launch into a 4 KB device buffer
// (concurrent launches overwrite each other's stamps: it is a debugging aid).  Compiled in by
// default (-DTFC_A0_STAMPS=0 removes them) because the kernel is FASTER with them: A/B on one box,
// two builds alternated twice, 0.760 against 0.876 ms at the bmshj2018 bench shape with identical
// results (profiles/r4h_trace_ab.txt) — ptxas schedules the epilogue's two phases differently around
// the clock reads; what exactly it does better has not been tracked down.
#ifndef TFC_A0_STAMPS
#define TFC_A0_STAMPS 1
#endif
#if TFC_A0_STAMPS
__device__ unsigned long long g_a0_trace[8 * 64];
#define TR(it, slot) do { if (blockIdx.x == 0 && (it) >= 8 && (it) < 16) g_a0_trace[((it) - 8) * 64 + (slot)] = clock64(); } while (0)
#else
#define TR(it, slot) do { } while (0)
#endif

namespace fold {
__host__ __device__ constexpr int tail(int K) { return K - (K - 1) / 64 * 64; }
__host__ __device__ constexpr int beta_slice(int K) { return (tail(K) + 5) / 16 * 16; }
__host__ __device__ constexpr bool fits(int K) {
  return tail(K) + 3 <= (tail(K) + 15) / 16 * 16 && beta_slice(K) + 16 <= 64;
}
}  // namespace fold

struct Params {
  int32_t n, h, w;                  // input [n, h, w, CIN]
  int32_t ho, wo, cout, npad;
  int32_t sd, pt, pl;
  int32_t th, tw, tiles_h, tiles_w, num_tiles, chunk;
  int32_t stages;
  int32_t patch;                    // image patches staged by TMA (else the builders use LDG)
  int32_t patch_pitch;              // elements per patch row
  int32_t patch_bytes;              // TMA box bytes
  int32_t patch_slot;               // bytes reserved per patch buffer
  int32_t vec;                      // patch reads: elements per shared-memory load (1 or 2)
  int32_t per16;                    // image elements per 16 bytes
  int32_t act, out_f32, out_wide, in_dtype;   // in_dtype: tfc_dtype (uint8: byte / 255)
  int32_t esize;                    // bytes per image element
  int32_t out_tma;                  // GDN output leaves by TMA tile stores
  int32_t debug;
  uint32_t rcp_per_img, rcp_tiles_w;  // ceil(2^32 / divisor) when exact for every tile, else 0
  const void* in;
  const float* bias;
  const float* beta;
  void* out;
};

struct __align__(8) Barriers {
  uint64_t a_full[kMaxStages], a_empty[kMaxStages];
  uint64_t w_full;
  uint64_t p_full[2], p_empty[2];
  uint64_t d1_full[2], d1_empty[2];
  uint64_t a2_full, d2_full[1];
  uint64_t y_full, y_free;          // TS: the staged y tile is complete / has been read by its store
  uint32_t tmem_base;
  uint32_t pad;
};

// image element -> float.  A uint8 image stands for byte / 255 (read_png, examples/
// bmshj2018.py:53-57); float(b) * (1/255.f) rounds to the same bfloat16 operand as float(b) / 255
// for all 256 bytes (checked exhaustively), so the layer's result does not depend on which of the
// two forms the caller feeds.
constexpr float kInv255 = 1.0f / 255.0f;
__device__ __forceinline__ float ldf(const float* p) { return __ldg(p); }
__device__ __forceinline__ float ldf(const __nv_bfloat16* p) {
  return __bfloat162float(*p);
}
__device__ __forceinline__ float ldf(const uint8_t* p) { return (float)__ldg(p) * kInv255; }
// the same from shared memory (TMA-staged patch)
__device__ __forceinline__ float lds1(const float* p) { return *p; }
__device__ __forceinline__ float lds1(const __nv_bfloat16* p) { return __bfloat162float(*p); }
__device__ __forceinline__ float lds1(const uint8_t* p) { return (float)(*p) * kInv255; }

// One tile row (= output pixel) of k-block KBI: K-columns [64*KBI, 64*KBI + 16*NM) of the im2col
// matrix, k = (a*KW + b)*CIN + c.  `src` points at the top-left tap of the pixel's window.
template <int KH, int KW, int CIN, int KBI, int NM, bool CHECK, bool FOLD, typename T>
__device__ __forceinline__ void build_row(const T* __restrict__ src, int64_t row_stride, int y0,
                                          int x0, int h, int w, uint8_t* row, int r) {
  constexpr int K = KH * KW * CIN;
  constexpr int L = KW * CIN;
  float v[NM * 16];
#pragma unroll
  for (int e = 0; e < NM * 16; ++e) {
    const int k = KBI * 64 + e;            // compile-time after unrolling
    const int a = k / L, t = k - a * L, b = t / CIN;
    if (k < K) {
      if (CHECK) {
        const bool ok = (unsigned)(y0 + a) < (unsigned)h && (unsigned)(x0 + b) < (unsigned)w;
        v[e] = ok ? ldf(src + a * row_stride + t) : 0.f;
      } else {
        v[e] = ldf(src + a * row_stride + t);
      }
    } else {
      v[e] = (FOLD && k < K + 3) ? 1.f : 0.f;      // the columns that meet the bias rows of W
    }
  }
#pragma unroll
  for (int c = 0; c < NM * 2; ++c) {
    const int physc = c ^ (r & 7);
    *reinterpret_cast<uint4*>(row + physc * 16) =
        make_uint4(epi::pack_bf16(v[8 * c], v[8 * c + 1]), epi::pack_bf16(v[8 * c + 2], v[8 * c + 3]),
                   epi::pack_bf16(v[8 * c + 4], v[8 * c + 5]), epi::pack_bf16(v[8 * c + 6], v[8 * c + 7]));
  }
}

__device__ __forceinline__ void lds_vec(const float* p, float (&o)[2]) {
  const float2 v = *reinterpret_cast<const float2*>(p);
  o[0] = v.x; o[1] = v.y;
}
__device__ __forceinline__ void lds_vec(const __nv_bfloat16* p, float (&o)[2]) {
  const float2 v = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(p));
  o[0] = v.x; o[1] = v.y;
}
__device__ __forceinline__ void lds_vec(const uint8_t* p, float (&o)[2]) {
  const uchar2 v = *reinterpret_cast<const uchar2*>(p);
  o[0] = (float)v.x * kInv255; o[1] = (float)v.y * kInv255;
}

// The same row from the TMA-staged patch in shared memory (zero padding already applied by the
// TMA unit).  `src` points at the top-left tap inside the patch; V = 2 reads element pairs
// (LDS.64 / LDS.32), which needs an even sd*CIN.
template <int KH, int KW, int CIN, int KBI, int NM, int V, bool FOLD, typename T>
__device__ __forceinline__ void build_row_patch(const T* src, int pitch, uint8_t* row, int r) {
  constexpr int K = KH * KW * CIN;
  constexpr int L = KW * CIN;
  constexpr int lo = KBI * 64, hi = (lo + NM * 16 < K) ? lo + NM * 16 : K;
  float v[NM * 16];
#pragma unroll
  for (int e = 0; e < NM * 16; ++e)
    v[e] = (FOLD && lo + e >= K && lo + e < K + 3) ? 1.f : 0.f;   // 1.0 against the bias rows of W
#pragma unroll
  for (int a = 0; a < KH; ++a) {
#pragma unroll
    for (int q = 0; q < (L + V - 1) / V; ++q) {
      const int k0 = a * L + q * V;            // compile-time after unrolling
      if (k0 + V > lo && k0 < hi) {
        if (V == 2) {
          float pr[2];
          lds_vec(src + a * pitch + q * 2, pr);
          if (k0 >= lo && k0 < hi) v[k0 - lo] = pr[0];
          if (q * 2 + 1 < L && k0 + 1 >= lo && k0 + 1 < hi) v[k0 + 1 - lo] = pr[1];
        } else {
          v[k0 - lo] = lds1(src + a * pitch + q);
        }
      }
    }
  }
#pragma unroll
  for (int c = 0; c < NM * 2; ++c) {
    const int physc = c ^ (r & 7);
    *reinterpret_cast<uint4*>(row + physc * 16) =
        make_uint4(epi::pack_bf16(v[8 * c], v[8 * c + 1]), epi::pack_bf16(v[8 * c + 2], v[8 * c + 3]),
                   epi::pack_bf16(v[8 * c + 4], v[8 * c + 5]), epi::pack_bf16(v[8 * c + 6], v[8 * c + 7]));
  }
}

template <int KH, int KW, int CIN, int KBI, bool FOLD = false, typename T>
__device__ __forceinline__ void build_block_patch(const Params& p, const T* patch, int t,
                                                  uint8_t* sa) {
  constexpr int NM = (KH * KW * CIN - KBI * 64 >= 64) ? 4 : (KH * KW * CIN - KBI * 64 + 15) / 16;
#pragma unroll 1
  for (int rr = 0; rr < 2; ++rr) {
    const int r = t + rr * kBuilders;
    const int li = r / p.tw, lj = r - li * p.tw;
    const T* src = patch + li * p.sd * p.patch_pitch + lj * p.sd * CIN;
    uint8_t* row = sa + r * 128;
    if (p.vec == 2)
      build_row_patch<KH, KW, CIN, KBI, NM, 2, FOLD>(src, p.patch_pitch, row, r);
    else
      build_row_patch<KH, KW, CIN, KBI, NM, 1, FOLD>(src, p.patch_pitch, row, r);
  }
}

template <int K, int KBI>
struct BlockMmas {   // 16-wide MMA K-steps that k-block KBI really needs
  static constexpr int rest = K - KBI * 64;
  static constexpr int value = rest >= 64 ? 4 : (rest + 15) / 16;
};

template <int KH, int KW, int CIN, int KBI, bool FOLD = false, typename T>
__device__ __forceinline__ void build_block(const Params& p, const T* img, int n, int i0, int j0,
                                            int t, uint8_t* sa) {
  constexpr int NM = BlockMmas<KH * KW * CIN, KBI>::value;
#pragma unroll 1
  for (int rr = 0; rr < 2; ++rr) {
    const int r = t + rr * kBuilders;
    const int li = r / p.tw, lj = r - li * p.tw;
    const int y0 = (i0 + li) * p.sd - p.pt, x0 = (j0 + lj) * p.sd - p.pl;
    const int64_t row_stride = (int64_t)p.w * CIN;
    const T* src = img + ((int64_t)n * p.h + y0) * row_stride + (int64_t)x0 * CIN;
    uint8_t* row = sa + r * 128;
    const bool inside = y0 >= 0 && x0 >= 0 && y0 + KH <= p.h && x0 + KW <= p.w;
    if (inside)
      build_row<KH, KW, CIN, KBI, NM, false, FOLD>(src, row_stride, y0, x0, p.h, p.w, row, r);
    else
      build_row<KH, KW, CIN, KBI, NM, true, FOLD>(src, row_stride, y0, x0, p.h, p.w, row, r);
  }
}

// MODE 0: bias (+ReLU).  MODE 1: GDN with cout == 32 * NC.
//
// TS (GDN with a bfloat16 output; the default): this layer has K = 75, so there is no long main
// GEMM to hide the epilogue's chain  D1 -> x^2 -> [GEMM2 round trip] -> D2 -> y -> store  behind:
// a tile takes 4 us against 0.4 us of tensor-pipe work, and warp-state sampling of the 8-warp
// epilogue (profiles/r3f_*) puts 23 % of its time on the GEMM2 round trip, 14 % on waiting for
// the previous tile's TMA store to leave the buffer x^2 is about to be written to, 8 % on the
// tile-coordinate divisions, and shows the bias / beta loads' latency exposed in front of every
// FADD.  The TS form removes those:
//   * x^2 goes to TENSOR MEMORY (tcgen05.st, 96 columns of packed bfloat16 next to the two
//     accumulators) and GEMM2 takes its A operand from there (tcgen05.mma with A in TMEM): no x^2
//     tile in shared memory, so the 48 KB buffer only stages y and the wait for the previous
//     store moves behind the GEMM2 round trip, where it is free;
//   * bias and beta are FOLDED INTO THE GEMMS: the im2col matrix carries 1.0 in three of its
//     padding columns (K = 75 is padded to 80 anyway) against three bfloat16 pieces of the bias
//     (hi + mid + lo = the float32 value to 2^-24) in the matching rows of W, and GEMM2 gets one
//     more K = 16 step whose A is a constant block of tensor memory and whose B rows hold the
//     three pieces of beta in an unused corner of W's last k-block — the epilogue has no
//     shared-memory loads and no adds left;
//   * sixteen epilogue warps (four per tensor-memory lane quadrant, a quarter of the channels
//     each) instead of eight, pipelined over two tiles — square(i+1) between the end of GEMM2(i)
//     and normalise(i), so that GEMM2(i+1) runs behind normalise(i); the stores issued by the
//     patch-loading thread; tile coordinates by multiply-high instead of division.
// 1.28 -> 0.76 ms at the bmshj2018 bench shape (0.88 without the phase stamps below:
// profiles/r3g_a0_stall_breakdown.md, r3o_a0_trace.txt, r4h_trace_ab.txt).
// Rounds 1-2 tried a split by phase and a software-pipelined 8-warp epilogue first (slower:
// profiles/r2i_*, r2n_*); the older epilogue stays as the float32-output / debug-bit-32 form.
template <int MODE, int NC, int KH, int KW, int CIN, bool TS, int KO = 0>
__global__ void __launch_bounds__(TS ? kThreadsTs : kThreads, 1)
conv_thin_kernel(const __grid_constant__ CUtensorMap map_w,
                 const __grid_constant__ CUtensorMap map_g,
                 const __grid_constant__ CUtensorMap map_img,
                 const __grid_constant__ CUtensorMap map_out, const Params p) {
  constexpr bool GDN = MODE != 0;
  static_assert(!TS || GDN, "the tensor-memory epilogue is the GDN epilogue");
  constexpr int K = KH * KW * CIN;
  constexpr int KB = (K + 63) / 64;
  constexpr int gdn_kblocks = GDN ? NC / 2 : 0;
  constexpr int kEpi = TS ? kEpiThreadsTs : kEpiThreads;
  constexpr int kAll = TS ? kThreadsTs : kThreads;
  // tensor memory (TS): [accumulator 0][accumulator 1][x^2 as bf16 pairs][constant block]
  constexpr int ACC = TS ? NC * 32 : 256;
  constexpr int X2COL = 2 * ACC, ONESCOL = X2COL + NC * 16;
  static_assert(!TS || ONESCOL + 8 <= 512, "tensor memory");
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // carve: [A ring: stages x 16 KB] [W: KB x npad*128 B] [gamma: cout/64 x npad*128 B]
  //        [A2 / y staging: cout/64 x 16 KB] [2 image patches] [bias|beta] [barriers]
  // aligned with pointer arithmetic on the __shared__ array (not through an integer): the
  // compiler then knows every pointer derived from `smem` is a shared-memory address and emits
  // LDS / STS instead of generic LD / ST (twice the wavefronts for a broadcast 128-bit load)
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  const int b_bytes = p.npad * 128;
  const int b_slot = (b_bytes + 1023) & ~1023;
  uint8_t* w_res = smem + p.stages * kABytes;
  uint8_t* g_res = w_res + KB * b_slot;
  uint8_t* a2 = g_res + gdn_kblocks * b_slot;
  uint8_t* patches = a2 + gdn_kblocks * kABytes;
  float* s_bias = reinterpret_cast<float*>(patches + (p.patch ? 2 * p.patch_slot : 0));
  float* s_beta = s_bias + 256;
  Barriers* bars = reinterpret_cast<Barriers*>(s_beta + 256);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if constexpr (!TS) {
    for (int i = threadIdx.x; i < 256; i += kAll) {
      s_bias[i] = (p.bias && i < p.cout) ? p.bias[i] : 0.f;
      s_beta[i] = (GDN && i < p.cout) ? p.beta[i] : 1.f;
    }
  }
  if (warp == 0 && lane == 0) {
    prefetch_tmap(&map_w);
    if (GDN) prefetch_tmap(&map_g);
    for (int s = 0; s < kMaxStages; ++s) {
      mbar_init(&bars->a_full[s], kBuilders / 32);
      mbar_init(&bars->a_empty[s], 1);
    }
    mbar_init(&bars->w_full, 1);
    for (int s = 0; s < 2; ++s) { mbar_init(&bars->p_full[s], 1); mbar_init(&bars->p_empty[s], kBuilders / 32); }
    if (p.patch) prefetch_tmap(&map_img);
    constexpr int kEpiArrivals = kEpi;
    for (int s = 0; s < 2; ++s) { mbar_init(&bars->d1_full[s], 1); mbar_init(&bars->d1_empty[s], kEpiArrivals); }
    mbar_init(&bars->a2_full, kEpiArrivals);
    mbar_init(&bars->d2_full[0], 1);
    mbar_init(&bars->y_full, kEpi);
    mbar_init(&bars->y_free, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(&bars->tmem_base, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = bars->tmem_base;
  const int tile0 = p.chunk ? (int)blockIdx.x * p.chunk : (int)blockIdx.x;
  const int tile1 = p.chunk ? min(tile0 + p.chunk, p.num_tiles) : p.num_tiles;
  const int tile_step = p.chunk ? 1 : (int)gridDim.x;
  const uint32_t idesc = make_idesc_bf16(128, p.npad);

  // tile -> (image, first output row, first output column); the quotients by multiply-high with
  // host-made reciprocals (exact for every tile index: checked by the host before it sets them)
  auto tile_coords = [&](int tile, int& n, int& i0, int& j0) {
    const int per_img = p.tiles_h * p.tiles_w;
    n = p.rcp_per_img ? (int)__umulhi((uint32_t)tile, p.rcp_per_img) : tile / per_img;
    const int r = tile - n * per_img;
    const int ti = p.rcp_tiles_w ? (int)__umulhi((uint32_t)r, p.rcp_tiles_w) : r / p.tiles_w;
    i0 = ti * p.th;
    j0 = (r - ti * p.tiles_w) * p.tw;
  };

  if (warp >= 4) {
    if constexpr (TS) {
      // ===== epilogue (TS): 16 warps, thread <-> TMEM lane <-> tile pixel, four column parts =====
      constexpr int CP = NC * 8;                      // channels per thread
      constexpr int NPC = CP / 16;                    // 16-column pieces
      const int quad = warp & 3;
      const int part = (warp - 4) >> 2;
      const int r = quad * 32 + lane;
      const uint32_t lane_addr = (uint32_t)(quad * 32) << 16;
      const bool tr = quad == 0 && lane == 0;
      const int ts0 = part * 12;
      if (part == 0) {
        // A of the beta step of GEMM2: 1.0 at K positions 13, 14, 15 of its 16 (written once)
        const uint32_t ones[8] = {0u, 0u, 0u, 0u, 0u, 0u, 0x3F800000u, 0x3F803F80u};
        tmem_st_32x8(tmem + ONESCOL + lane_addr, ones);
        tmem_st_wait();
      }
      const int ntiles = tile1 > tile0 ? (tile1 - tile0 + tile_step - 1) / tile_step : 0;
      // phase 1 of tile number `it`: x = D1 (bias already in it) -> registers (packed bf16),
      // x^2 -> tensor memory; 8 columns at a time so that the x of two tiles fits in registers
      auto square = [&](uint32_t it, uint32_t (&xs)[CP / 2]) {
        const int acc = it & 1;
        const uint32_t dacc = tmem + acc * ACC + lane_addr + part * CP;
        mbar_wait(&bars->d1_full[acc], (it >> 1) & 1);
        tc_fence_after();
        if (tr) TR(it, ts0 + 0);
#pragma unroll
        for (int pc = 0; pc < CP / 8; ++pc) {
          uint32_t v[8], sq[4];
          tmem_ld_32x8(dacc + pc * 8, v);
          tmem_ld_wait();
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const float x0 = __uint_as_float(v[2 * e]), x1 = __uint_as_float(v[2 * e + 1]);
            xs[pc * 4 + e] = epi::pack_bf16(x0, x1);
            sq[e] = epi::pack_bf16(x0 * x0, x1 * x1);
          }
          tmem_st_32x4(tmem + X2COL + lane_addr + part * (CP / 2) + pc * 4, sq);
        }
        tmem_st_wait();
        tc_fence_before();
        mbar_arrive(&bars->a2_full);
        if (tr) TR(it, ts0 + 2);
      };
      // phase 2 of tile number `it` (its d2_full has been waited for): y = x * rsqrt(D2) (beta
      // already in it), computed IN PLACE over x; only then does the thread need the staging
      // tile, so the previous tile's TMA store drains behind square() and this arithmetic
      auto normalise = [&](uint32_t it, uint32_t (&xs)[CP / 2]) {
        const int acc = it & 1;
        const uint32_t dacc = tmem + acc * ACC + lane_addr + part * CP;
        tc_fence_after();
        if (tr) TR(it, ts0 + 3);
#pragma unroll
        for (int pc = 0; pc < CP / 8; ++pc) {
          uint32_t g[8];
          tmem_ld_32x8(dacc + pc * 8, g);
          tmem_ld_wait();
          if (pc == CP / 8 - 1) {
            // D2 is in registers: hand the accumulator back now
            tc_fence_before();
            mbar_arrive(&bars->d1_empty[acc]);
            if (tr) TR(it, ts0 + 4);
          }
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const uint32_t pr = xs[pc * 4 + e];
            const float y0 = __uint_as_float(pr << 16) *
                             epi::rsqrt_approx(__uint_as_float(g[2 * e]));
            const float y1 = __uint_as_float(pr & 0xFFFF0000u) *
                             epi::rsqrt_approx(__uint_as_float(g[2 * e + 1]));
            xs[pc * 4 + e] = epi::pack_bf16(y0, y1);
          }
        }
        if (tr) TR(it, ts0 + 5);
        if (it > 0) mbar_wait(&bars->y_free, (it - 1) & 1);
        if (tr) TR(it, ts0 + 6);
#pragma unroll
        for (int h = 0; h < CP / 8; ++h) {            // 16-byte chunks of 8 channels
          const int gc = part * (CP / 8) + h;         // chunk of the 128-byte-swizzled tile row
          *reinterpret_cast<uint4*>(a2 + (gc >> 3) * kABytes + r * 128 + (((gc & 7) ^ (r & 7)) * 16)) =
              make_uint4(xs[h * 4], xs[h * 4 + 1], xs[h * 4 + 2], xs[h * 4 + 3]);
        }
        fence_proxy_async_smem();
        mbar_arrive(&bars->y_full);
        if (tr) TR(it, ts0 + 7);
      };
      // square(i + 1) goes between the end of GEMM2(i) and normalise(i): GEMM2(i + 1) then runs
      // while this tile is normalised, and no epilogue warp waits for the tensor pipe
      uint32_t xa[CP / 2], xb[CP / 2];
      if (ntiles > 0) square(0, xa);
      for (int it = 0; it < ntiles; it += 2) {
        mbar_wait(&bars->d2_full[0], it & 1);         // also: GEMM2(it) has read x^2
        if (it + 1 < ntiles) square(it + 1, xb);
        normalise(it, xa);
        if (it + 1 < ntiles) {
          mbar_wait(&bars->d2_full[0], (it + 1) & 1);
          if (it + 2 < ntiles) square(it + 2, xa);
          normalise(it + 1, xb);
        }
      }
    } else {
    // ===== epilogue: 8 warps, thread <-> TMEM lane <-> tile pixel, two column halves =====
    const int quad = warp & 3;
    const int half = (warp - 4) >> 2;
    const int r = quad * 32 + lane;
    const uint32_t lane_addr = (uint32_t)(quad * 32) << 16;
    const int li = r / p.tw, lj = r - li * p.tw;
    uint32_t it = 0;
    for (int tile = tile0; tile < tile1; tile += tile_step, ++it) {
      int n, i0, j0;
      tile_coords(tile, n, i0, j0);
      const int acc = it & 1;
      const int i2 = i0 + li, j2 = j0 + lj;
      const bool valid = i2 < p.ho && j2 < p.wo && !(p.debug & 2);
      const int64_t elem0 = (((int64_t)n * p.ho + i2) * p.wo + j2) * p.cout;
      const uint32_t dacc = tmem + acc * 256 + lane_addr;
      mbar_wait(&bars->d1_full[acc], (it >> 1) & 1);
      tc_fence_after();
      if constexpr (GDN)
        epi::gdn_tile<MODE, NC / 2, KO>(dacc, a2, r, half * (NC / 2), s_bias, s_beta, &bars->a2_full,
                                    &bars->d2_full[0], it, valid, p.out, p.out_f32 != 0,
                                    p.out_wide != 0, elem0, p.out_tma ? &map_out : nullptr,
                                    epi::TileOrigin{j0, i0, n}, threadIdx.x == 128, NC / 2);
      else
        epi::bias_tile(dacc, half, p.npad, p.cout, s_bias, p.act == TFC_ACT_RELU, valid, p.out,
                       p.out_f32 != 0, p.out_wide != 0, elem0);
      tc_fence_before();
      mbar_arrive(&bars->d1_empty[acc]);
    }
    if (GDN && p.out_tma && threadIdx.x == 128) bulk_wait_read();
    }
  } else if (warp == 0) {
    // ===== one-time TMA: resident weights and gamma =====
    if (lane == 0) {
      // TS: this thread also sends the staged y tiles off (issuing the three tensor stores takes
      // ~500 cycles and seeing them leave shared memory another ~1200: kept off the epilogue
      // warps)
      auto store_tile = [&](uint32_t j) {
        int n, i0, j0;
        tile_coords(tile0 + (int)j * tile_step, n, i0, j0);
        TR(j, 61);
        for (int kc = 0; kc < gdn_kblocks; ++kc)
          tma_store_4d(&map_out, a2 + kc * kABytes, kc * 64, j0, i0, n);
        bulk_commit();
        TR(j, 62);
        bulk_wait_read();
        mbar_arrive(&bars->y_free);
        TR(j, 60);
      };
      mbar_expect_tx(&bars->w_full, (KB + gdn_kblocks) * b_bytes);
      for (int kb = 0; kb < KB; ++kb)
        tma_load_3d(w_res + kb * b_slot, &map_w, &bars->w_full, 0, 0, kb);
      for (int kc = 0; kc < gdn_kblocks; ++kc)
        tma_load_2d(g_res + kc * b_slot, &map_g, &bars->w_full, kc * 64, 0);
      const uint32_t ntiles = tile1 > tile0 ? (tile1 - tile0 + tile_step - 1) / tile_step : 0;
      auto load_patch = [&](uint32_t it) {
        // ===== image patches: one TMA box per tile, borders zero-filled by the TMA unit =====
        int n, i0, j0;
        tile_coords(tile0 + (int)it * tile_step, n, i0, j0);
        const int pb = it & 1;
        mbar_expect_tx(&bars->p_full[pb], p.patch_bytes);
        // the box must start on a 16-byte boundary of the image row: round the first column
        // down, the builders skip the difference
        int c0 = (j0 * p.sd - p.pl) * CIN;
        c0 -= c0 & (p.per16 - 1);
        tma_load_3d(patches + pb * p.patch_slot, &map_img, &bars->p_full[pb], c0,
                    i0 * p.sd - p.pt, n);
      };
      if constexpr (TS) {
        // two duties, whichever is ready: the patch of the next tile (two buffers ahead of the
        // builders) and the store of the next staged y tile
        uint32_t np = p.patch ? 0 : ntiles, ns = 0;
        while (ns < ntiles || np < ntiles) {
          if (np < ntiles && mbar_test(&bars->p_empty[np & 1], ((np >> 1) & 1) ^ 1)) {
            load_patch(np);
            ++np;
          }
          if (ns < ntiles && mbar_test(&bars->y_full, ns & 1)) {
            store_tile(ns);
            ++ns;
          }
        }
      } else if (p.patch) {
        for (uint32_t it = 0; it < ntiles; ++it) {
          mbar_wait(&bars->p_empty[it & 1], ((it >> 1) & 1) ^ 1);
          load_patch(it);
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer =====
    if (lane == 0) {
      int stage = 0;
      uint32_t ph = 0, it = 0;
      auto issue_gdn = [&](uint32_t j) {
        mbar_wait(&bars->a2_full, j & 1);
        tc_fence_after();
        TR(j, 50);
        const uint32_t d2 = tmem + (j & 1) * ACC;
        if constexpr (TS) {
          // A from tensor memory; the last step adds beta: the constant A block against the beta
          // rows in W's last k-block.  (Issued in two or four column parts with a commit each, so
          // that the epilogue could start on the first columns early, GEMM2 takes as long per
          // part as it does whole — 89 cycles per MMA at N = 96 as at N = 192: r3m_*.)
          for (int kc = 0; kc < gdn_kblocks; ++kc) {
            const uint64_t db = make_desc_sw128(smem_u32(g_res + kc * b_slot));
#pragma unroll
            for (int k = 0; k < 4; ++k)
              umma_bf16_ts(d2, tmem + X2COL + (kc * 4 + k) * 8, db + (k * 2), idesc, (kc | k) != 0);
          }
          const uint64_t db = make_desc_sw128(smem_u32(w_res + (KB - 1) * b_slot));
          umma_bf16_ts(d2, tmem + ONESCOL, db + (uint64_t)(fold::beta_slice(K) / 8), idesc, 1);
          umma_commit(&bars->d2_full[0]);
          TR(j, 51);
        } else {
          for (int kc = 0; kc < gdn_kblocks; ++kc) {
            const uint64_t da = make_desc_sw128(smem_u32(a2 + kc * kABytes));
            const uint64_t db = make_desc_sw128(smem_u32(g_res + kc * b_slot));
#pragma unroll
            for (int k = 0; k < 4; ++k)
              umma_bf16(d2, da + (k * 2), db + (k * 2), idesc, (kc | k) != 0);
          }
          umma_commit(&bars->d2_full[0]);
        }
      };
      mbar_wait(&bars->w_full, 0);
      auto main_gemm = [&](uint32_t it) {
        const int acc = it & 1;
        mbar_wait(&bars->d1_empty[acc], ((it >> 1) & 1) ^ 1);
        tc_fence_after();
        TR(it, 52);
        const uint32_t d1 = tmem + acc * ACC;
#pragma unroll
        for (int kb = 0; kb < KB; ++kb) {
          const int nm = (K - kb * 64 >= 64) ? 4 : (K - kb * 64 + 15) / 16;
          mbar_wait(&bars->a_full[stage], ph);
          tc_fence_after();
          const uint64_t da = make_desc_sw128(smem_u32(smem + stage * kABytes));
          const uint64_t db = make_desc_sw128(smem_u32(w_res + kb * b_slot));
#pragma unroll
          for (int k = 0; k < 4; ++k)
            if (k < nm) umma_bf16(d1, da + (k * 2), db + (k * 2), idesc, (kb | k) != 0);
          umma_commit(&bars->a_empty[stage]);
          if (++stage == p.stages) { stage = 0; ph ^= 1; }
        }
        umma_commit(&bars->d1_full[acc]);
        TR(it, 53);
      };
      const uint32_t ntiles = tile1 > tile0 ? (tile1 - tile0 + tile_step - 1) / tile_step : 0;
      if constexpr (TS) {
        // In the order the epilogue makes things ready (square(it+1), normalise(it), ...):
        // GEMM2(it+1) as soon as tile it+1 is squared, the main GEMM of tile it+2 once tile it's
        // accumulator has been read by normalise().
        if (ntiles > 0) main_gemm(0);
        if (ntiles > 1) main_gemm(1);
        if (ntiles > 0) issue_gdn(0);
        for (it = 0; it < ntiles; ++it) {
          if (it + 1 < ntiles) issue_gdn(it + 1);
          if (it + 2 < ntiles) main_gemm(it + 2);
        }
      } else {
        for (it = 0; it < ntiles; ++it) {
          main_gemm(it);
          // GEMM2 of the previous tile goes AFTER this tile's main GEMM: the main GEMM is short
          // and its operands are ready long before the epilogue has squared the previous tile
          if (GDN && it > 0) issue_gdn(it - 1);
        }
        if (GDN && it > 0) issue_gdn(it - 1);
      }
    }
  } else {
    // ===== im2col builders (warps 2-3) =====
    const int t = threadIdx.x - 64;
    int stage = 0;
    uint32_t ph = 0;
    const uint8_t* patch = nullptr;
    auto block = [&](auto kbi, int n, int i0, int j0) {
      constexpr int KBI = decltype(kbi)::value;
      if constexpr (KBI < KB) {
        mbar_wait(&bars->a_empty[stage], ph ^ 1);
        uint8_t* sa = smem + stage * kABytes;
        if (p.debug & 1) {
        } else if (p.patch) {
          if (p.in_dtype == TFC_F32)
            build_block_patch<KH, KW, CIN, KBI, TS>(p, reinterpret_cast<const float*>(patch), t, sa);
          else if (p.in_dtype == TFC_U8)
            build_block_patch<KH, KW, CIN, KBI, TS>(p, patch, t, sa);
          else
            build_block_patch<KH, KW, CIN, KBI, TS>(p, reinterpret_cast<const __nv_bfloat16*>(patch), t, sa);
        } else if (p.in_dtype == TFC_F32)
          build_block<KH, KW, CIN, KBI, TS>(p, static_cast<const float*>(p.in), n, i0, j0, t, sa);
        else if (p.in_dtype == TFC_U8)
          build_block<KH, KW, CIN, KBI, TS>(p, static_cast<const uint8_t*>(p.in), n, i0, j0, t, sa);
        else
          build_block<KH, KW, CIN, KBI, TS>(p, static_cast<const __nv_bfloat16*>(p.in), n, i0, j0, t, sa);
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive(&bars->a_full[stage]);
        if (++stage == p.stages) { stage = 0; ph ^= 1; }
      }
    };
    uint32_t it = 0;
    for (int tile = tile0; tile < tile1; tile += tile_step, ++it) {
      int n, i0, j0;
      tile_coords(tile, n, i0, j0);
      const int pb = it & 1;
      if (p.patch) {
        mbar_wait(&bars->p_full[pb], (it >> 1) & 1);
        const int skip = ((j0 * p.sd - p.pl) * CIN) & (p.per16 - 1);
        patch = patches + pb * p.patch_slot + skip * p.esize;
      }
      block(std::integral_constant<int, 0>{}, n, i0, j0);
      block(std::integral_constant<int, 1>{}, n, i0, j0);
      block(std::integral_constant<int, 2>{}, n, i0, j0);
      block(std::integral_constant<int, 3>{}, n, i0, j0);
      if (p.patch) {
        __syncwarp();
        if (lane == 0) mbar_arrive(&bars->p_empty[pb]);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 512);
}

// w fp32 [K][cout] -> bf16 [KB][npad][64] (K-major blocks, zero padded)
// `folded`: rows K..K+2 hold the bias and rows beta_k..beta_k+2 beta, each as three bfloat16
// pieces hi + mid + lo (namespace fold)
__device__ __forceinline__ float bf16_piece(float v, int i) {
  for (int j = 0; j < i; ++j) v -= __bfloat162float(__float2bfloat16_rn(v));
  return v;
}
__global__ void repack_thin_weights_kernel(const float* __restrict__ w, int K, int cout, int npad,
                                           int kblocks, __nv_bfloat16* __restrict__ out,
                                           const float* __restrict__ bias,
                                           const float* __restrict__ beta, int folded, int beta_k) {
  const int total = kblocks * npad * 64;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int kk = i & 63;
    const int o = (i >> 6) % npad;
    const int k = (i / (64 * npad)) * 64 + kk;
    float v = (k < K && o < cout) ? w[(int64_t)k * cout + o] : 0.f;
    if (folded && o < cout) {
      if (k >= K && k < K + 3) v = bias ? bf16_piece(bias[o], k - K) : 0.f;
      if (k >= beta_k && k < beta_k + 3) v = bf16_piece(beta[o], k - beta_k);
    }
    out[i] = __float2bfloat16_rn(v);
  }
}

// gamma fp32 [k][c] -> bf16 [c (npad rows)][k]
__global__ void repack_gamma_kernel(const float* __restrict__ g, int c_count, int npad,
                                    __nv_bfloat16* __restrict__ out) {
  const int total = npad * c_count;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int k = i % c_count, c = i / c_count;
    out[i] = __float2bfloat16_rn(c < c_count ? g[(int64_t)k * c_count + c] : 0.f);
  }
}

// =========================================================================================
// Filter gradient of the thin layers on tcgen05 (training step, examples/bls2017.py:156-185)
// =========================================================================================
// In the normalised form of tfc_b200.h the gradient of a layer with 3 channels on one side is
//     G[(a, b, c)][o] = sum_{n,i,j} big[n, s*i + a - p, s*j + b - q, c] * small[n, i, j, o]
// (big = the 3-channel tensor: the image for the first analysis layer, the output gradient for
// the last synthesis layer; small = the 128-channel one on the coarse grid): a GEMM
//     G[K = kh*kw*3, cs] = im2col(big)^T [K x pixels] . small [pixels x cs]
// whose reduction dimension is the pixel index.  Its A operand is the SAME shared-memory tile the
// forward kernel above feeds its main GEMM with — 128 pixels x 64 im2col columns per 16 KB slab,
// 128-byte rows, 128-byte swizzle — read MN-major instead of K-major (the major bit of the
// instruction descriptor), so the forward kernel's builders are reused as they are; `small`
// arrives by TMA exactly as in conv_wgrad_umma.cu.  K = 243 needs two M = 128 accumulators
// (2 x cs <= 512 TMEM columns), which live for the whole CTA; at the end four warps add them to
// dw with float atomics (148 CTAs meet there).  Replaces the mma.sync kernel of conv_wgrad.cu for
// these layers: 0.46 ms -> see profiles/ for the 9x9x3 <-> 128 layers of bls2017 at batch 64.
constexpr int kWgThreads = 256;       // TMA, MMA, 2 builder warps, 4 epilogue warps

struct WgParams {
  Params b;                 // geometry of the forward-like view: (h, w) = big, (ho, wo) = small
  float* dw;
  int32_t cs;               // channels of `small`
  int32_t transposed;       // 0: dw[k][o] = G[k][o];  1: dw[(taps-1-tap)][o][c] = G[(tap, c)][o]
  int32_t ntiles_cta_step;
};

struct __align__(8) WgBarriers {
  uint64_t full[2], empty[2], p_full[2], p_empty[2];
  uint64_t d_full;
  uint32_t tmem_base;
  uint32_t pad;
};

__device__ __forceinline__ uint64_t make_desc_mn_sw128(uint32_t smem_addr, uint32_t lbo) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFFu) >> 4);
  d |= (uint64_t)((lbo >> 4) & 0x3FFFu) << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}

template <int KH, int KW, int CIN>
__global__ void __launch_bounds__(kWgThreads, 1)
conv_thin_wgrad_kernel(const __grid_constant__ CUtensorMap map_img,
                       const __grid_constant__ CUtensorMap map_small, const WgParams q) {
  constexpr int K = KH * KW * CIN;
  constexpr int KB = (K + 63) / 64;            // 16 KB slabs of the A tile
  constexpr int MH = (KB + 1) / 2;             // M = 128 halves of the accumulator
  const Params& p = q.b;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  const int ns = q.cs / 64;
  const int stage_bytes = (2 * MH + ns) * kABytes;
  uint8_t* patches = smem + 2 * stage_bytes;
  WgBarriers* bars = reinterpret_cast<WgBarriers*>(patches + 2 * p.patch_slot);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t tmem_cols = MH * q.cs <= 64 ? 64 : (MH * q.cs <= 128 ? 128 : (MH * q.cs <= 256 ? 256 : 512));

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&map_img);
    prefetch_tmap(&map_small);
    for (int s = 0; s < 2; ++s) {
      mbar_init(&bars->full[s], 1 + kBuilders / 32);
      mbar_init(&bars->empty[s], 1);
      mbar_init(&bars->p_full[s], 1);
      mbar_init(&bars->p_empty[s], kBuilders / 32);
    }
    mbar_init(&bars->d_full, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(&bars->tmem_base, tmem_cols);
  // slabs the builders never write (odd KB) must not hold NaN patterns: zero the A tiles once
  for (int i = threadIdx.x * 16; i < 2 * stage_bytes; i += kWgThreads * 16)
    *reinterpret_cast<uint4*>(smem + i) = make_uint4(0, 0, 0, 0);
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = bars->tmem_base;
  const int tile0 = (int)blockIdx.x, tile_step = (int)gridDim.x;

  auto tile_coords = [&](int tile, int& n, int& i0, int& j0) {
    const int per_img = p.tiles_h * p.tiles_w;
    n = tile / per_img;
    const int r = tile - n * per_img;
    const int ti = r / p.tiles_w;
    i0 = ti * p.th;
    j0 = (r - ti * p.tiles_w) * p.tw;
  };

  if (warp == 0) {
    if (lane == 0) {
      uint32_t it = 0;
      for (int tile = tile0; tile < p.num_tiles; tile += tile_step, ++it) {
        int n, i0, j0;
        tile_coords(tile, n, i0, j0);
        const int sb = it & 1;
        const uint32_t par = ((it >> 1) & 1) ^ 1;
        mbar_wait(&bars->p_empty[sb], par);
        mbar_expect_tx(&bars->p_full[sb], p.patch_bytes);
        int c0 = (j0 * p.sd - p.pl) * CIN;
        c0 -= c0 & (p.per16 - 1);
        tma_load_3d(patches + sb * p.patch_slot, &map_img, &bars->p_full[sb], c0, i0 * p.sd - p.pt, n);
        mbar_wait(&bars->empty[sb], par);
        uint8_t* sbase = smem + sb * stage_bytes + 2 * MH * kABytes;
        mbar_expect_tx(&bars->full[sb], (uint32_t)(ns * kABytes));
        for (int c = 0; c < ns; ++c) {
          asm volatile(
              "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes"
              " [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(smem_u32(sbase + c * kABytes)),
              "l"(reinterpret_cast<uint64_t>(&map_small)), "r"(smem_u32(&bars->full[sb])),
              "r"(c * 64), "r"(j0), "r"(i0), "r"(n)
              : "memory");
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      // kind::f16, bf16 x bf16 -> f32, A and B MN-major (bits 15, 16)
      const uint32_t idesc = make_idesc_bf16(128, q.cs) | (1u << 15) | (1u << 16);
      uint32_t it = 0;
      for (int tile = tile0; tile < p.num_tiles; tile += tile_step, ++it) {
        const int sb = it & 1;
        mbar_wait(&bars->full[sb], (it >> 1) & 1);
        tc_fence_after();
        const uint32_t sa = smem_u32(smem + sb * stage_bytes);
        const uint64_t db = make_desc_mn_sw128(sa + 2 * MH * kABytes, kABytes);
#pragma unroll
        for (int half = 0; half < MH; ++half) {
          const uint64_t da = make_desc_mn_sw128(sa + half * 2 * kABytes, kABytes);
#pragma unroll
          for (int k = 0; k < 8; ++k)          // 16 pixel rows = 2048 bytes per step
            umma_bf16(tmem + half * q.cs, da + (uint64_t)(k * 128), db + (uint64_t)(k * 128), idesc,
                      it != 0 || k != 0);
        }
        umma_commit(&bars->empty[sb]);
      }
      umma_commit(&bars->d_full);
    }
  } else if (warp < 4) {
    // ===== im2col builders: the forward kernel's, all KB blocks of a tile into one stage =====
    const int t = threadIdx.x - 64;
    uint32_t it = 0;
    for (int tile = tile0; tile < p.num_tiles; tile += tile_step, ++it) {
      int n, i0, j0;
      tile_coords(tile, n, i0, j0);
      const int sb = it & 1;
      mbar_wait(&bars->p_full[sb], (it >> 1) & 1);
      mbar_wait(&bars->empty[sb], ((it >> 1) & 1) ^ 1);
      const int skip = ((j0 * p.sd - p.pl) * CIN) & (p.per16 - 1);
      const __nv_bfloat16* patch = reinterpret_cast<const __nv_bfloat16*>(
          patches + sb * p.patch_slot + skip * p.esize);
      uint8_t* sa = smem + sb * stage_bytes;
      auto block = [&](auto kbi) {
        constexpr int KBI = decltype(kbi)::value;
        if constexpr (KBI < KB)
          build_block_patch<KH, KW, CIN, KBI>(p, patch, t, sa + KBI * kABytes);
      };
      block(std::integral_constant<int, 0>{});
      block(std::integral_constant<int, 1>{});
      block(std::integral_constant<int, 2>{});
      block(std::integral_constant<int, 3>{});
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(&bars->full[sb]);
        mbar_arrive(&bars->p_empty[sb]);
      }
    }
  } else {
    // ===== epilogue: warp w reads TMEM lanes 32*(w%4) .. +31 (= im2col column), all cs columns =====
    if (tile0 < p.num_tiles) {
      const int quad = warp & 3;
      mbar_wait(&bars->d_full, 0);
      tc_fence_after();
      constexpr int taps = KH * KW;
      for (int half = 0; half < MH; ++half) {
        const int m = half * 128 + quad * 32 + lane;          // (tap, c) = (m / CIN, m % CIN)
        for (int c0 = 0; c0 < q.cs; c0 += 32) {
          uint32_t v[32];
          tmem_ld_32x32(tmem + ((uint32_t)(quad * 32) << 16) + (uint32_t)(half * q.cs + c0), v);
          tmem_ld_wait();
          if (m < K) {
            const int tap = m / CIN, c = m - tap * CIN;
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              const int o = c0 + j;
              const int64_t at = q.transposed ? ((int64_t)(taps - 1 - tap) * q.cs + o) * CIN + c
                                              : (int64_t)m * q.cs + o;
              atomicAdd(q.dw + at, __uint_as_float(v[j]));
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


struct Plan {
  int kblocks, npad, stages;
  int th, tw;
  bool patch;               // TMA-staged image patches
  int patch_rows, patch_pitch, patch_slot, vec;
  size_t smem;
  bool gdn;
  bool ts;                  // GDN epilogue through tensor memory (debug bit 32 turns it off)
};

bool shape_ok(const tfc_conv_desc& d) {
  return d.cin == 3 && d.kh == d.kw && (d.kh == 5 || d.kh == 9);
}

bool make_plan(const tfc_conv_desc& d, const void* in, Plan* pl) {
  pl->gdn = d.act == TFC_ACT_GDN;
  const int K = d.kh * d.kw * d.cin;
  pl->ts = pl->gdn && d.out_dtype == TFC_BF16 && d.cout % 64 == 0 && fold::fits(K) &&
           !(tfc::debug_bits() & (32 | 16));
  pl->kblocks = (K + 63) / 64;
  pl->npad = (d.cout + 15) / 16 * 16;
  tfc::conv_pick_tile(d.ho, d.wo, &pl->th, &pl->tw);
  const size_t b_slot = ((size_t)pl->npad * 128 + 1023) & ~size_t(1023);
  const size_t fixed = 1024 + pl->kblocks * b_slot +
                       (pl->gdn ? (size_t)(d.cout / 64) * (b_slot + kABytes) : 0) +
                       2 * 256 * sizeof(float) + sizeof(Barriers) + 64;
  // image patch of one tile: ((th-1)*sd + kh) rows of ((tw-1)*sd + kw)*cin elements, the row
  // starting on a 16-byte boundary and padded to 16 bytes (+ one element so that a pair load may
  // straddle the end)
  const int esize = d.in_dtype == TFC_F32 ? 4 : (d.in_dtype == TFC_U8 ? 1 : 2);
  const int per16 = 16 / esize;
  pl->patch_rows = (pl->th - 1) * d.sd + d.kh;
  pl->patch_pitch = (((pl->tw - 1) * d.sd + d.kw) * d.cin + 1 + (per16 - 1) + per16 - 1) / per16 * per16;
  pl->patch_slot = (pl->patch_rows * pl->patch_pitch * esize + 16 + 127) & ~127;
  pl->vec = ((d.sd * d.cin) % 2 == 0 && (d.pad_left * d.cin) % 2 == 0) ? 2 : 1;
  // TMA needs 16-byte aligned rows in global memory and boxes of at most 256 elements per dim;
  // `in == nullptr` (workspace / support queries) assumes an aligned image
  const bool tma_ok = !(tfc::debug_bits() & 4) && ((size_t)d.w * d.cin * esize) % 16 == 0 && pl->patch_pitch <= 256 &&
                      pl->patch_rows <= 256 && (reinterpret_cast<uintptr_t>(in) & 15) == 0;
  for (int with_patch = tma_ok ? 1 : 0; with_patch >= 0; --with_patch) {
    const size_t extra = with_patch ? 2 * (size_t)pl->patch_slot : 0;
    for (int s = kMaxStages; s >= 2; --s) {
      if (fixed + extra + (size_t)s * kABytes <= 227 * 1024) {
        pl->stages = s;
        pl->patch = with_patch != 0;
        pl->smem = fixed + extra + (size_t)s * kABytes;
        return true;
      }
    }
  }
  return false;
}

size_t align256(size_t x) { return (x + 255) & ~size_t(255); }

template <int KH, int KW>
int launch(const tfc_conv_desc& d, const Plan& pl, const CUtensorMap& map_w,
           const CUtensorMap& map_g, const CUtensorMap& map_img, const CUtensorMap& map_out,
           const Params& p, int grid, cudaStream_t st) {
  void (*kern)(const CUtensorMap, const CUtensorMap, const CUtensorMap, const CUtensorMap,
               const Params) = nullptr;
  int threads = kThreads;
  if (!pl.gdn) {
    kern = conv_thin_kernel<0, 0, KH, KW, 3, false>;
  } else if (pl.ts) {
    threads = kThreadsTs;
    switch (d.cout / 32) {
      case 4: kern = conv_thin_kernel<1, 4, KH, KW, 3, true>; break;
      case 6: kern = conv_thin_kernel<1, 6, KH, KW, 3, true>; break;
      default: return tfc::fail(TFC_ERR_UNIMPLEMENTED, "thin-input GDN needs 128 or 192 filters");
    }
  } else {
    switch (d.cout / 32) {
      case 4: kern = conv_thin_kernel<1, 4, KH, KW, 3, false>; break;
      case 6: kern = conv_thin_kernel<1, 6, KH, KW, 3, false>; break;
      default: return tfc::fail(TFC_ERR_UNIMPLEMENTED, "thin-input GDN needs 128 or 192 filters");
    }
#ifdef TFC_A0_KNOCKOUT
    if (KH == 5 && d.cout == 192) {
      switch (tfc::debug_bits() >> 16) {
        case 1: kern = conv_thin_kernel<1, 6, KH, KW, 3, false, 1>; break;
        case 4: kern = conv_thin_kernel<1, 6, KH, KW, 3, false, 4>; break;
        case 8: kern = conv_thin_kernel<1, 6, KH, KW, 3, false, 8>; break;
        case 16: kern = conv_thin_kernel<1, 6, KH, KW, 3, false, 16>; break;
        case 32: kern = conv_thin_kernel<1, 6, KH, KW, 3, false, 32>; break;
        case 5: kern = conv_thin_kernel<1, 6, KH, KW, 3, false, 5>; break;
        case 21: kern = conv_thin_kernel<1, 6, KH, KW, 3, false, 21>; break;
        case 61: kern = conv_thin_kernel<1, 6, KH, KW, 3, false, 61>; break;
        default: break;
      }
    }
#endif
  }
  TFC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem));
  kern<<<grid, threads, pl.smem, st>>>(map_w, map_g, map_img, map_out, p);
  TFC_LAUNCHED("conv_thin_kernel");
  return TFC_OK;
}

}  // namespace

namespace tfc {

bool conv_thin_supported(const tfc_conv_desc& d, const char** why) {
  const char* dummy;
  if (!why) why = &dummy;
  if (d.su != 1) { *why = "thin-input path needs strides_up == 1"; return false; }
  if (d.out_dtype == TFC_U8) { *why = "thin-input path does not write uint8"; return false; }
  if (!shape_ok(d)) { *why = "thin-input path is built for 3 channels and 5x5 / 9x9 kernels"; return false; }
  if (d.act == TFC_ACT_IGDN) { *why = "thin-input path has no IGDN epilogue"; return false; }
  if (d.act == TFC_ACT_GDN && d.cout != 128 && d.cout != 192) {
    *why = "thin-input GDN needs 128 or 192 filters"; return false;
  }
  if (d.cout > 256) { *why = "cout must be <= 256"; return false; }
  if (d.sd > 8) { *why = "stride too large"; return false; }
  Plan pl;
  if (!make_plan(d, nullptr, &pl)) { *why = "shared memory"; return false; }
  if ((int64_t)d.n * d.ho * d.wo / 128 > (1 << 30)) { *why = "too many tiles"; return false; }
  return true;
}

size_t conv_thin_workspace(const tfc_conv_desc& d) {
  Plan pl;
  if (!make_plan(d, nullptr, &pl)) return 0;
  return align256((size_t)pl.kblocks * pl.npad * 64 * 2) + align256((size_t)pl.npad * pl.npad * 2) + 256;
}

int conv_thin(const tfc_conv_desc& d, const void* in, const float* w, const float* bias,
              const float* beta, const float* gamma, void* out, void* ws, size_t ws_bytes,
              cudaStream_t st) {
  Plan pl;
  if (!make_plan(d, in, &pl)) return fail(TFC_ERR_UNIMPLEMENTED, "thin-input conv: shared memory");
  if (ws == nullptr || ws_bytes < conv_thin_workspace(d))
    return fail(TFC_ERR_INVALID_ARGUMENT, "conv workspace too small: need %zu bytes",
                conv_thin_workspace(d));
  if ((reinterpret_cast<uintptr_t>(ws) & 255) || (reinterpret_cast<uintptr_t>(out) & 15))
    return fail(TFC_ERR_INVALID_ARGUMENT, "conv buffers must be 16-byte (workspace 256-byte) aligned");
  Params p{};
  p.n = d.n; p.h = d.h; p.w = d.w; p.ho = d.ho; p.wo = d.wo; p.cout = d.cout; p.npad = pl.npad;
  p.sd = d.sd; p.pt = d.pad_top; p.pl = d.pad_left;
  p.act = d.act; p.out_f32 = d.out_dtype == TFC_F32; p.in_dtype = d.in_dtype;
  p.esize = d.in_dtype == TFC_F32 ? 4 : (d.in_dtype == TFC_U8 ? 1 : 2);
  p.out_wide = epi::wide_stores_ok(out, d.cout, p.out_f32 != 0);
  p.in = in; p.bias = bias; p.beta = beta; p.out = out;
  p.stages = pl.stages;
  p.debug = debug_bits();
  p.th = pl.th; p.tw = pl.tw;
  p.patch = pl.patch; p.patch_pitch = pl.patch_pitch; p.patch_slot = pl.patch_slot; p.vec = pl.vec;
  p.per16 = 16 / p.esize;
  p.patch_bytes = pl.patch_rows * pl.patch_pitch * p.esize;
  p.tiles_h = (d.ho + p.th - 1) / p.th;
  p.tiles_w = (d.wo + p.tw - 1) / p.tw;
  p.num_tiles = d.n * p.tiles_h * p.tiles_w;
  {
    // q = umulhi(x, ceil(2^32 / d)) equals x / d for every x with x * d < 2^32
    auto rcp = [&](int divisor, int64_t max_x) -> uint32_t {
      if (divisor <= 1 || max_x * divisor >= (int64_t(1) << 32)) return 0;
      return (uint32_t)(((uint64_t(1) << 32) + divisor - 1) / divisor);
    };
    p.rcp_per_img = rcp(p.tiles_h * p.tiles_w, p.num_tiles);
    p.rcp_tiles_w = rcp(p.tiles_w, p.tiles_h * p.tiles_w);
  }

  const int K = d.kh * d.kw * d.cin;
  __nv_bfloat16* wb = static_cast<__nv_bfloat16*>(ws);
  const size_t wbytes = align256((size_t)pl.kblocks * pl.npad * 64 * 2);
  __nv_bfloat16* gb = reinterpret_cast<__nv_bfloat16*>(static_cast<uint8_t*>(ws) + wbytes);
  if (!(d.flags & TFC_CONV_WEIGHTS_STAGED)) {
    repack_thin_weights_kernel<<<(pl.kblocks * pl.npad * 64 + 255) / 256, 256, 0, st>>>(
        w, K, d.cout, pl.npad, pl.kblocks, wb, bias, beta, (pl.gdn && fold::fits(K)) ? 1 : 0,
        (pl.kblocks - 1) * 64 + fold::beta_slice(K) + 13);
    TFC_LAUNCHED("repack_thin_weights_kernel");
    if (pl.gdn) {
      repack_gamma_kernel<<<(pl.npad * d.cout + 255) / 256, 256, 0, st>>>(gamma, d.cout, pl.npad, gb);
      TFC_LAUNCHED("repack_gamma_kernel");
    }
  }
  CUtensorMap map_w, map_g, map_img, map_out;
  {
    cuuint64_t dims[3] = {64, (cuuint64_t)pl.npad, (cuuint64_t)pl.kblocks};
    cuuint64_t str[2] = {128, (cuuint64_t)pl.npad * 128};
    cuuint32_t box[3] = {64, (cuuint32_t)pl.npad, 1};
    int rc = make_tensor_map_bf16(&map_w, wb, 3, dims, str, box);
    if (rc) return rc;
  }
  if (pl.gdn) {
    cuuint64_t dims[2] = {(cuuint64_t)d.cout, (cuuint64_t)pl.npad};
    cuuint64_t str[1] = {(cuuint64_t)d.cout * 2};
    cuuint32_t box[2] = {64, (cuuint32_t)pl.npad};
    int rc = make_tensor_map_bf16(&map_g, gb, 2, dims, str, box);
    if (rc) return rc;
  } else {
    map_g = map_w;
  }
  if (pl.patch) {
    const cuuint64_t es = p.esize;
    cuuint64_t dims[3] = {(cuuint64_t)d.w * d.cin, (cuuint64_t)d.h, (cuuint64_t)d.n};
    cuuint64_t str[2] = {(cuuint64_t)d.w * d.cin * es, (cuuint64_t)d.h * d.w * d.cin * es};
    cuuint32_t box[3] = {(cuuint32_t)pl.patch_pitch, (cuuint32_t)pl.patch_rows, 1};
    int rc = make_tensor_map_plain(&map_img, d.in_dtype, const_cast<void*>(in), 3, dims, str, box);
    if (rc) return rc;
  } else {
    map_img = map_w;
  }
  p.out_tma = pl.gdn && !p.out_f32 && d.cout % 64 == 0 && !(debug_bits() & 16);
  if (p.out_tma) {
    cuuint64_t dims[4] = {(cuuint64_t)d.cout, (cuuint64_t)d.wo, (cuuint64_t)d.ho, (cuuint64_t)d.n};
    cuuint64_t str[3] = {(cuuint64_t)d.cout * 2, (cuuint64_t)d.wo * d.cout * 2,
                         (cuuint64_t)d.ho * d.wo * d.cout * 2};
    cuuint32_t box[4] = {64, (cuuint32_t)p.tw, (cuuint32_t)p.th, 1};
    int rc = make_tensor_map_bf16(&map_out, out, 4, dims, str, box);
    if (rc) return rc;
  } else {
    map_out = map_w;
  }
  // the resident operands cost one 50-120 KB load per CTA (and a tile takes ~4 us): never fewer
  // than 32 tiles per CTA
  p.chunk = tiles_per_cta() ? std::max(tiles_per_cta(), 32) : 0;
  const int grid = p.chunk ? (p.num_tiles + p.chunk - 1) / p.chunk
                           : std::min(p.num_tiles, persistent_grid());
  if (d.kh == 5) return launch<5, 5>(d, pl, map_w, map_g, map_img, map_out, p, grid, st);
  return launch<9, 9>(d, pl, map_w, map_g, map_img, map_out, p, grid, st);
}

}  // namespace tfc

// ---- thin-layer filter gradient (kernel above) -------------------------------------------
namespace {

struct WgPlan {
  tfc_conv_desc fwd;        // the forward-like view the builders walk
  const void* big;
  const void* small;
  int cs, transposed;
  Plan pl;
  size_t smem;
};

bool make_wg_plan(const tfc_conv_desc& d, const void* x, const void* dy, WgPlan* w) {
  if (d.sd > 1 && d.su > 1) return false;
  const bool up = d.su > 1;
  const int cb = up ? d.cout : d.cin, cs = up ? d.cin : d.cout;
  if (cb != 3 || d.kh != d.kw || (d.kh != 5 && d.kh != 9)) return false;
  if (cs % 64 || cs < 64 || cs > 256) return false;
  const int K = d.kh * d.kw * 3, MH = ((K + 63) / 64 + 1) / 2;
  if (MH * cs > 512) return false;
  tfc_conv_desc f = d;
  f.h = up ? d.ho : d.h;   f.w = up ? d.wo : d.w;
  f.ho = up ? d.h : d.ho;  f.wo = up ? d.w : d.wo;
  f.cin = 3; f.cout = cs;
  f.sd = up ? d.su : d.sd; f.su = 1;
  f.pad_top = up ? d.kh - 1 - d.pad_top : d.pad_top;
  f.pad_left = up ? d.kw - 1 - d.pad_left : d.pad_left;
  f.act = TFC_ACT_NONE; f.in_dtype = TFC_BF16; f.out_dtype = TFC_F32;
  if (f.sd > 8 || (int64_t)f.n * f.ho * f.wo < 128) return false;
  w->fwd = f;
  w->big = up ? dy : x;
  w->small = up ? x : dy;
  w->cs = cs;
  w->transposed = up ? 1 : 0;
  if (!make_plan(f, w->big, &w->pl) || !w->pl.patch) return false;
  const size_t stage = (size_t)(2 * MH + cs / 64) * kABytes;
  w->smem = 1024 + 2 * stage + 2 * (size_t)w->pl.patch_slot + sizeof(WgBarriers) + 64;
  return w->smem <= 227 * 1024;
}

template <int KH, int KW>
int launch_wg(const WgPlan& w, const CUtensorMap& map_img, const CUtensorMap& map_small,
              const WgParams& q, int grid, cudaStream_t st) {
  auto kern = conv_thin_wgrad_kernel<KH, KW, 3>;
  TFC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)w.smem));
  kern<<<grid, kWgThreads, w.smem, st>>>(map_img, map_small, q);
  TFC_LAUNCHED("conv_thin_wgrad_kernel");
  return TFC_OK;
}

}  // namespace

namespace tfc {

bool conv_thin_wgrad_supported(const tfc_conv_desc& d, const void* x, const void* dy) {
  WgPlan w;
  return make_wg_plan(d, x, dy, &w);
}

// dw must be zero on entry (tfc_signal_conv2d_wgrad clears it): the CTAs add their partial sums
int conv_thin_wgrad(const tfc_conv_desc& d, const void* x, const void* dy, float* dw,
                    cudaStream_t st) {
  WgPlan w;
  if (!make_wg_plan(d, x, dy, &w))
    return fail(TFC_ERR_UNIMPLEMENTED, "thin-layer filter gradient: shape not supported");
  const tfc_conv_desc& f = w.fwd;
  const Plan& pl = w.pl;
  WgParams q{};
  Params& p = q.b;
  p.n = f.n; p.h = f.h; p.w = f.w; p.ho = f.ho; p.wo = f.wo; p.cout = f.cout; p.npad = f.cout;
  p.sd = f.sd; p.pt = f.pad_top; p.pl = f.pad_left;
  p.in_dtype = TFC_BF16; p.esize = 2; p.per16 = 8;
  p.th = pl.th; p.tw = pl.tw;
  p.patch = 1; p.patch_pitch = pl.patch_pitch; p.patch_slot = pl.patch_slot; p.vec = pl.vec;
  p.patch_bytes = pl.patch_rows * pl.patch_pitch * p.esize;
  p.tiles_h = (f.ho + p.th - 1) / p.th;
  p.tiles_w = (f.wo + p.tw - 1) / p.tw;
  p.num_tiles = f.n * p.tiles_h * p.tiles_w;
  q.dw = dw; q.cs = w.cs; q.transposed = w.transposed;

  CUtensorMap map_img, map_small;
  {
    cuuint64_t dims[3] = {(cuuint64_t)f.w * 3, (cuuint64_t)f.h, (cuuint64_t)f.n};
    cuuint64_t str[2] = {(cuuint64_t)f.w * 3 * 2, (cuuint64_t)f.h * f.w * 3 * 2};
    cuuint32_t box[3] = {(cuuint32_t)pl.patch_pitch, (cuuint32_t)pl.patch_rows, 1};
    int rc = make_tensor_map_plain(&map_img, TFC_BF16, const_cast<void*>(w.big), 3, dims, str, box);
    if (rc) return rc;
  }
  {
    const cuuint64_t C = w.cs;
    cuuint64_t dims[4] = {C, (cuuint64_t)f.wo, (cuuint64_t)f.ho, (cuuint64_t)f.n};
    cuuint64_t str[3] = {C * 2, (cuuint64_t)f.wo * C * 2, (cuuint64_t)f.ho * f.wo * C * 2};
    cuuint32_t box[4] = {64, (cuuint32_t)p.tw, (cuuint32_t)p.th, 1};
    int rc = make_tensor_map_bf16(&map_small, const_cast<void*>(w.small), 4, dims, str, box);
    if (rc) return rc;
  }
  const int grid = std::min(p.num_tiles, num_sms());
  if (d.kh == 5) return launch_wg<5, 5>(w, map_img, map_small, q, grid, st);
  return launch_wg<9, 9>(w, map_img, map_small, q, grid, st);
}

}  // namespace tfc

extern "C" int tfc_debug_a0_trace(unsigned long long* dst, int n) {
#if TFC_A0_STAMPS
  if (n > 8 * 64) n = 8 * 64;
  return (int)cudaMemcpyFromSymbol(dst, g_a0_trace, sizeof(unsigned long long) * n);
#else
  (void)dst; (void)n;
  return (int)cudaErrorNotSupported;   // built with -DTFC_A0_STAMPS=0
#endif
}
