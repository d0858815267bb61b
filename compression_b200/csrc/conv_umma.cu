// tcgen05 / TMEM / TMA implicit-GEMM SignalConv2D with fused bias + ReLU / GDN / IGDN epilogue.
//
// Replaces, for the layers that carry the codecs' FLOPs, the stock TF kernels the reference
// dispatches to (/root/reference/tensorflow_compression/python/layers/signal_conv.py:516-518
// tf.nn.conv2d, :693-695 tf.nn.conv2d_backprop_input, :804 bias_add) together with the GDN
// activation that follows them (gdn.py:183-202: 1x1 conv of x^2 with gamma, bias_add beta,
// rsqrt/sqrt, multiply) — one kernel per layer instead of six TF ops.
//
// Formulation.  GEMM view per CTA tile: D[128 pixels, N=cout] +=
// A[128, 64] * B[64, N] for every (kernel tap, 64-channel chunk).
//   * A tiles come straight from the NHWC bf16 activation by TMA: the tensor map views it as
//     (s*C, W/s, s, H/s, N) for a stride-s correlation, so a tap is a dense box of one parity
//     plane and out-of-image taps are zero-filled by the TMA unit (no im2col buffer, no
//     padding copy).  An upsampled (transposed) convolution is split into su*su output
//     phases, each a stride-1 correlation over the un-stuffed input with its own tap subset —
//     no multiplications by stuffed zeros.
//   * B tiles are the weights repacked once per call to bf16 [tap][cout][cin] (K-major).
//   * Accumulators live in TMEM (fp32).  Warp roles: warp 0 TMA producer, warp 1 MMA issuer
//     (one elected lane issues tcgen05.mma), warps 2-5 epilogue (tcgen05.ld).
//   * GDN: the epilogue squares x = D1 + bias into a K-major bf16 tile in shared memory, the
//     MMA warp runs a second GEMM D2 = x^2 * gamma (gamma tiles ride the same TMA ring), and
//     the epilogue finishes y = x * rsqrt(beta + D2) (IGDN: * sqrt).
#include <cuda.h>
#include <cuda_bf16.h>

#include <algorithm>
#include <mutex>
#include <vector>

#include "common.cuh"
#include "conv_common.cuh"
#include "conv_epilogue.cuh"
#include "umma_ptx.cuh"

namespace {

using namespace tfc::ptx;
namespace epi = tfc::epi;

constexpr int kTileM = 128;          // output pixels per CTA tile (UMMA M)
constexpr int kBlockK = 64;          // channels per k-block: 64 bf16 = one 128-byte swizzle row
constexpr int kStages = 6;          // most stages of the operand ring (Params::stages of them used)
constexpr int kMaxTaps = 81;         // 9x9
constexpr int kMaxPhases = 16;       // 4x4 upsampling
constexpr int kThreads = 64 + tfc::epi::kEpilogueThreads;   // producer warp, MMA warp, 8 epilogue warps
constexpr int kABytes = tfc::epi::kABytes;     // 16 KB

struct Tap {            // one (kernel tap) of one output phase
  int8_t dh, dw;        // offset of the input box relative to the tile origin (in box rows/cols)
  int8_t ph, pw;        // parity plane (stride-s correlation): coordinates in the `s` dims
  int16_t widx;         // index of the weight slab (a*kw + b)
  int16_t pad;
};

struct Params {
  int32_t n, ho, wo, cout;          // output tensor [n, ho, wo, cout]
  int32_t cin;
  int32_t su;                       // output phases per dimension
  int32_t hs, ws;                   // per-phase output sub-grid (ho/su, wo/su)
  int32_t th, tw;                   // tile shape, th*tw == 128
  int32_t tiles_h, tiles_w;
  int32_t num_tiles;
  int32_t chunk;                    // tiles per CTA (0: persistent grid)
  int32_t cs;                       // cluster size: the CTAs of a cluster share the B loads
  int32_t stages;                   // stages of the TMA operand ring in use (2..kStages)
  int32_t pair;                     // CTA-pair MMAs (cta_group::2, M = 256 over the cluster of 2)
  int32_t npad;                     // UMMA N (cout rounded up to 16)
  int32_t act;
  int32_t out_f32;
  int32_t out_wide;                 // 256-bit stores allowed
  int32_t out_tma;                  // GDN output leaves by TMA tile stores
  int32_t phase_start[kMaxPhases + 1];
  Tap taps[kMaxTaps];
  const float* bias;
  const float* beta;
  void* out;
};

struct OutMaps {          // one output tensor map per phase of an upsampling layer (su <= 2)
  CUtensorMap m[4];
};

struct __align__(8) Barriers {
  uint64_t full[kStages], empty[kStages];
  uint64_t d1_full[2], d1_empty[2];
  uint64_t a2_full, d2_full;
  uint32_t tmem_base;
  uint32_t pad;
};

constexpr int kGdnSlot = 4;   // the GDN GEMM of tile i is issued before k-block 4 of tile i+1
                              // (2 / 6 / 8 / 10 measured the same: gpurun r3x, A1 / A2 / S1 / S2)

// MODE 0: bias (+ReLU).  MODE 1 / 2: GDN / IGDN with cout == 32 * NC.
//
// GDN pipeline (per CTA, tiles i = 0, 1, ...; two TMEM accumulators used alternately):
//   MMA      : main(i) -> acc[i&1];   GEMM2(i-1) = x^2 * gamma -> acc[(i-1)&1] IN PLACE, slotted
//              in before k-block kGdnSlot of main(i)
//   epilogue : phase 1(i): x = D1 + bias kept as packed bf16 in registers, x^2 -> smem (K-major)
//              phase 2(i): y = x * rsqrt(beta + D2)   (IGDN: * sqrt), overlapping main(i+1)
// PAIR: CTA-pair MMAs (tcgen05.mma.cta_group::2, see umma_ptx.cuh).  A separate instantiation: a
// kernel that contains cta_group::2 instructions can only be launched as a cluster.
template <int MODE, int NC, bool PAIR>
__global__ void __launch_bounds__(kThreads, 1)
conv_umma_kernel(const __grid_constant__ CUtensorMap map_a,
                 const __grid_constant__ CUtensorMap map_w,
                 const __grid_constant__ CUtensorMap map_g,
                 const __grid_constant__ OutMaps maps_out, const Params p) {
  constexpr bool GDN = MODE != 0;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // carve: [stages x (A 16 KB | B npad*128 B)] [A2: cout/64 x 16 KB] [bias|beta] [barriers]
  // aligned with pointer arithmetic on the __shared__ array (not through an integer): the
  // compiler then knows every pointer derived from `smem` is a shared-memory address and emits
  // LDS / STS instead of generic LD / ST (twice the wavefronts for a broadcast 128-bit load)
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  // B bytes one CTA holds per stage: all N rows, or its half of them when the pair shares B
  const int b_bytes = (PAIR ? p.npad / 2 : p.npad) * kBlockK * 2;
  const int stage_bytes = kABytes + ((b_bytes + 1023) & ~1023);
  uint8_t* a2 = smem + p.stages * stage_bytes;
  constexpr int gdn_kblocks = GDN ? NC / 2 : 0;
  float* s_bias = reinterpret_cast<float*>(a2 + gdn_kblocks * kABytes);
  float* s_beta = s_bias + 256;
  Barriers* bars = reinterpret_cast<Barriers*>(s_beta + 256);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int kchunks = p.cin / kBlockK;

  for (int i = threadIdx.x; i < 256; i += kThreads) {
    s_bias[i] = (p.bias && i < p.cout) ? p.bias[i] : 0.f;
    s_beta[i] = (GDN && i < p.cout) ? p.beta[i] : 1.f;
  }
  if (warp == 0 && lane == 0) {
    prefetch_tmap(&map_a);
    prefetch_tmap(&map_w);
    if (GDN) prefetch_tmap(&map_g);
    // pair mode: only the leader's MMA warp commits (one multicast arrival per CTA), and the
    // leader's d1_empty / a2_full collect the epilogue threads of BOTH CTAs
    const int epi_arrivals = epi::kEpilogueThreads * (PAIR ? 2 : 1);
    for (int s = 0; s < kStages; ++s) { mbar_init(&bars->full[s], 1); mbar_init(&bars->empty[s], PAIR ? 1 : p.cs); }
    for (int s = 0; s < 2; ++s) { mbar_init(&bars->d1_full[s], 1); mbar_init(&bars->d1_empty[s], epi_arrivals); }
    mbar_init(&bars->a2_full, epi_arrivals);
    mbar_init(&bars->d2_full, 1);
    fence_barrier_init();
  }
  if (warp == 1) { if (PAIR) tmem_alloc_pair(&bars->tmem_base, 512); else tmem_alloc(&bars->tmem_base, 512); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = bars->tmem_base;
  // Clusters (cs > 1): the cs CTAs of a cluster walk cs ADJACENT tiles in lockstep — same output
  // phase, hence the same taps in the same order — so each loads npad/cs rows of every weight
  // (and gamma) tile and multicasts them to all: the B share of the L2 -> shared-memory traffic
  // drops by cs.  A stage may be refilled once the MMAs of ALL CTAs have read it: every MMA warp
  // commits to the `empty` barrier of every CTA (count cs).
  const int rank = p.cs > 1 ? (int)cluster_ctarank() : 0;
  const uint16_t mc_mask = (uint16_t)((1u << p.cs) - 1);
  const int b_rows = p.npad / p.cs;
  if (p.cs > 1) cluster_sync();      // the peers' barriers are initialised
  // tile groups g, g + step, ... (a group = cs adjacent tiles, this CTA takes tile g + rank).
  // persistent (chunk == 0): one group per cluster and round; chunked: `chunk` consecutive groups
  // per cluster, many more CTAs than SMs, so the hardware places them on whatever SMs are free
  // (used when the range coder of another batch occupies part of the GPU).  A tile index past
  // the end (ragged last group) is a dummy: it repeats the last tile with its stores masked.
  const int cluster_id = (int)blockIdx.x / p.cs;
  const int tile0 = p.chunk ? cluster_id * p.chunk * p.cs : cluster_id * p.cs;
  const int tile1 = p.chunk ? min(tile0 + p.chunk * p.cs, p.num_tiles) : p.num_tiles;
  const int tile_step = p.chunk ? p.cs : (int)gridDim.x;
  const uint32_t idesc = make_idesc_bf16(PAIR ? 2 * kTileM : kTileM, p.npad);

  auto tile_coords = [&](int tile, int& phase, int& n, int& i0, int& j0) {
    const int per_img = p.tiles_h * p.tiles_w;
    const int per_phase = per_img * p.n;
    phase = tile / per_phase;
    int r = tile - phase * per_phase;
    n = r / per_img;
    r -= n * per_img;
    const int ti = r / p.tiles_w;
    i0 = ti * p.th;
    j0 = (r - ti * p.tiles_w) * p.tw;
  };

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      int stage = 0;
      uint32_t ph = 0, it = 0;
      auto load_gamma = [&]() {
        for (int kc = 0; kc < gdn_kblocks; ++kc) {
          mbar_wait(&bars->empty[stage], ph ^ 1);
          uint8_t* sa = smem + stage * stage_bytes;
          if (PAIR) {
            // each CTA loads ITS half of gamma's rows; both complete on the leader's barrier
            if (rank == 0) mbar_expect_tx(&bars->full[stage], 2 * b_bytes);
            tma_load_2d_pair(sa + kABytes, &map_g, leader_u32(&bars->full[stage]), kc * kBlockK,
                             rank * b_rows);
            if (++stage == p.stages) { stage = 0; ph ^= 1; }
            continue;
          }
          mbar_expect_tx(&bars->full[stage], b_bytes);
          if (p.cs > 1)
            tma_load_2d_mc(sa + kABytes + rank * b_rows * 128, &map_g, &bars->full[stage],
                           kc * kBlockK, rank * b_rows, mc_mask);
          else
            tma_load_2d(sa + kABytes, &map_g, &bars->full[stage], kc * kBlockK, 0);
          if (++stage == p.stages) { stage = 0; ph ^= 1; }
        }
      };
      for (int g = tile0; g < tile1; g += tile_step, ++it) {
        int phase, n, i0, j0;
        tile_coords(min(g + rank, p.num_tiles - 1), phase, n, i0, j0);
        const int nk = (p.phase_start[phase + 1] - p.phase_start[phase]) * kchunks;
        const int slot = min(kGdnSlot, nk);
        int kb = 0;
        for (int t = p.phase_start[phase]; t < p.phase_start[phase + 1]; ++t) {
          const Tap tap = p.taps[t];
          for (int kc = 0; kc < kchunks; ++kc, ++kb) {
            if (GDN && it > 0 && kb == slot) load_gamma();
            mbar_wait(&bars->empty[stage], ph ^ 1);
            uint8_t* sa = smem + stage * stage_bytes;
            if (PAIR) {
              // this CTA's 128 pixels of A and its half of the weight rows, signalled to the leader
              const uint32_t lb = leader_u32(&bars->full[stage]);
              if (rank == 0) mbar_expect_tx(&bars->full[stage], 2 * (kABytes + b_bytes));
              tma_load_5d_pair(sa, &map_a, lb, tap.pw * p.cin + kc * kBlockK, j0 + tap.dw, tap.ph,
                               i0 + tap.dh, n);
              tma_load_3d_pair(sa + kABytes, &map_w, lb, kc * kBlockK, rank * b_rows, tap.widx);
              if (++stage == p.stages) { stage = 0; ph ^= 1; }
              continue;
            }
            mbar_expect_tx(&bars->full[stage], kABytes + b_bytes);
            tma_load_5d(sa, &map_a, &bars->full[stage], tap.pw * p.cin + kc * kBlockK,
                        j0 + tap.dw, tap.ph, i0 + tap.dh, n);
            if (p.cs > 1)
              tma_load_3d_mc(sa + kABytes + rank * b_rows * 128, &map_w, &bars->full[stage],
                             kc * kBlockK, rank * b_rows, tap.widx, mc_mask);
            else
              tma_load_3d(sa + kABytes, &map_w, &bars->full[stage], kc * kBlockK, 0, tap.widx);
            if (++stage == p.stages) { stage = 0; ph ^= 1; }
          }
        }
        if (GDN && it > 0 && slot == nk) load_gamma();
      }
      if (GDN && it > 0) load_gamma();
    }
  } else if (warp == 1) {
    // ===== MMA issuer (pair mode: the leader CTA issues for both) =====
    if (lane == 0 && (!PAIR || rank == 0)) {
      int stage = 0;
      uint32_t ph = 0, it = 0;
      const uint16_t pair_mask = 0x3;
      auto mma = [&](uint32_t d, uint64_t da, uint64_t db, uint32_t accumulate) {
        if (PAIR) umma_bf16_pair(d, da, db, idesc, accumulate);
        else umma_bf16(d, da, db, idesc, accumulate);
      };
      // stage consumed: the producer(s) may refill it
      auto release = [&](uint64_t* bar) {
        if (PAIR) umma_commit_pair(bar, pair_mask);
        else if (p.cs > 1) umma_commit_mc(bar, mc_mask);
        else umma_commit(bar);
      };
      // accumulator finished: the epilogue(s) may read it
      auto publish = [&](uint64_t* bar) {
        if (PAIR) umma_commit_pair(bar, pair_mask); else umma_commit(bar);
      };
      // barriers the epilogue threads arrive on (those of the peer CTA too in pair mode)
      auto wait_epilogue = [&](uint64_t* bar, uint32_t parity) {
        if (PAIR) mbar_wait_cluster(bar, parity); else mbar_wait(bar, parity);
      };
      // GEMM2 of tile j: acc[j & 1] = x^2 (smem a2) * gamma, overwriting D1 (already consumed)
      auto issue_gdn = [&](uint32_t j) {
        wait_epilogue(&bars->a2_full, j & 1);
        tc_fence_after();
        const uint32_t d2 = tmem + (j & 1) * 256;
        for (int kc = 0; kc < gdn_kblocks; ++kc) {
          mbar_wait(&bars->full[stage], ph);
          tc_fence_after();
          const uint32_t sb = smem_u32(smem + stage * stage_bytes) + kABytes;
          const uint64_t da = make_desc_sw128(smem_u32(a2 + kc * kABytes));
          const uint64_t db = make_desc_sw128(sb);
#pragma unroll
          for (int k = 0; k < kBlockK / 16; ++k)
            mma(d2, da + (k * 2), db + (k * 2), (kc | k) != 0);
          release(&bars->empty[stage]);
          if (++stage == p.stages) { stage = 0; ph ^= 1; }
        }
        publish(&bars->d2_full);
      };
      for (int g = tile0; g < tile1; g += tile_step, ++it) {
        int phase, n, i0, j0;
        tile_coords(min(g + rank, p.num_tiles - 1), phase, n, i0, j0);
        const int acc = it & 1;
        wait_epilogue(&bars->d1_empty[acc], ((it >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t d1 = tmem + acc * 256;
        const int nk = (p.phase_start[phase + 1] - p.phase_start[phase]) * kchunks;
        const int slot = min(kGdnSlot, nk);
        for (int kb = 0; kb < nk; ++kb) {
          if (GDN && it > 0 && kb == slot) issue_gdn(it - 1);
          mbar_wait(&bars->full[stage], ph);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + stage * stage_bytes);
          const uint64_t da = make_desc_sw128(sa), db = make_desc_sw128(sa + kABytes);
#pragma unroll
          for (int k = 0; k < kBlockK / 16; ++k)
            mma(d1, da + (k * 2), db + (k * 2), (kb | k) != 0);
          release(&bars->empty[stage]);
          if (++stage == p.stages) { stage = 0; ph ^= 1; }
        }
        if (GDN && it > 0 && slot == nk) issue_gdn(it - 1);
        publish(&bars->d1_full[acc]);
      }
      if (GDN && it > 0) issue_gdn(it - 1);
    }
  } else {
    // ===== epilogue: 8 warps, thread <-> TMEM lane <-> tile pixel, two column halves =====
    const int quad = warp & 3;
    const int half = (warp - 2) >> 2;
    const int r = quad * 32 + lane;                  // pixel within the tile / TMEM lane
    const uint32_t lane_addr = (uint32_t)(quad * 32) << 16;
    uint32_t it = 0;
    for (int g = tile0; g < tile1; g += tile_step, ++it) {
      int phase, n, i0, j0;
      const bool dummy = g + rank >= p.num_tiles;
      tile_coords(min(g + rank, p.num_tiles - 1), phase, n, i0, j0);
      const int acc = it & 1;
      const int li = r / p.tw, lj = r - li * p.tw;
      const int i2 = i0 + li, j2 = j0 + lj;
      const bool valid = i2 < p.hs && j2 < p.ws && !dummy;
      const int pi = phase / p.su, pj = phase - pi * p.su;
      const int64_t pix = ((int64_t)n * p.ho + (i2 * p.su + pi)) * p.wo + (j2 * p.su + pj);
      const uint32_t dacc = tmem + acc * 256 + lane_addr;
      mbar_wait(&bars->d1_full[acc], (it >> 1) & 1);
      tc_fence_after();
      const int64_t elem0 = pix * p.cout;
      if constexpr (GDN)
        epi::gdn_tile<MODE, NC / 2>(dacc, a2, r, half * (NC / 2), s_bias, s_beta, &bars->a2_full,
                                    &bars->d2_full, it, valid, p.out, p.out_f32 != 0,
                                    p.out_wide != 0, elem0,
                                    p.out_tma ? &maps_out.m[phase] : nullptr,
                                    epi::TileOrigin{j0, i0, dummy ? -1 : n}, threadIdx.x == 64,
                                    NC / 2, PAIR ? leader_u32(&bars->a2_full) : 0u);
      else
        epi::bias_tile(dacc, half, p.npad, p.cout, s_bias, p.act == TFC_ACT_RELU, valid, p.out,
                       p.out_f32 != 0, p.out_wide != 0, elem0);
      tc_fence_before();
      if (PAIR) mbar_arrive_cluster(leader_u32(&bars->d1_empty[acc]));
      else mbar_arrive(&bars->d1_empty[acc]);
    }
    if (GDN && p.out_tma && threadIdx.x == 64) bulk_wait_read();
  }
  tc_fence_before();
  __syncthreads();
  if (PAIR) {
    cluster_sync();                  // both CTAs are done with the pair's tensor memory
    if (warp == 1) tmem_dealloc_pair(tmem, 512);
  } else if (warp == 1) {
    tmem_dealloc(tmem, 512);
  }
  if (p.cs > 1) cluster_sync();      // no CTA leaves while a peer may still signal its barriers
}


// ---- weight repack: fp32 [tap][cin][cout] -> bf16 [tap][npad][cin]; gamma [k][c] -> [c][k] -----
__global__ void repack_weights_kernel(const float* __restrict__ w, int taps, int cin, int cout,
                                      int npad, __nv_bfloat16* __restrict__ out) {
  const int64_t total = (int64_t)taps * npad * cin;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % cin);
    const int o = (int)((i / cin) % npad);
    const int t = (int)(i / ((int64_t)cin * npad));
    const float v = o < cout ? w[((int64_t)t * cin + c) * cout + o] : 0.f;
    out[i] = __float2bfloat16_rn(v);
  }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, []() {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) ==
            cudaSuccess && q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

int make_map_any(CUtensorMap* m, CUtensorMapDataType dtype, CUtensorMapSwizzle swizzle, void* base,
                 int rank, const cuuint64_t* dims, const cuuint64_t* strides_bytes,
                 const cuuint32_t* box) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) return tfc::fail(TFC_ERR_CUDA, "cuTensorMapEncodeTiled entry point not found");
  cuuint32_t ones[5] = {1, 1, 1, 1, 1};
  CUresult r = fn(m, dtype, (cuuint32_t)rank, base, dims, strides_bytes, box, ones,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return tfc::fail(TFC_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  return TFC_OK;
}

int make_map(CUtensorMap* m, void* base, int rank, const cuuint64_t* dims,
             const cuuint64_t* strides_bytes, const cuuint32_t* box) {
  return make_map_any(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, CU_TENSOR_MAP_SWIZZLE_128B, base, rank,
                      dims, strides_bytes, box);
}

inline int floor_div(int a, int b) { return (a >= 0) ? a / b : -((-a + b - 1) / b); }

// Taps of every output phase (see the header comment).  Returns false if unsupported.
bool build_taps(const tfc_conv_desc& d, Params* p) {
  const int su = d.su, sd = d.sd;
  int count = 0;
  p->su = su;
  for (int pi = 0; pi < su; ++pi) {
    for (int pj = 0; pj < su; ++pj) {
      p->phase_start[pi * su + pj] = count;
      for (int a = 0; a < d.kh; ++a) {
        for (int b = 0; b < d.kw; ++b) {
          Tap t{};
          if (su > 1) {  // sd == 1: tap is live for this phase iff it lands on the lattice
            const int ya = pi + a - d.pad_top, xb = pj + b - d.pad_left;
            if (((ya % su) + su) % su != 0 || ((xb % su) + su) % su != 0) continue;
            t.dh = (int8_t)floor_div(ya, su);
            t.dw = (int8_t)floor_div(xb, su);
            t.ph = t.pw = 0;
          } else {       // stride-sd correlation: split the offset into (box shift, parity)
            const int ya = a - d.pad_top, xb = b - d.pad_left;
            const int qh = floor_div(ya, sd), qw = floor_div(xb, sd);
            t.dh = (int8_t)qh; t.dw = (int8_t)qw;
            t.ph = (int8_t)(ya - qh * sd); t.pw = (int8_t)(xb - qw * sd);
          }
          t.widx = (int16_t)(a * d.kw + b);
          if (count >= kMaxTaps) return false;
          p->taps[count++] = t;
        }
      }
    }
  }
  p->phase_start[su * su] = count;
  return true;
}

void pick_tile(int hs, int ws, int* th, int* tw) {
  const int cand[][2] = {{8, 16}, {4, 32}, {16, 8}, {2, 64}, {32, 4}, {1, 128}, {64, 2}, {128, 1}};
  double best = -1;
  for (auto& c : cand) {
    const double cover = (double)((hs + c[0] - 1) / c[0] * c[0]) * ((ws + c[1] - 1) / c[1] * c[1]);
    const double util = (double)hs * ws / cover;
    if (util > best + 1e-9) { best = util; *th = c[0]; *tw = c[1]; }
  }
}

size_t smem_bytes(int npad, int cout, bool gdn, int stages = 4, bool pair = false) {
  const int b_bytes = (pair ? npad / 2 : npad) * kBlockK * 2;
  const int stage = kABytes + ((b_bytes + 1023) & ~1023);
  return 1024 + (size_t)stages * stage + (gdn ? (size_t)(cout / kBlockK) * kABytes : 0) +
         2 * 256 * sizeof(float) + sizeof(Barriers) + 64;
}

// stages of the operand ring: tfc_configure(TFC_CFG_CONV_STAGES), else as many as fit (<= 4)
int pick_stages(int npad, int cout, bool gdn, bool pair) {
  int want = tfc::g_conv_stages.load(std::memory_order_relaxed);
  // a CTA of a pair holds half of B per stage (28 KB instead of 40 KB at 192 filters): the same
  // shared memory carries five stages
  if (want < 2 || want > kStages) want = pair ? 5 : 4;
  while (want > 2 && smem_bytes(npad, cout, gdn, want, pair) > 227 * 1024) --want;
  return want;
}

}  // namespace

namespace tfc {

int make_tensor_map_plain(CUtensorMap* m, int dtype, void* base, int rank, const cuuint64_t* dims,
                          const cuuint64_t* strides_bytes, const cuuint32_t* box) {
  const CUtensorMapDataType t = dtype == TFC_F32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32
                                : dtype == TFC_U8 ? CU_TENSOR_MAP_DATA_TYPE_UINT8
                                                  : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  return make_map_any(m, t, CU_TENSOR_MAP_SWIZZLE_NONE, base, rank, dims, strides_bytes, box);
}

void conv_pick_tile(int hs, int ws, int* th, int* tw) { pick_tile(hs, ws, th, tw); }

int make_tensor_map_bf16(CUtensorMap* m, void* base, int rank, const cuuint64_t* dims,
                         const cuuint64_t* strides_bytes, const cuuint32_t* box) {
  return make_map(m, base, rank, dims, strides_bytes, box);
}

bool conv_umma_supported(const tfc_conv_desc& d, const char** why) {
  const char* dummy;
  if (!why) why = &dummy;
  const bool gdn = d.act == TFC_ACT_GDN || d.act == TFC_ACT_IGDN;
  if (d.in_dtype != TFC_BF16) { *why = "input must be bfloat16"; return false; }
  if (d.cin % kBlockK) { *why = "cin must be a multiple of 64"; return false; }
  if (d.cout > 256) { *why = "cout must be <= 256"; return false; }
  if (gdn && d.cout % kBlockK) { *why = "GDN needs cout to be a multiple of 64"; return false; }
  if (d.sd > 1 && d.su > 1) { *why = "strides_down and strides_up together"; return false; }
  if (d.su > 4 || d.sd > 8) { *why = "stride too large"; return false; }
  if (d.h % d.sd || d.w % d.sd) { *why = "input size must be a multiple of strides_down"; return false; }
  if (d.ho % d.su || d.wo % d.su) { *why = "output size must be a multiple of strides_up"; return false; }
  if (d.kh * d.kw > kMaxTaps || d.kh > 64 || d.kw > 64) { *why = "kernel too large"; return false; }
  const int npad = (d.cout + 15) / 16 * 16;
  if (smem_bytes(npad, d.cout, gdn, 2) > 227 * 1024) { *why = "shared memory"; return false; }
  if ((int64_t)d.n * d.ho * d.wo / 128 > (1 << 30)) { *why = "too many tiles"; return false; }
  return true;
}

size_t conv_umma_workspace(const tfc_conv_desc& d) {
  const int npad = (d.cout + 15) / 16 * 16;
  size_t w = (size_t)d.kh * d.kw * npad * d.cin * 2;
  w = (w + 255) & ~size_t(255);
  return w + (size_t)npad * npad * 2 + 256;
}

int conv_umma(const tfc_conv_desc& d, const void* in, const float* w, const float* bias,
              const float* beta, const float* gamma, void* out, void* ws, size_t ws_bytes,
              cudaStream_t st) {
  const bool gdn = d.act == TFC_ACT_GDN || d.act == TFC_ACT_IGDN;
  if (ws == nullptr || ws_bytes < conv_umma_workspace(d))
    return fail(TFC_ERR_INVALID_ARGUMENT, "conv workspace too small: need %zu bytes",
                conv_umma_workspace(d));
  if ((reinterpret_cast<uintptr_t>(in) & 15) || (reinterpret_cast<uintptr_t>(ws) & 255) ||
      (reinterpret_cast<uintptr_t>(out) & 15))
    return fail(TFC_ERR_INVALID_ARGUMENT, "conv buffers must be 16-byte (workspace 256-byte) aligned");
  Params p{};
  p.n = d.n; p.ho = d.ho; p.wo = d.wo; p.cout = d.cout; p.cin = d.cin;
  p.npad = (d.cout + 15) / 16 * 16;
  p.act = d.act; p.out_f32 = d.out_dtype == TFC_F32;
  p.out_wide = epi::wide_stores_ok(out, d.cout, p.out_f32 != 0);
  p.bias = bias; p.beta = beta; p.out = out;
  if (!build_taps(d, &p)) return fail(TFC_ERR_UNIMPLEMENTED, "too many taps");
  p.hs = d.ho / d.su; p.ws = d.wo / d.su;
  pick_tile(p.hs, p.ws, &p.th, &p.tw);
  p.tiles_h = (p.hs + p.th - 1) / p.th;
  p.tiles_w = (p.ws + p.tw - 1) / p.tw;
  p.num_tiles = d.su * d.su * d.n * p.tiles_h * p.tiles_w;
  // cluster size: groups of cs adjacent tiles must not straddle an output phase, and every CTA
  // loads npad/cs weight rows (whole 8-row swizzle atoms)
  p.cs = 1;
  {
    const int per_phase = d.n * p.tiles_h * p.tiles_w;
    const int want = conv_cluster();
    // automatic = off: the main loop is bound by the tensor pipe and the per-SM shared-memory
    // fill, neither of which multicast relieves (measured: A1 2.50 ms paired vs 2.56 ms)
    if ((want == 2 || want == 4) && per_phase % want == 0 && p.npad % (8 * want) == 0 &&
        p.num_tiles >= 2 * want)
      p.cs = want;
    // CTA pairs (tfc_configure(TFC_CFG_CONV_PAIR)): one M = 256 tcgen05.mma per two adjacent tiles,
    // each CTA holding half of B.  Needs an even number of tiles per output phase and N a multiple
    // of 32 (each half a multiple of 16 rows = whole swizzle atoms).
    // Automatic = on from 64 tiles (measured at batch 64, 768x512, profiles/r2d_*: A1 2.56 -> 2.40
    // ms, A2 0.644 -> 0.604, HS2 0.077 -> 0.072, S2 3.01 -> 2.94, S1 0.795 -> 0.789; ncu tensor
    // pipe of A1 56.4 -> 58.9 %, L2 throughput 59 -> 48 %).
    const int pair = tfc::g_conv_pair.load(std::memory_order_relaxed);
    const bool pair_wanted = pair == 2 || (pair == 0 && p.num_tiles >= 64 && want <= 1);
    if (pair_wanted && per_phase % 2 == 0 && p.npad % 32 == 0 && p.num_tiles >= 4 &&
        (!gdn || d.cout == 192)) {     // the pair kernel is built for bias-only and 192-filter GDN
      p.cs = 2;
      p.pair = 1;
    }
  }

  // weights (+ gamma^T) -> bf16 K-major
  __nv_bfloat16* wb = static_cast<__nv_bfloat16*>(ws);
  const int taps = d.kh * d.kw;
  size_t wbytes = ((size_t)taps * p.npad * d.cin * 2 + 255) & ~size_t(255);
  __nv_bfloat16* gb = reinterpret_cast<__nv_bfloat16*>(static_cast<uint8_t*>(ws) + wbytes);
  if (!(d.flags & TFC_CONV_WEIGHTS_STAGED)) {
    repack_weights_kernel<<<std::min<int64_t>(((int64_t)taps * p.npad * d.cin + 255) / 256, 2048),
                            256, 0, st>>>(w, taps, d.cin, d.cout, p.npad, wb);
    TFC_LAUNCHED("repack_weights_kernel");
    if (gdn) {
      // gamma[k][c] is "[1 tap][cin=k][cout=c]" -> [c][k]
      repack_weights_kernel<<<(d.cout * d.cout + 255) / 256, 256, 0, st>>>(gamma, 1, d.cout,
                                                                           d.cout, p.npad, gb);
      TFC_LAUNCHED("repack_weights_kernel");
    }
  }

  CUtensorMap map_a, map_w, map_g;
  const cuuint64_t C = d.cin, s = d.sd;
  {
    cuuint64_t dims[5] = {s * C, (cuuint64_t)d.w / s, s, (cuuint64_t)d.h / s, (cuuint64_t)d.n};
    cuuint64_t str[4] = {s * C * 2, (cuuint64_t)d.w * C * 2, s * d.w * C * 2,
                         (cuuint64_t)d.h * d.w * C * 2};
    cuuint32_t box[5] = {kBlockK, (cuuint32_t)p.tw, 1, (cuuint32_t)p.th, 1};
    int rc = make_map(&map_a, const_cast<void*>(in), 5, dims, str, box);
    if (rc) return rc;
  }
  {
    cuuint64_t dims[3] = {C, (cuuint64_t)p.npad, (cuuint64_t)taps};
    cuuint64_t str[2] = {C * 2, (cuuint64_t)p.npad * C * 2};
    cuuint32_t box[3] = {kBlockK, (cuuint32_t)(p.npad / p.cs), 1};
    int rc = make_map(&map_w, wb, 3, dims, str, box);
    if (rc) return rc;
  }
  if (gdn) {
    cuuint64_t dims[2] = {(cuuint64_t)d.cout, (cuuint64_t)p.npad};
    cuuint64_t str[1] = {(cuuint64_t)d.cout * 2};
    cuuint32_t box[2] = {kBlockK, (cuuint32_t)(p.npad / p.cs)};
    int rc = make_map(&map_g, gb, 2, dims, str, box);
    if (rc) return rc;
  } else {
    map_g = map_w;
  }
  p.stages = pick_stages(p.npad, d.cout, gdn, p.pair != 0);
  const size_t smem = smem_bytes(p.npad, d.cout, gdn, p.stages, p.pair != 0);
  p.chunk = tiles_per_cta();
  const int groups = (p.num_tiles + p.cs - 1) / p.cs;
  const int grid = p.cs * (p.chunk ? (groups + p.chunk - 1) / p.chunk
                                   : std::max(1, std::min(groups, persistent_grid() / p.cs)));
  OutMaps maps_out;
  p.out_tma = gdn && !p.out_f32 && d.cout % 64 == 0 && d.su <= 2 && !(debug_bits() & 16);
  for (int ph = 0; ph < 4; ++ph) {
    if (p.out_tma && ph < d.su * d.su) {
      // phase (pi, pj): pixels (su*i + pi, su*j + pj) of the output
      const int pi = ph / d.su, pj = ph % d.su;
      const cuuint64_t co2 = (cuuint64_t)d.cout * 2;
      cuuint64_t dims[4] = {(cuuint64_t)d.cout, (cuuint64_t)p.ws, (cuuint64_t)p.hs, (cuuint64_t)d.n};
      cuuint64_t str[3] = {d.su * co2, (cuuint64_t)d.su * d.wo * co2, (cuuint64_t)d.ho * d.wo * co2};
      cuuint32_t box[4] = {kBlockK, (cuuint32_t)p.tw, (cuuint32_t)p.th, 1};
      void* base = static_cast<uint8_t*>(out) + ((size_t)pi * d.wo + pj) * co2;
      int rc = make_map(&maps_out.m[ph], base, 4, dims, str, box);
      if (rc) return rc;
    } else {
      maps_out.m[ph] = map_w;
    }
  }
  void (*kern)(const CUtensorMap, const CUtensorMap, const CUtensorMap, const OutMaps,
               const Params) = nullptr;
  if (!gdn) {
    kern = p.pair ? conv_umma_kernel<0, 0, true> : conv_umma_kernel<0, 0, false>;
  } else {
    const bool inv = d.act == TFC_ACT_IGDN;
    switch (d.cout / 32) {
      case 2: kern = inv ? conv_umma_kernel<2, 2, false> : conv_umma_kernel<1, 2, false>; break;
      case 4: kern = inv ? conv_umma_kernel<2, 4, false> : conv_umma_kernel<1, 4, false>; break;
      case 6:
        if (p.pair) kern = inv ? conv_umma_kernel<2, 6, true> : conv_umma_kernel<1, 6, true>;
        else kern = inv ? conv_umma_kernel<2, 6, false> : conv_umma_kernel<1, 6, false>;
        break;
      case 8: kern = inv ? conv_umma_kernel<2, 8, false> : conv_umma_kernel<1, 8, false>; break;
      default: return fail(TFC_ERR_UNIMPLEMENTED, "GDN needs cout in {64, 128, 192, 256}");
    }
  }
  TFC_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  if (p.cs > 1) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(kThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = p.cs;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    TFC_CUDA(cudaLaunchKernelEx(&cfg, kern, map_a, map_w, map_g, maps_out, p));
  } else {
    kern<<<grid, kThreads, smem, st>>>(map_a, map_w, map_g, maps_out, p);
  }
  TFC_LAUNCHED("conv_umma_kernel");
  return TFC_OK;
}

}  // namespace tfc
