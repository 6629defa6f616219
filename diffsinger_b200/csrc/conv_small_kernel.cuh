// k-tap dilated Conv1d along time for NARROW channel counts (C_in, C_out in {16, 32}) on tcgen05 -- the two widest-in-time
// stages of the NSF-HiFiGAN generator (modules/nsf_hifigan/models.py:35-67, 285-296: 32 and 16 channels at 256x / 512x the
// mel rate), where padding channels to the 64-wide K blocks of ds_gemm_kernel wastes 2-4x of both HBM traffic and MMA work.
//
// Per tile of 128 output samples ONE activation halo tile (128 + (k-1) * dil rows) is loaded; every tap reads it through
// a shifted shared-memory descriptor.  That needs a layout in which "row + s" is a plain address offset: K-major WITHOUT
// swizzle, 8-channel planes [C/8][rows][16 B] (core matrices of 8 rows x 16 B packed densely: SBO = 128 B, so row r of a
// plane sits at r * 16 B; LBO = plane stride steps to the next 8 channels).  TMA writes each plane with a (8 ch, rows) box.
// One UMMA (M = 128, N = C_out, K = 16) per tap and 16-channel slice; accumulators double-buffered in TMEM.
// HBM-bound by design: per output sample C_in + C_out (+ residual, + activated copy) bf16 values.
//
// Warp roles: 0 = TMA producer, 1 = MMA issuer, 2..5 and 6..9 = two epilogue groups (even / odd tiles).  Persistent over tiles.
#pragma once
#include <cuda.h>

#include "aux_kernels.cuh"   // pack_bf16x8 / unpack_bf16x8
#include "layer_kernel.cuh"  // fence_proxy_async_smem
#include "ptx.cuh"

namespace dsb {

struct ConvSmallParams {
  CUtensorMap tmA;            // bf16 (C_in, T, B), box (8, rows, 1), no swizzle
  const __nv_bfloat16* w;     // [taps][C_in / 8][C_out][8] bf16: the shared-memory image of the weights
  const float* bias;          // [C_out]
  const __nv_bfloat16* res;   // optional residual input [B][T][C_out]
  void* out;                  // bf16 [B][T][C_out]; fp32 [B][T] (column 0 only) when act == 4
  __nv_bfloat16* out2;        // optional LeakyReLU(out, slope) copy
  int T, B, taps, dil, rows;  // rows = halo rows per tile (multiple of 8)
  int tiles_t, total_tiles;
  int act;                    // 0 none, 3 LeakyReLU(slope), 4 tanh -> fp32 column 0
  float slope;
};

constexpr int kConvSmallThreads = 320;  // 2 control warps + two epilogue groups of 4 warps (one per TMEM accumulator buffer)
constexpr int kConvSmallStages = 4;
constexpr int kConvSmallCtasPerSm = 2;  // ~40-60 KB of shared memory and 64 TMEM columns each
constexpr int kConvSmallMaxRows = 192;  // 128 + 10 * 5 = 178 -> 184
constexpr int kConvSmallMaxTaps = 11;

template <int CIN, int COUT>
struct ConvSmallCfg {
  static constexpr uint32_t kStageBytes = (CIN / 8) * kConvSmallMaxRows * 16;
  static constexpr uint32_t kWBytes = kConvSmallMaxTaps * CIN * COUT * 2;
  static constexpr uint32_t kSmemBytes = 1024 + kConvSmallStages * kStageBytes + kWBytes + 256;
};

// K-major, no swizzle: LBO = distance between the two 8-element K halves, SBO = distance between 8-row groups.
__device__ __forceinline__ uint64_t make_kmajor_noswz_desc(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr & 0x3FFFF) >> 4);
  d |= static_cast<uint64_t>(lbo_bytes >> 4) << 16;
  d |= static_cast<uint64_t>(sbo_bytes >> 4) << 32;
  d |= static_cast<uint64_t>(1) << 46;  // descriptor version (sm_100); layout type 0 = no swizzle
  return d;
}

template <int CIN, int COUT>
__global__ void __launch_bounds__(kConvSmallThreads, kConvSmallCtasPerSm) ds_conv_small_kernel(const __grid_constant__ ConvSmallParams p) {
  using Cfg = ConvSmallCfg<CIN, COUT>;
  constexpr int S = kConvSmallStages;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024u - (raw_addr & 1023u)) & 1023u);
  uint8_t* smem_a = smem;
  uint8_t* smem_w = smem + S * Cfg::kStageBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_w + Cfg::kWBytes);
  uint64_t* full = bars;
  uint64_t* empty = bars + S;
  uint64_t* tmem_full = bars + 2 * S;
  uint64_t* tmem_empty = bars + 2 * S + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * S + 4);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t plane_bytes = static_cast<uint32_t>(p.rows) * 16;

  if (warp == 0 && lane == 0) tma_prefetch_desc(&p.tmA);
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < S; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tmem_full[i], 1);
      mbar_init(&tmem_empty[i], 128);
    }
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc<64>(tmem_slot);
  {  // weights: global image -> shared memory (generic proxy), made visible to the tensor core's async proxy
    const int n16 = p.taps * CIN * COUT * 2 / 16;
    const uint4* src = reinterpret_cast<const uint4*>(p.w);
    uint4* dst = reinterpret_cast<uint4*>(smem_w);
    for (int i = threadIdx.x; i < n16; i += kConvSmallThreads) dst[i] = __ldg(src + i);
    fence_proxy_async_smem();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int halo = (p.taps / 2) * p.dil;

  if (warp == 0) {
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x) {
        const int tt = tile % p.tiles_t, b = tile / p.tiles_t;
        mbar_wait(&empty[stage], phase ^ 1);
        mbar_arrive_expect_tx(&full[stage], (CIN / 8) * plane_bytes);
        uint8_t* sa = smem_a + stage * Cfg::kStageBytes;
#pragma unroll
        for (int pl = 0; pl < CIN / 8; ++pl) tma_load_3d(&p.tmA, &full[stage], sa + pl * plane_bytes, pl * 8, tt * 128 - halo, b);
        if (++stage == S) {
          stage = 0;
          phase ^= 1;
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16_f32(COUT);
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x, ++it) {
        const int as = it & 1;
        mbar_wait(&tmem_empty[as], ((it >> 1) & 1) ^ 1);
        mbar_wait(&full[stage], phase);
        tc_fence_after();
        const uint32_t a_base = smem_u32(smem_a + stage * Cfg::kStageBytes);
        const uint32_t w_base = smem_u32(smem_w);
        const uint32_t tmem_d = tmem_base + as * 32;
        for (int tap = 0; tap < p.taps; ++tap) {
#pragma unroll
          for (int s = 0; s < CIN / 16; ++s) {
            const uint64_t da = make_kmajor_noswz_desc(a_base + 2 * s * plane_bytes + tap * p.dil * 16, plane_bytes, 128);
            const uint64_t db = make_kmajor_noswz_desc(w_base + ((tap * (CIN / 8) + 2 * s) * COUT) * 16, COUT * 16, 128);
            umma_bf16(tmem_d, da, db, idesc, (tap | s) != 0);
          }
        }
        umma_commit(&empty[stage]);
        umma_commit(&tmem_full[as]);
        if (++stage == S) {
          stage = 0;
          phase ^= 1;
        }
      }
    }
  } else {
    // epilogue group g = accumulator buffer g: the two groups alternate tiles, so one tile's residual loads / stores
    // (a global-memory round trip per tile) overlap the other's
    const int quad = warp & 3;
    const int grp = (warp - 2) >> 2;
    float bias[COUT];
#pragma unroll
    for (int j = 0; j < COUT; ++j) bias[j] = __ldg(p.bias + j);
    for (int it = grp; blockIdx.x + static_cast<long>(it) * gridDim.x < p.total_tiles; it += 2) {
      const int tile = blockIdx.x + it * gridDim.x;
      const int tt = tile % p.tiles_t, b = tile / p.tiles_t;
      const int as = it & 1;
      const int t = tt * 128 + quad * 32 + lane;
      const bool valid = t < p.T;
      const size_t row = static_cast<size_t>(b) * p.T + (valid ? t : 0);
      uint4 rq[COUT / 8];  // the residual row does not depend on the MMAs: in flight while this tile's accumulator completes
      if (p.res) {
        const uint4* rp = reinterpret_cast<const uint4*>(p.res + row * COUT);
#pragma unroll
        for (int i = 0; i < COUT / 8; ++i) rq[i] = __ldg(rp + i);
      }
      mbar_wait(&tmem_full[as], (it >> 1) & 1);
      tc_fence_after();
      uint32_t acc[32];
      tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(quad * 32) << 16) + as * 32, acc);
      float v[COUT];
      if (p.res) {
#pragma unroll
        for (int i = 0; i < COUT / 8; ++i) {
          float r8[8];
          unpack_bf16x8(rq[i], r8);
#pragma unroll
          for (int j = 0; j < 8; ++j) v[8 * i + j] = r8[j] + bias[8 * i + j];
        }
      } else {
#pragma unroll
        for (int j = 0; j < COUT; ++j) v[j] = bias[j];
      }
      tmem_ld_wait();
      tc_fence_before();
      mbar_arrive(&tmem_empty[as]);
#pragma unroll
      for (int j = 0; j < COUT; ++j) v[j] += __uint_as_float(acc[j]);
      if (p.act == 4) {
        if (valid) reinterpret_cast<float*>(p.out)[row] = tanhf(v[0]);
        continue;
      }
      if (p.act == 3) {
#pragma unroll
        for (int j = 0; j < COUT; ++j) v[j] = v[j] > 0.0f ? v[j] : v[j] * p.slope;
      }
      if (valid) {
        uint4* op = reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(p.out) + row * COUT);
#pragma unroll
        for (int i = 0; i < COUT / 8; ++i) {
          float o8[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) o8[j] = v[8 * i + j];
          op[i] = pack_bf16x8(o8);
        }
        if (p.out2) {
          uint4* op2 = reinterpret_cast<uint4*>(p.out2 + row * COUT);
#pragma unroll
          for (int i = 0; i < COUT / 8; ++i) {
            float o8[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) o8[j] = v[8 * i + j] > 0.0f ? v[8 * i + j] : v[8 * i + j] * p.slope;
            op2[i] = pack_bf16x8(o8);
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc<64>(tmem_base);
  }
}

}  // namespace dsb
