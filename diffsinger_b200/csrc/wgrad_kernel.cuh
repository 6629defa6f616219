// Weight-gradient GEMM of the training path:
//
//   out[m][n] = sum_{b, t} G[b][t][g_c0 + m] * X[b][t + shift][x_c0 + n]          (fp32, one 128 x BN tile per CTA)
//
// e.g. d(dilated_conv.weight)[:, :, tap] = dz^T x~ shifted by (tap - 1) * dilation (wavenet.py:36-38 under autograd),
// d(conditioner_projection.weight) = dz^T cond, d(output_projection.weight) = [h | dskip]^T y.
//
// Both operands are time-major activations [B][T][channels] -- the reduction runs over FRAMES, so each operand tile
// arrives as [64 frames][64 channels] TMA boxes (128-byte rows, SWIZZLE_128B) and is consumed by tcgen05.mma as an
// MN-major operand (instruction-descriptor bits 15 / 16; canonical layout ((8,n),(8,k)):((1,LBO),(8,SBO)) in 16-byte
// units: SBO = 1024 B between 8-frame groups, LBO = 8192 B between 64-channel boxes).  The time shift of a conv tap
// and the utterance boundaries are TMA coordinates: frames outside [0, T) are zero filled, exactly the conv's padding,
// so there is no im2col, no per-utterance batched product and no reduction over utterances afterwards.
//
// One CTA = one output tile, K loop over all B * ceil(T / 64) frame blocks; tiles of one launch come from a job table
// (all layers, taps and weight tensors at once: up to four (G, X) operand pairs per launch, so that one backward pass
// is ~10 full waves instead of several launches with a partial last wave each), ordered so that tiles sharing operand
// columns run concurrently and share L2.
// Warp roles: 0 = TMA producer, 1 = MMA issuer, 2..5 = epilogue (TMEM -> fp32 global).
#pragma once
#include <cuda.h>

#include "ptx.cuh"

namespace dsb {

struct WgradJob {
  int g_c0, g_b0;    // first channel / first batch row of the G operand (M side)
  int x_c0, x_b0;    // same for the X operand (N side)
  int shift;         // X frame = G frame + shift
  int m_valid;       // rows of this tile that exist (<= 128)
  int n_valid;       // columns of this tile that exist (<= BN)
  int ld_out;        // fp32 elements between output rows
  long long out_off; // element offset of out[0][0] of this tile
  int pair;          // which (G, X) operand pair of the launch this tile reads
  int pad;
};

constexpr int kWgradMaxPairs = 4;

struct WgradParams {
  CUtensorMap tmG[kWgradMaxPairs], tmX[kWgradMaxPairs];  // bf16 (channels, T, batch rows), box (64, 64, 1), SWIZZLE_128B
  const WgradJob* jobs;
  float* out;
  int T, B;
};

constexpr int kWgradThreads = 192;
constexpr uint32_t kWgradBox = 64 * 128;  // one [64 frames][64 channels] box

template <int BN>
struct WgradCfg {
  static constexpr uint32_t kABytes = 2 * kWgradBox;
  static constexpr uint32_t kBBytes = (BN / 64) * kWgradBox;
  static constexpr int kStages = BN == 256 ? 4 : (BN == 128 ? 6 : 8);
  static constexpr uint32_t kSmemBytes = 1024 + kStages * (kABytes + kBBytes) + 256;
};

__device__ __forceinline__ uint64_t make_mnmajor_sw128_desc(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr & 0x3FFFF) >> 4);  // [0,14)  start address >> 4
  d |= static_cast<uint64_t>(kWgradBox >> 4) << 16;        // [16,30) leading byte offset: next 64-channel box
  d |= static_cast<uint64_t>(1024 >> 4) << 32;             // [32,46) stride byte offset: next 8-frame group
  d |= static_cast<uint64_t>(1) << 46;                     // [46,48) descriptor version (sm_100)
  d |= static_cast<uint64_t>(2) << 61;                     // [61,64) SWIZZLE_128B
  return d;
}

template <int BN>
__global__ void __launch_bounds__(kWgradThreads, 1) ds_wgrad_kernel(const __grid_constant__ WgradParams p) {
  using Cfg = WgradCfg<BN>;
  constexpr int S = Cfg::kStages;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024u - (raw_addr & 1023u)) & 1023u);
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + S * Cfg::kABytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + S * (Cfg::kABytes + Cfg::kBBytes));
  uint64_t* full = bars;
  uint64_t* empty = bars + S;
  uint64_t* tmem_full = bars + 2 * S;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * S + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const WgradJob job = p.jobs[blockIdx.x];
  const CUtensorMap* tmG = &p.tmG[job.pair];
  const CUtensorMap* tmX = &p.tmX[job.pair];
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(tmG);
    tma_prefetch_desc(tmX);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < S; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    mbar_init(tmem_full, 1);
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc<BN>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int tiles_t = (p.T + 63) / 64;
  const int num_kb = p.B * tiles_t;

  if (warp == 0) {
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int b = 0; b < p.B; ++b) {
        for (int tt = 0; tt < tiles_t; ++tt) {
          const int t0 = tt * 64;
          mbar_wait(&empty[stage], phase ^ 1);
          mbar_arrive_expect_tx(&full[stage], Cfg::kABytes + Cfg::kBBytes);
          uint8_t* sa = smem_a + stage * Cfg::kABytes;
          uint8_t* sb = smem_b + stage * Cfg::kBBytes;
          tma_load_3d(tmG, &full[stage], sa, job.g_c0, t0, job.g_b0 + b);
          tma_load_3d(tmG, &full[stage], sa + kWgradBox, job.g_c0 + 64, t0, job.g_b0 + b);
#pragma unroll
          for (int j = 0; j < BN / 64; ++j)
            tma_load_3d(tmX, &full[stage], sb + j * kWgradBox, job.x_c0 + j * 64, t0 + job.shift, job.x_b0 + b);
          if (++stage == S) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16_f32(BN) | (1u << 15) | (1u << 16);  // A and B MN-major
      int stage = 0;
      uint32_t phase = 0;
      for (int kb = 0; kb < num_kb; ++kb) {
        mbar_wait(&full[stage], phase);
        tc_fence_after();
        const uint64_t da = make_mnmajor_sw128_desc(smem_u32(smem_a + stage * Cfg::kABytes));
        const uint64_t db = make_mnmajor_sw128_desc(smem_u32(smem_b + stage * Cfg::kBBytes));
#pragma unroll
        for (int k = 0; k < 4; ++k)  // 4 x (K = 16 frames): +2048 B on the start address
          umma_bf16(tmem_base, da + k * (2048 >> 4), db + k * (2048 >> 4), idesc, (kb | k) != 0);
        umma_commit(&empty[stage]);
        if (++stage == S) {
          stage = 0;
          phase ^= 1;
        }
      }
      umma_commit(tmem_full);
    }
  } else {
    const int quad = warp & 3;
    mbar_wait(tmem_full, 0);
    tc_fence_after();
    const int m = quad * 32 + lane;
    const uint32_t taddr = tmem_base + (static_cast<uint32_t>(quad * 32) << 16);
    float* orow = p.out + job.out_off + static_cast<long long>(m) * job.ld_out;
#pragma unroll 1
    for (int c = 0; c < BN / 32; ++c) {
      uint32_t v[32];
      tmem_ld_32x32(taddr + c * 32, v);
      tmem_ld_wait();
      if (m < job.m_valid) {
        if (c * 32 + 32 <= job.n_valid) {
#pragma unroll
          for (int i = 0; i < 8; ++i)
            reinterpret_cast<float4*>(orow + c * 32)[i] = make_float4(__uint_as_float(v[4 * i]), __uint_as_float(v[4 * i + 1]),
                                                                      __uint_as_float(v[4 * i + 2]), __uint_as_float(v[4 * i + 3]));
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (c * 32 + j < job.n_valid) orow[c * 32 + j] = __uint_as_float(v[j]);
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc<BN>(tmem_base);
  }
}

}  // namespace dsb
