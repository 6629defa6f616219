// Depthwise Conv1d along time as per-channel Toeplitz products on the warp-level tensor-core path (mma.sync m16n8k16, bf16 x bf16
// -> fp32).  lynxnet.py:59 (k = 31), convnext.py:22 (k = 7); the fp32-FMA kernel in aux_kernels.cuh spends 31 FFMA2 + ~50 other
// issue slots per output pair and is issue-bound at 1.6 TB/s (profiles/r01_prim_bench.txt); here a channel's 128 outputs cost
// 2 ldmatrix + 3 mma.
//
// For one channel c and a tile of 128 output frames starting at t0:   y[t0 + 16n + i] = sum_k W[i][k] * X[k][n]
//   W[i][k] = w_c[k - i] (0 <= k-i < K)  16 x (16 NK) Toeplitz operand A: constant over time, lives in registers (4 NK per channel);
//   X[k][n] = x[t0 - pad + 16n + k]      (16 NK) x 8 operand B: overlapping windows of the channel's time series.  With the tile
//                                        stored channel-major in shared memory a window starts every 32 bytes, so the B fragments
//                                        are plain ldmatrix loads of 16-byte chunks; even and odd chunks live in separate halves
//                                        of the channel row, which makes every 8-row ldmatrix phase conflict-free.
// Activations arrive time-major ([T][CH]), so every tile is transposed once on its way in: ONE TMA box (32 channels x kTW frames,
// SWIZZLE_64B; frames outside [0, T) are zero-filled = the conv's zero padding) per tile into a natural-layout staging tile under an
// mbarrier, then LDS.128 of two consecutive frames, 8 PRMTs and 8 STS.32 per thread unit into the channel-major tile.  Outputs go back through a time-major tile to 16-byte coalesced stores.
// A CTA (8 warps x 4 channels) walks `chunks` consecutive 128-frame tiles of one utterance; the TMA loads run two tiles ahead
// (double-buffered staging tile).  Weights are rounded to bf16 (as in every GEMM of the same model); accumulation is fp32.
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cstdint>

#include "ptx.cuh"

namespace dsb {

__device__ __forceinline__ void ldmatrix_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void ldmatrix_x2(uint32_t addr, uint32_t& r0, uint32_t& r1) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x2.shared.b16 {%0, %1}, [%2];" : "=r"(r0), "=r"(r1) : "r"(addr));
}
__device__ __forceinline__ void mma_bf16_16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t prmt(uint32_t a, uint32_t b, uint32_t sel) {
  uint32_t r;
  asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(sel));
  return r;
}
template <int kAct>
__device__ __forceinline__ float dw_act(float v, float slope) {
  if (kAct == 1) return fmaxf(v, 0.f);
  if (kAct == 2) return v > 0.f ? v : slope * v;
  if (kAct == 3) return v / (1.f + __expf(-v));
  return v;
}

template <int K>
struct DwMma {
  static constexpr int kNK = (16 + K - 1 + 15) / 16;   // k16 steps covering the 16 + K - 1 inputs of 16 consecutive outputs
  static constexpr int kTW = 16 * 7 + 16 * kNK;        // frames the B windows of one 128-frame tile touch (160 for K = 31)
  static constexpr int kStageRow = 64;                 // bytes per staged frame: 32 channels, 16-byte chunks XOR-swizzled by (frame / 2) % 4 (TMA SWIZZLE_64B)
  static constexpr int kStageBytes = kTW * kStageRow;
  static constexpr int kHalf = 12;                     // 16-byte slot of the first odd chunk: 12 * 4 words = 16 (mod 32 banks)
  static constexpr int kXtRow = 24 * 16;               // bytes per channel row: even chunks in slots 0.., odd chunks in slots 12..
  static constexpr int kXtBytes = 32 * kXtRow;
  static constexpr int kOutBytes = 128 * 64;           // [frame][32 channels], 8-byte groups XOR-swizzled per frame
  static constexpr int kSmemBytes = 2 * kStageBytes + kXtBytes + kOutBytes;  // staging tile double buffered
  static_assert(kTW / 8 <= 2 * kHalf && (kTW / 8 + 1) / 2 <= kHalf, "channel row too short");
};

// grid = (CH / 32, ceil(ceil(T / 128) / chunks), B), block = 256, dynamic smem = DwMma<K>::kSmemBytes
template <int K, int kAct>
__global__ void __launch_bounds__(256) dwconv_mma_kernel(const __grid_constant__ CUtensorMap tmX, __nv_bfloat16* __restrict__ y,
                                                         const float* __restrict__ w, const float* __restrict__ bias,
                                                         const float* __restrict__ slope, int T, int CH, int pad, int chunks) {
  using S = DwMma<K>;
  extern __shared__ __align__(1024) uint8_t dwm_smem[];
  __shared__ uint64_t dwm_bars[2];
  grid_dep_launch();
  const uint32_t stage = smem_u32(dwm_smem), xt = stage + 2 * S::kStageBytes, ot = xt + S::kXtBytes;
  const int tid = threadIdx.x, lane = tid & 31, wq = tid >> 5, g = lane >> 2, q = lane & 3;
  const int c0 = blockIdx.x * 32, b = blockIdx.z;
  const int n_tiles = (T + 127) / 128;
  const int tile0 = blockIdx.y * chunks, tile1 = min(tile0 + chunks, n_tiles);
  __nv_bfloat16* yb = y + (static_cast<size_t>(b) * T) * CH + c0;

  // Toeplitz A fragments of this warp's 4 channels: aw[s] = {W[g][16s+2q..], W[g+8][16s+2q..], W[g][16s+2q+8..], W[g+8][16s+2q+8..]}
  uint32_t aw[4][S::kNK][4];
  float bv[4], sv[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int c = c0 + wq * 4 + j;
    const float* wc = w + static_cast<size_t>(c) * K;
#pragma unroll
    for (int s = 0; s < S::kNK; ++s)
#pragma unroll
      for (int h = 0; h < 4; ++h) {
        const int k = 16 * s + 2 * q + 8 * (h >> 1) - g - 8 * (h & 1);  // tap index of the lower element
        const float lo = (k >= 0 && k < K) ? __ldg(wc + k) : 0.f;
        const float hi = (k + 1 >= 0 && k + 1 < K) ? __ldg(wc + k + 1) : 0.f;
        aw[j][s][h] = pack_bf16(lo, hi);
      }
    bv[j] = bias ? __ldg(bias + c) : 0.f;
    sv[j] = (kAct == 2) ? __ldg(slope + c) : 0.f;
  }
  grid_dep_wait();

  // tile -> staging buffer (tile - tile0) % 2: one TMA box per tile, completion on the buffer's mbarrier
  if (tid == 0) {
    tma_prefetch_desc(&tmX);
    mbar_init(&dwm_bars[0], 1);
    mbar_init(&dwm_bars[1], 1);
    fence_barrier_init();
  }
  __syncthreads();
  auto issue_loads = [&](int tile) {  // thread 0 only
    if (tile < tile1) {
      const int buf = (tile - tile0) & 1;
      mbar_arrive_expect_tx(&dwm_bars[buf], S::kStageBytes);
      tma_load_3d(&tmX, &dwm_bars[buf], dwm_smem + buf * S::kStageBytes, c0, tile * 128 - pad, b);
    }
  };
  if (tid == 0) {
    issue_loads(tile0);
    issue_loads(tile0 + 1);
  }

  // ldmatrix lane address inside a channel row: matrix mi = lane / 8 holds chunks 2r + mi (+ 2 per k16 step), r = lane % 8
  const int mi = lane >> 3, r = lane & 7;
  const uint32_t b_lane = xt + (wq * 4) * S::kXtRow + (r + (mi >> 1) + (mi & 1) * S::kHalf) * 16;

  for (int tile = tile0; tile < tile1; ++tile) {
    mbar_wait(&dwm_bars[(tile - tile0) & 1], ((tile - tile0) >> 1) & 1);  // staged tile landed
    const uint8_t* sg = dwm_smem + ((tile - tile0) & 1) * S::kStageBytes;
    // transpose: unit = (frame pair tp, channel octet o) -> 8 words (c @ 2tp, c @ 2tp + 1) of the channel-major tile
    for (int u = tid; u < (S::kTW / 2) * 4; u += 256) {
      const int o = u / (S::kTW / 2), tp = u - o * (S::kTW / 2);
      const int oc = (o ^ (tp & 3)) << 4;  // SWIZZLE_64B: chunk ^= (frame / 2) % 4, the same for frames 2tp and 2tp + 1
      const uint4 r0 = *reinterpret_cast<const uint4*>(sg + (2 * tp) * S::kStageRow + oc);
      const uint4 r1 = *reinterpret_cast<const uint4*>(sg + (2 * tp + 1) * S::kStageRow + oc);
      const uint32_t a[4] = {r0.x, r0.y, r0.z, r0.w}, c[4] = {r1.x, r1.y, r1.z, r1.w};
      const int ch = tp >> 2;  // 16-byte chunk (8 frames) -> slot (ch / 2) + (ch % 2) * kHalf
      const uint32_t dst = xt + (o * 8) * S::kXtRow + ((ch >> 1) + (ch & 1) * S::kHalf) * 16 + (tp & 3) * 4;
#pragma unroll
      for (int p = 0; p < 4; ++p) {
        asm volatile("st.shared.b32 [%0], %1;" ::"r"(dst + (2 * p) * S::kXtRow), "r"(prmt(a[p], c[p], 0x5410)) : "memory");
        asm volatile("st.shared.b32 [%0], %1;" ::"r"(dst + (2 * p + 1) * S::kXtRow), "r"(prmt(a[p], c[p], 0x7632)) : "memory");
      }
    }
    __syncthreads();  // channel-major tile complete, this staging buffer free; the previous tile's output tile has been read out
    if (tid == 0) {
      fence_proxy_async_smem();  // generic-proxy reads of the staging buffer before its async-proxy refill
      issue_loads(tile + 2);
    }

    // 4 channels x (ldmatrix.x4 [+ .x2] + kNK mma); D[g (+8)][2q (+1)] = frame 16 (2q (+1)) + g (+8)
    uint32_t ow[4][2];  // [frame slot v][channel pair] packed bf16
#pragma unroll
    for (int jp = 0; jp < 2; ++jp) {
      float d[2][4];
#pragma unroll
      for (int jj = 0; jj < 2; ++jj) {
        const int j = 2 * jp + jj;
#pragma unroll
        for (int v = 0; v < 4; ++v) d[jj][v] = bv[j];
        uint32_t bx[2 * ((S::kNK + 1) / 2) * 2];
#pragma unroll
        for (int s2 = 0; s2 < S::kNK / 2; ++s2)  // two k16 steps per ldmatrix.x4 (chunks advance by 4 = 2 slots per parity)
          ldmatrix_x4(b_lane + j * S::kXtRow + s2 * 32, bx[4 * s2], bx[4 * s2 + 1], bx[4 * s2 + 2], bx[4 * s2 + 3]);
        if (S::kNK & 1)  // last k16 step alone: matrices 0 / 1 of lanes 0..15
          ldmatrix_x2(b_lane + j * S::kXtRow + (S::kNK / 2) * 32, bx[2 * (S::kNK - 1)], bx[2 * (S::kNK - 1) + 1]);
#pragma unroll
        for (int s = 0; s < S::kNK; ++s) mma_bf16_16816(d[jj], aw[j][s], bx[2 * s], bx[2 * s + 1]);
#pragma unroll
        for (int v = 0; v < 4; ++v) d[jj][v] = dw_act<kAct>(d[jj][v], sv[j]);
      }
#pragma unroll
      for (int v = 0; v < 4; ++v) ow[v][jp] = pack_bf16(d[0][v], d[1][v]);
    }
    // 8 B (this warp's 4 channels) per frame into the time-major output tile; 8-byte group index XOR (q | (g / 2 % 2) << 2)
#pragma unroll
    for (int v = 0; v < 4; ++v) {
      const int t = 32 * q + 16 * (v & 1) + g + 8 * (v >> 1);
      const int sw = ((t >> 5) & 3) | (((t >> 1) & 1) << 2);
      asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(ot + t * 64 + ((wq ^ sw) << 3)), "r"(ow[v][0]), "r"(ow[v][1]) : "memory");
    }
    __syncthreads();  // output tile complete
    const int tb = tile * 128;
#pragma unroll
    for (int it = 0; it < 2; ++it) {
      const int idx = tid + 256 * it, t = idx >> 2, o = idx & 3;
      if (tb + t < T) {
        const int sw = ((t >> 5) & 3) | (((t >> 1) & 1) << 2);
        uint4 v = *reinterpret_cast<const uint4*>(dwm_smem + 2 * S::kStageBytes + S::kXtBytes + t * 64 + ((o ^ (sw >> 1)) << 4));
        if (sw & 1) v = make_uint4(v.z, v.w, v.x, v.y);  // the two 8-byte groups of the chunk sit swapped
        *reinterpret_cast<uint4*>(yb + static_cast<size_t>(tb + t) * CH + o * 8) = v;
      }
    }
  }
}

}  // namespace dsb
