// CTA-pair variant of ds_gemm_kernel<256, EPI> (gemm_kernel.cuh): the same GEMM, epilogues and parameter block, issued as
// tcgen05.mma.cta_group::2 over a cluster of two CTAs.
//
// Why: with one CTA per 128 x 256 tile every UMMA (K = 16, 128 cycles) reads 4 KB of A and 8 KB of B from shared memory while
// TMA writes the same 12 KB for a later one -- 192 B/clk against the 128 B/clk the shared-memory port delivers, so the
// single-CTA kernel cannot pass ~2/3 of the tensor peak (measured 1.1 PFLOP/s, profiles/r01_launches_lynxnet.txt).  A pair
// owns a 256-frame x 256-column tile: each CTA stages its own 128 frames of A and HALF of the weight rows (128), the leader's
// M = 256 UMMA reads both halves, and the per-CTA traffic drops to 64 + 64 B/clk (the fused layer kernel's arrangement,
// layer_kernel.cuh, which reaches 88 % tensor-pipe activity with it).
//
// One pipeline stage = TWO 64-wide K blocks (8 UMMAs) under one full / empty barrier pair, three stages of 64 KB.
// Accumulators: 2 x 256 fp32 TMEM columns per CTA (its own 128 frames x 256 columns), drained by the unchanged
// epilogue_tile<256, EPI> of gemm_epilogue.cuh, whose bf16 outputs leave through warp-private staging tiles and TMA stores (p.tmOut;
// the per-lane STG.128 epilogue alone took 64 of the 241 us of the LYNXNet C -> 2I product, tools/gemm_bench.py).  p.tmBh is the weight map with a 128-row box; p.tiles_t counts 256-frame tiles.
#pragma once
#include "gemm_epilogue.cuh"
#include "layer_kernel.cuh"  // kPairBBytes

namespace dsb {

constexpr int kGpGroup = 2;   // K blocks per pipeline stage (1 x 6 stages measured the same on LYNXNet, 2.5 % slower in training)
constexpr int kGpStages = 6 / kGpGroup;
constexpr uint32_t kGpABytes = 128 * 128;                                // 128 frames x 64 bf16
constexpr uint32_t kGpStageBytes = kGpGroup * (kGpABytes + kPairBBytes);  // [A blk0][A blk1][B blk0][B blk1]
constexpr uint32_t kGpWarpStage = 2048;                                  // per epilogue warp: 32 frames x 32 channels bf16, SWIZZLE_64B
constexpr uint32_t kGpSmemBytes = 1024 + kGpStages * kGpStageBytes + 8 * kGpWarpStage + 256;

template <int EPI>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kGemmThreads, 1) ds_gemm_pair_kernel(const __grid_constant__ GemmParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024u - (raw_addr & 1023u)) & 1023u);  // SWIZZLE_128B needs 1024-B alignment
  uint8_t* warp_stage = smem + kGpStages * kGpStageBytes;  // 8 x kGpWarpStage, 2 KB aligned
  uint64_t* bars = reinterpret_cast<uint64_t*>(warp_stage + 8 * kGpWarpStage);
  uint64_t* full = bars;                  // leader: 2 producer arrivals + the bytes of both CTAs
  uint64_t* empty = full + kGpStages;     // per CTA: released by the leader's multicast commit
  uint64_t* tmem_full = empty + kGpStages;
  uint64_t* tmem_empty = tmem_full + 2;   // leader: 16 warp arrivals (8 per CTA)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);  // warp-uniform to the compiler: converged single-issuer roles (gemm_kernel.cuh)
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&p.tmA0);
    tma_prefetch_desc(&p.tmBh);
    if (p.kb_a1 > 0) tma_prefetch_desc(&p.tmA1);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < kGpStages; ++i) {
      mbar_init(&full[i], 2);
      mbar_init(&empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tmem_full[i], 1);
      mbar_init(&tmem_empty[i], 16);
    }
    fence_barrier_init();
  }
  cluster_sync_all();  // peer barriers are initialised before any remote arrive / multicast commit
  if (warp == 2) tmem_alloc_pair<512>(tmem_slot);
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  grid_dep_launch();
  grid_dep_wait();

  const int num_kb = p.n_taps * p.kb_a0 + p.kb_a1;
  const int pair_id = blockIdx.x >> 1;
  const int n_pairs = gridDim.x >> 1;

  if (warp == 0) {
    {
      // ------------------------------------------------------------ TMA producer (both CTAs)
      const bool lead = elect_one();
      int st = 0;
      uint32_t ph = 0;
      const int kb_taps = p.n_taps * p.kb_a0;
      for (int tile = pair_id; tile < p.total_tiles; tile += n_pairs) {
        const int nt = tile % p.n_tiles;
        const int tt = (tile / p.n_tiles) % p.tiles_t;
        const int b = tile / (p.n_tiles * p.tiles_t);
        const int t0 = tt * 256 + rank * 128;
        const int row0 = p.b_row0 + nt * 256 + rank * 128;
        for (int kb0 = 0; kb0 < num_kb; kb0 += kGpGroup) {
          const int n = num_kb - kb0 < kGpGroup ? num_kb - kb0 : kGpGroup;
          mbar_wait(&empty[st], ph ^ 1);
          if (lead) {
            if (leader) mbar_arrive_expect_tx(&full[st], 2u * n * (kGpABytes + kPairBBytes));
            else mbar_arrive_cluster(&full[st], 0);
          }
          uint8_t* s = smem + st * kGpStageBytes;
          for (int j = 0; j < n && lead; ++j) {
            const int kb = kb0 + j;
            if (kb < kb_taps) {
              const int tap = kb / p.kb_a0;
              const int cb = kb - tap * p.kb_a0;
              tma_load_3d_pair(&p.tmA0, &full[st], s + j * kGpABytes, p.a0_c0 + cb * 64, t0 + (tap - p.n_taps / 2) * p.dil, b);
            } else {
              tma_load_3d_pair(&p.tmA1, &full[st], s + j * kGpABytes, (kb - kb_taps) * 64, t0, b);
            }
            tma_load_2d_pair(&p.tmBh, &full[st], s + kGpGroup * kGpABytes + j * kPairBBytes, p.b_k0 + kb * 64, row0);
          }
          if (++st == kGpStages) { st = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (leader) {
      // ------------------------------------------------------------ MMA issuer (leader CTA only)
      const bool lead = elect_one();
      constexpr uint32_t idesc = make_idesc_bf16_f32_m(256, 256);
      int st = 0;
      uint32_t ph = 0;
      int it = 0;
      for (int tile = pair_id; tile < p.total_tiles; tile += n_pairs, ++it) {
        const uint32_t as = it & 1;
        mbar_wait(&tmem_empty[as], ((it >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + as * 256;
        for (int kb0 = 0; kb0 < num_kb; kb0 += kGpGroup) {
          const int n = num_kb - kb0 < kGpGroup ? num_kb - kb0 : kGpGroup;
          mbar_wait(&full[st], ph);
          tc_fence_after();
          const uint32_t s = smem_u32(smem + st * kGpStageBytes);
          for (int j = 0; j < n && lead; ++j) {
            const uint64_t da = make_kmajor_sw128_desc(s + j * kGpABytes);
            const uint64_t db = make_kmajor_sw128_desc(s + kGpGroup * kGpABytes + j * kPairBBytes);
#pragma unroll
            for (int k = 0; k < 4; ++k) umma_bf16_pair(tmem_d, da + 2 * k, db + 2 * k, idesc, ((kb0 + j) | k) != 0);
          }
          if (lead) {
            umma_commit_pair(&empty[st]);
            if (kb0 + n >= num_kb) umma_commit_pair(&tmem_full[as]);
          }
          if (++st == kGpStages) { st = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp >= 4) {
    // -------------------------------------------------------------- epilogue (both CTAs; 256 threads each)
    const int quad = warp & 3;  // TMEM lane quadrant this warp may read
    const int half = (warp - 4) >> 2;
    int it = 0;
    for (int tile = pair_id; tile < p.total_tiles; tile += n_pairs, ++it) {
      const int nt = tile % p.n_tiles;
      const int tt = (tile / p.n_tiles) % p.tiles_t;
      const int b = tile / (p.n_tiles * p.tiles_t);
      const uint32_t as = it & 1;
      mbar_wait(&tmem_full[as], (it >> 1) & 1);
      tc_fence_after();
      const int t = tt * 256 + rank * 128 + quad * 32 + lane;
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(quad * 32) << 16) + as * 256;
      EpiStage es;
      es.sw = smem_u32(warp_stage + (warp - 4) * kGpWarpStage);
      es.lane = lane;
      es.tw0 = tt * 256 + rank * 128 + quad * 32;
      es.on = p.has_out_map != 0;
      if (p.vec_bstride != -12345) epilogue_tile<256, EPI, false>(p, taddr, b, t, t < p.T, nt, half, es);  // (timing hook of tools/gemm_bench.py)
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {  // release the accumulator on the LEADER's barrier: one arrival per warp
        if (leader) mbar_arrive(&tmem_empty[as]);
        else mbar_arrive_cluster(&tmem_empty[as], 0);
      }
    }
  }
  if (warp >= 4) bulk_wait_all();  // staging tiles are read out before the CTA's shared memory goes away
  tc_fence_before();
  cluster_sync_all();  // the leader's MMAs read this CTA's shared memory and write its TMEM until the very end
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc_pair<512>(tmem_base);
  }
}

}  // namespace dsb
