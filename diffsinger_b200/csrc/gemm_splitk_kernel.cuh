// Split-K variant of ds_gemm_kernel<BN, EPI> (BN = 64 / 128) for the latency-bound small problems (BASELINE configs #1 / #4:
// batch 1, a few hundred frames).  There a GEMM has 4-32 output tiles, so most of the 148 SMs idle while each busy CTA
// streams the whole K extent of its tile through one TMA pipeline (gate GEMM at T = 500, C = 512: 32 CTAs x 896 KB, 14.8 us;
// profiles/r01_launches_small.txt).  Here a cluster of kSplit = 4 CTAs owns one 128-frame x BN tile and each CTA multiplies
// a quarter of the K blocks into its own TMEM accumulator.  The three peers then dump their fp32 partial tiles into their own
// shared memory ([column / 4][frame][4]: conflict-free 16-byte stores), one cluster barrier publishes them, and the leader
// adds them to its accumulator over distributed shared memory (ld.shared::cluster.v4), writes the sums back to TMEM
// (tcgen05.st) and runs the unchanged fused epilogue (gemm_epilogue.cuh).  One tile per cluster, no persistence: the kernel
// exists for problems with fewer tiles than SMs.
// Measured at B1 x T500, C = 512 (ncu, cold cache): deferred skip GEMM (160 K blocks) ~50 -> 16 us; gate GEMM (28 K blocks)
// 14.8 -> 15.5 us and residual GEMM (8 K blocks) 8.7 -> 10.7 us -- a 2-K-block GEMM of the same tile count already takes 7 us,
// so the launcher takes this kernel only for K >= 64 blocks.
#pragma once
#include "gemm_epilogue.cuh"

namespace dsb {

constexpr int kSplitK = 4;

template <int BN>
struct SplitKCfg {
  static constexpr int kStages = BN == 128 ? 4 : 6;
  static constexpr uint32_t kABytes = 128 * 128;
  static constexpr uint32_t kBBytes = BN * 128;
  static constexpr uint32_t kPartBytes = 128 * BN * 4;  // this CTA's fp32 partial tile
  static constexpr uint32_t kSmemBytes = 1024 + kStages * (kABytes + kBBytes) + kPartBytes + 256;
};

__device__ __forceinline__ void tmem_st_32x32(uint32_t taddr, const uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
      ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]),
        "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]), "r"(v[16]), "r"(v[17]), "r"(v[18]), "r"(v[19]),
        "r"(v[20]), "r"(v[21]), "r"(v[22]), "r"(v[23]), "r"(v[24]), "r"(v[25]), "r"(v[26]), "r"(v[27]), "r"(v[28]), "r"(v[29]),
        "r"(v[30]), "r"(v[31])
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ uint32_t dsmem_addr(uint32_t local_addr, uint32_t cta) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_addr), "r"(cta));
  return r;
}
__device__ __forceinline__ float4 ld_dsmem_v4(uint32_t addr) {
  float4 v;
  asm volatile("ld.shared::cluster.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr) : "memory");
  return v;
}

// grid = kSplitK * total_tiles (cluster = the kSplitK CTAs of one tile); requires n_taps * kb_a0 + kb_a1 >= kSplitK.
template <int BN, int EPI>
__global__ void __cluster_dims__(kSplitK, 1, 1) __launch_bounds__(kGemmThreads, 1) ds_gemm_splitk_kernel(const __grid_constant__ GemmParams p) {
  using Cfg = SplitKCfg<BN>;
  constexpr int S = Cfg::kStages;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024u - (raw_addr & 1023u)) & 1023u);  // SWIZZLE_128B needs 1024-B alignment
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + S * Cfg::kABytes;
  uint8_t* part = smem + S * (Cfg::kABytes + Cfg::kBBytes);
  uint64_t* bars = reinterpret_cast<uint64_t*>(part + Cfg::kPartBytes);
  uint64_t* full = bars;
  uint64_t* empty = bars + S;
  uint64_t* tmem_full = bars + 2 * S;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * S + 1);

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);  // warp-uniform to the compiler: converged single-issuer roles (gemm_kernel.cuh)
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&p.tmA0);
    tma_prefetch_desc(&p.tmB);
    if (p.kb_a1 > 0) tma_prefetch_desc(&p.tmA1);
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
  grid_dep_launch();
  grid_dep_wait();

  const int tile = blockIdx.x / kSplitK;
  const int nt = tile % p.n_tiles;
  const int tt = (tile / p.n_tiles) % p.tiles_t;
  const int b = tile / (p.n_tiles * p.tiles_t);
  const int num_kb = p.n_taps * p.kb_a0 + p.kb_a1;
  const int kb_begin = static_cast<int>(rank) * num_kb / kSplitK, kb_end = static_cast<int>(rank + 1) * num_kb / kSplitK;

  if (warp == 0) {
    {
      // ------------------------------------------------------------ TMA producer: this CTA's K blocks
      const bool lead = elect_one();
      int stage = 0;
      uint32_t phase = 0;
      const int t0 = tt * 128;
      const int row0 = p.b_row0 + nt * BN;
      const int kb_taps = p.n_taps * p.kb_a0;
      for (int kb = kb_begin; kb < kb_end; ++kb) {
        mbar_wait(&empty[stage], phase ^ 1);
        uint8_t* sa = smem_a + stage * Cfg::kABytes;
        uint8_t* sb = smem_b + stage * Cfg::kBBytes;
        if (lead) {
          mbar_arrive_expect_tx(&full[stage], Cfg::kABytes + Cfg::kBBytes);
          if (kb < kb_taps) {
            const int tap = kb / p.kb_a0;
            const int cb = kb - tap * p.kb_a0;
            tma_load_3d(&p.tmA0, &full[stage], sa, p.a0_c0 + cb * 64, t0 + (tap - p.n_taps / 2) * p.dil, b);
          } else {
            tma_load_3d(&p.tmA1, &full[stage], sa, (kb - kb_taps) * 64, t0, b);
          }
          tma_load_2d(&p.tmB, &full[stage], sb, p.b_k0 + kb * 64, row0);
        }
        if (++stage == S) {
          stage = 0;
          phase ^= 1;
        }
      }
    }
  } else if (warp == 1) {
    {
      // ------------------------------------------------------------ MMA issuer
      const bool lead = elect_one();
      constexpr uint32_t idesc = make_idesc_bf16_f32(BN);
      int stage = 0;
      uint32_t phase = 0;
      for (int kb = kb_begin; kb < kb_end; ++kb) {
        mbar_wait(&full[stage], phase);
        tc_fence_after();
        const uint64_t da = make_kmajor_sw128_desc(smem_u32(smem_a + stage * Cfg::kABytes));
        const uint64_t db = make_kmajor_sw128_desc(smem_u32(smem_b + stage * Cfg::kBBytes));
        if (lead) {
#pragma unroll
          for (int k = 0; k < 4; ++k) umma_bf16(tmem_base, da + 2 * k, db + 2 * k, idesc, (kb != kb_begin || k != 0));
          umma_commit(&empty[stage]);
          if (kb == kb_end - 1) umma_commit(tmem_full);
        }
        if (++stage == S) {
          stage = 0;
          phase ^= 1;
        }
      }
    }
  } else if (warp >= 4 && rank != 0) {
    // -------------------------------------------------------------- peers: accumulator -> own shared memory
    const int quad = warp & 3;
    const int half = (warp - 4) >> 2;
    const int row = quad * 32 + lane;
    mbar_wait(tmem_full, 0);
    tc_fence_after();
    const uint32_t taddr = tmem_base + (static_cast<uint32_t>(quad * 32) << 16);
    const uint32_t pbase = smem_u32(part) + row * 16;
    for (int c = half; c < BN / 32; c += 2) {
      uint32_t v[32];
      tmem_ld_32x32(taddr + c * 32, v);
      tmem_ld_wait();
#pragma unroll
      for (int g4 = 0; g4 < 8; ++g4)  // [column / 4][frame][4]
        st_shared_v4(pbase + (c * 8 + g4) * 2048, v[4 * g4], v[4 * g4 + 1], v[4 * g4 + 2], v[4 * g4 + 3]);
    }
  }
  tc_fence_before();
  cluster_sync_all();  // #1: the peers' partial tiles are complete and visible cluster-wide
  if (warp >= 4 && rank == 0) {
    // ---------------------------------------------------------------- leader: reduce into TMEM, then the fused epilogue
    const int quad = warp & 3;
    const int half = (warp - 4) >> 2;
    const int row = quad * 32 + lane;
    mbar_wait(tmem_full, 0);
    tc_fence_after();
    const uint32_t taddr = tmem_base + (static_cast<uint32_t>(quad * 32) << 16);
    uint32_t peer[kSplitK - 1];
#pragma unroll
    for (int r = 1; r < kSplitK; ++r) peer[r - 1] = dsmem_addr(smem_u32(part), r) + row * 16;
    for (int c = half; c < BN / 32; c += 2) {
      uint32_t v[32];
      tmem_ld_32x32(taddr + c * 32, v);
      float4 q[kSplitK - 1][8];
#pragma unroll
      for (int r = 0; r < kSplitK - 1; ++r)
#pragma unroll
        for (int g4 = 0; g4 < 8; ++g4) q[r][g4] = ld_dsmem_v4(peer[r] + (c * 8 + g4) * 2048);
      tmem_ld_wait();
#pragma unroll
      for (int g4 = 0; g4 < 8; ++g4) {
        float s0 = __uint_as_float(v[4 * g4]), s1 = __uint_as_float(v[4 * g4 + 1]), s2 = __uint_as_float(v[4 * g4 + 2]),
              s3 = __uint_as_float(v[4 * g4 + 3]);
#pragma unroll
        for (int r = 0; r < kSplitK - 1; ++r) { s0 += q[r][g4].x; s1 += q[r][g4].y; s2 += q[r][g4].z; s3 += q[r][g4].w; }
        v[4 * g4] = __float_as_uint(s0); v[4 * g4 + 1] = __float_as_uint(s1); v[4 * g4 + 2] = __float_as_uint(s2);
        v[4 * g4 + 3] = __float_as_uint(s3);
      }
      tmem_st_32x32(taddr + c * 32, v);
    }
    tmem_st_wait();
    tc_fence_before();
    asm volatile("bar.sync 1, 256;" ::: "memory");  // the epilogue's chunk order differs per EPI: all sums are in TMEM first
    tc_fence_after();
    const int t = tt * 128 + row;
    epilogue_tile<BN, EPI, false>(p, taddr, b, t, t < p.T, nt, half);
  }
  tc_fence_before();
  cluster_sync_all();  // #2: the leader has read the peers' shared memory
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc<BN>(tmem_base);
  }
}

}  // namespace dsb
