// Persistent, warp-specialised tcgen05 GEMM for the WaveNet denoiser (sm_100a).
//
//   D[frame, n] = sum_k A[frame, k] * W[n, k]        (frames are the UMMA M dimension, 128 per tile)
//
// A is "time-major" activation memory [B][T][channels] (bf16) read through 3-D TMA tensor maps, so
//   * a dilated-conv tap is just a shifted time coordinate (t0 + (tap-1)*dil) and TMA's out-of-bounds
//     zero fill IS the conv's zero padding (per utterance, no cross-batch leakage),
//   * the 1x1 conditioner projection is extra K blocks from a second tensor map (K-concatenation).
// W is a packed bf16 matrix [rows][K] (K-major) read through a 2-D tensor map.
// Accumulators live in TMEM (2 x BN fp32 columns, double buffered) and are drained by 4 epilogue warps
// (thread <-> frame/TMEM lane) that apply the fused epilogue and write global memory directly.
//
// Warp roles (384 threads): warp0 = TMA producer (1 elected lane), warp1 = MMA issuer (1 lane),
// warp2 = TMEM allocator, warp3 idle, warps4..11 = epilogue (two warps per TMEM lane quadrant).
#pragma once
#include <cuda.h>
#include "ptx.cuh"

namespace dsb {

constexpr int kGemmThreads = 384;  // 4 control warps + 8 epilogue warps (two per TMEM lane quadrant)
constexpr int kEpiThreads = 256;

enum : int { EPI_LINEAR = 0, EPI_GATE = 1, EPI_RES = 2, EPI_OUTF32 = 3, EPI_CPROJ = 4, EPI_SWIGLU = 5, EPI_GATEBWD = 6 };

struct GemmParams {
  CUtensorMap tmA0;  // primary activation source  (dims: channels, T, B)
  CUtensorMap tmA1;  // secondary activation source (conditioner), used when kb_a1 > 0
  CUtensorMap tmB;   // packed weights (dims: K, rows)
  CUtensorMap tmBh;  // the same weights with a 128-row box (CTA-pair kernel, gemm_pair_kernel.cuh); valid when has_pair
  int has_pair;
  CUtensorMap tmOut; // bf16 output as a store map (dims: ld_out channels, T, B; box 32 x 32, SWIZZLE_64B); valid when has_out_map
  int has_out_map;   // the pair kernel's epilogue warps then write through TMA stores instead of per-lane STG.128
  int T, B, tiles_t, n_tiles, total_tiles;
  const int* len;  // optional [B] (ragged batch, ds_session_set_lengths): bf16 outputs of frames t >= len[b] are written as ZEROS,
                   // so that the next dilated convolution reads the utterance's own zero padding there
  int a0_c0;   // channel offset inside A0 (selects a layer's slice of the gated-activation store)
  int kb_a0;   // 64-wide K blocks per tap from A0
  int n_taps;  // 1 (plain 1x1) or 3 (dilated conv)
  int dil;
  int kb_a1;   // 64-wide K blocks from A1
  int b_row0;  // first row of this GEMM inside the packed weight matrix
  int b_k0;    // first K column of this GEMM inside the packed weight matrix (hoisted conditioner projection: 3C)
  int relu;    // activation of the LINEAR / OUTF32 epilogues: 0 none, 1 ReLU, 2 GELU (exact, erf), 3 LeakyReLU(slope2), 4 tanh
  __nv_bfloat16* out2;  // EPI_RES, optional: second output LeakyReLU(out, slope2) with out's ld / offset
  float slope2;
  float scale;  // LINEAR / OUTF32: multiplies act(acc + bias); RES: multiplies (acc + res_in)  (1/sqrt(2) in WaveNet)
  const float* bias;     // indexed by tile-global n (packed order for EPI_GATE)
  const float* vec_sub;  // unused (kept for ABI stability of the debug hook)
  const float* vec_add;  // LINEAR: vector added after relu/scale; RES: folded (b - d_l)/sqrt(2) + d_{l+1}
  int vec_bstride;
  const __nv_bfloat16* res_in;  // EPI_RES: x~ of this layer, [B][T][ld_res]
  int ld_res;
  void* out;   // bf16 (LINEAR/GATE/RES) or fp32 (OUTF32), [B][T][ld_out]
  int ld_out;
  int out_c0;  // channel offset in out
  int n_real;  // EPI_OUTF32: number of valid output columns
  // fp32-accurate ("split") mode, kSplit instantiations only: every bf16 operand is stored as hi = bf16(v) followed,
  // `*_lo` elements further along the channel / K axis, by lo = bf16(v - hi); each logical 64-wide K block is issued
  // three times (hi*hi, lo*hi, hi*lo) into the same fp32 accumulator.
  int a0_lo, a1_lo, w_lo;  // offsets of the lo halves in A0 / A1 (channels) and W (K)
  int out_lo, res_lo;      // offsets of the lo halves in out / res_in (channels)
  // EPI_CPROJ over several layers in one launch: output column n belongs to group n / grp_n and is stored at
  // out + (n / grp_n) * grp_stride + row * ld_out + n % grp_n  (elements; grp_n = 0: one group)
  int grp_n;
  long grp_stride;
  // LayerNorm folded into the GEMMs around it (LYNXNet strong_cond, lynxnet.py:52-56 with the step vector added before the norm):
  //   producer (LINEAR / RES epilogues): per row and per (N tile, warp half) the partial sums {sum r, sum r^2, sum r d, 0} of the
  //   bf16-ROUNDED outputs r and the next norm's step vector d -> stats_out[row][parts] (float4), parts = 2 * n_tiles;
  //   consumer (SWIGLU epilogue): mean / rstd of (r + d) over ln_channels from those sums (+ sum d, sum d^2 stored behind e), and
  //   y[n] = rstd * (acc[n] + e[n] - mean * s[n]) + bias[n]  with  s = W~ 1,  e = W~ d,  W~ = W diag(gamma),  bias = W beta + b.
  float4* stats_out;
  const float* stats_d;  // [Bt][N] (row stride stats_d_bstride; 0 = one vector for every utterance); may be null (d = 0)
  int stats_d_bstride;
  const float4* ln_stats;
  int ln_parts, ln_channels;
  float ln_eps;
  const float* ln_s;     // [N] packed column order
  const float* ln_e;     // [Bt][N + 4]: e, then sum d, sum d^2 (row stride ln_e_bstride; 0 = shared)
  int ln_e_bstride, ln_n;
};

template <int BN>
struct GemmCfg {
  static constexpr int kStages = BN == 256 ? 4 : (BN == 128 ? 6 : 8);
  static constexpr uint32_t kABytes = 128 * 128;  // 128 frames x 64 bf16
  static constexpr uint32_t kBBytes = BN * 128;   // BN rows  x 64 bf16
  static constexpr uint32_t kTmemCols = 2 * BN;   // 128 / 256 / 512: powers of two
  static constexpr uint32_t kWarpStage = 2048;    // per epilogue warp: 32 frames x 32 channels bf16 (TMA-store staging, SWIZZLE_64B)
  static constexpr uint32_t kSmemBytes = 1024 + kStages * (kABytes + kBBytes) + 8 * kWarpStage + 256;
};

// Where an epilogue warp stages its 32 x 32 output sub-tiles for TMA stores (warp_store_tile); on = false: direct stores.
struct EpiStage {
  uint32_t sw = 0;  // shared-memory address of the warp's 2 KB staging tile
  int lane = 0;
  int tw0 = 0;      // first frame of the warp's 32 rows
  bool on = false;
};

template <int BN, int EPI, bool kSplit>
__device__ __forceinline__ void epilogue_tile(const GemmParams& p, uint32_t taddr, int b, int t, bool valid, int nt, int half,
                                              const EpiStage& es = EpiStage{});

template <int BN, int EPI, bool kSplit = false>
__global__ void __launch_bounds__(kGemmThreads, 1) ds_gemm_kernel(const __grid_constant__ GemmParams p) {
  constexpr int kSub = kSplit ? 3 : 1;  // sub-blocks per logical K block
  using Cfg = GemmCfg<BN>;
  constexpr int S = Cfg::kStages;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024u - (raw_addr & 1023u)) & 1023u);  // SWIZZLE_128B needs 1024-B alignment
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + S * Cfg::kABytes;
  uint8_t* warp_stage = smem + S * (Cfg::kABytes + Cfg::kBBytes);  // 8 x kWarpStage, 2 KB aligned
  uint64_t* bars = reinterpret_cast<uint64_t*>(warp_stage + 8 * Cfg::kWarpStage);
  uint64_t* full = bars;
  uint64_t* empty = bars + S;
  uint64_t* tmem_full = bars + 2 * S;
  uint64_t* tmem_empty = bars + 2 * S + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * S + 4);

  // warp index through a shuffle: the compiler then knows it is warp-uniform, and the single-lane roles below can run as CONVERGED
  // warps (waits and index arithmetic on all lanes, in uniform registers) with one elected lane issuing -- inside `if (lane == 0)`
  // ptxas wraps every tcgen05.mma / commit / TMA issue in a vector->uniform register waterfall loop (~55 cycles per MMA)
  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
  const int lane = threadIdx.x & 31;

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
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tmem_full[i], 1);
      mbar_init(&tmem_empty[i], kEpiThreads);
    }
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc<Cfg::kTmemCols>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  grid_dep_launch();  // the next kernel may begin its own prologue now ...
  grid_dep_wait();    // ... and this one touches global memory only after its predecessor has completed

  const int num_kb = p.n_taps * p.kb_a0 + p.kb_a1;

  if (warp == 0) {
    {
      // ------------------------------------------------------------ TMA producer
      const bool lead = elect_one();
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x) {
        const int nt = tile % p.n_tiles;
        const int tt = (tile / p.n_tiles) % p.tiles_t;
        const int b = tile / (p.n_tiles * p.tiles_t);
        const int t0 = tt * 128;
        const int row0 = p.b_row0 + nt * BN;
        const int kb_taps = p.n_taps * p.kb_a0;
        for (int kb = 0; kb < num_kb; ++kb) {
#pragma unroll 1
          for (int sub = 0; sub < kSub; ++sub) {  // split: (A hi, W hi), (A lo, W hi), (A hi, W lo)
            mbar_wait(&empty[stage], phase ^ 1);
            uint8_t* sa = smem_a + stage * Cfg::kABytes;
            uint8_t* sb = smem_b + stage * Cfg::kBBytes;
            if (lead) {
              mbar_arrive_expect_tx(&full[stage], Cfg::kABytes + Cfg::kBBytes);
              if (kb < kb_taps) {
                const int tap = kb / p.kb_a0;
                const int cb = kb - tap * p.kb_a0;
                tma_load_3d(&p.tmA0, &full[stage], sa, p.a0_c0 + cb * 64 + (sub == 1 ? p.a0_lo : 0),
                            t0 + (tap - p.n_taps / 2) * p.dil, b);
              } else {
                tma_load_3d(&p.tmA1, &full[stage], sa, (kb - kb_taps) * 64 + (sub == 1 ? p.a1_lo : 0), t0, b);
              }
              tma_load_2d(&p.tmB, &full[stage], sb, p.b_k0 + kb * 64 + (sub == 2 ? p.w_lo : 0), row0);
            }
            if (++stage == S) {
              stage = 0;
              phase ^= 1;
            }
          }
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
      int it = 0;
      for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x, ++it) {
        const int as = it & 1;
        const uint32_t aphase = (it >> 1) & 1;
        mbar_wait(&tmem_empty[as], aphase ^ 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + as * BN;
        for (int kb = 0; kb < num_kb * kSub; ++kb) {
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          const uint64_t da = make_kmajor_sw128_desc(smem_u32(smem_a + stage * Cfg::kABytes));
          const uint64_t db = make_kmajor_sw128_desc(smem_u32(smem_b + stage * Cfg::kBBytes));
          if (lead) {
#pragma unroll
            for (int k = 0; k < 4; ++k)  // 4 x (K = 16) per 64-wide block: +32 B on the start address
              umma_bf16(tmem_d, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0);
            umma_commit(&empty[stage]);
            if (kb == num_kb * kSub - 1) umma_commit(&tmem_full[as]);
          }
          if (++stage == S) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp >= 4) {
    // -------------------------------------------------------------- epilogue (256 threads)
    const int quad = warp & 3;  // TMEM lane quadrant this warp may read
    const int half = (warp - 4) >> 2;
    int it = 0;
    for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x, ++it) {
      const int nt = tile % p.n_tiles;
      const int tt = (tile / p.n_tiles) % p.tiles_t;
      const int b = tile / (p.n_tiles * p.tiles_t);
      const int as = it & 1;
      const uint32_t aphase = (it >> 1) & 1;
      mbar_wait(&tmem_full[as], aphase);
      tc_fence_after();
      const int t = tt * 128 + quad * 32 + lane;
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(quad * 32) << 16) + as * BN;
      EpiStage es;
      es.sw = smem_u32(warp_stage + (warp - 4) * Cfg::kWarpStage);
      es.lane = lane;
      es.tw0 = tt * 128 + quad * 32;
      es.on = !kSplit && p.has_out_map != 0;
      epilogue_tile<BN, EPI, kSplit>(p, taddr, b, t, t < p.T, nt, half, es);
      tc_fence_before();
      mbar_arrive(&tmem_empty[as]);
    }
    bulk_wait_all();  // staging tiles are read out before the CTA's shared memory goes away
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc<Cfg::kTmemCols>(tmem_base);
  }
}

}  // namespace dsb
