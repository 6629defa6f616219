// One WaveNet residual layer as ONE persistent tcgen05 kernel on CTA pairs (reference: ResidualBlock.forward,
// modules/backbones/wavenet.py:33-48; dilated_conv = Conv1d(C, 2C, 3, padding=dilation, dilation=dilation), wavenet.py:22-28).
//
// Two CTAs of a cluster -- the two SMs of a TPC -- work on 256 consecutive frames (128 each) and run two back-to-back GEMMs per tile
// without leaving the SM:
//   GEMM-1  Y[256 x 2C] = [x~(t-d) | x~(t) | x~(t+d) | cond] (K = 3C+H)  x  W1^T   in 2C/256 passes of N = 256 (128 gate + 128 filter
//           columns of the same channels); epilogue z = sigmoid(gate) * tanh(filter) -> bf16 -> the gated-activation store
//   GEMM-2  R[256 x C] = z (K = C) x W2^T (residual half of output_projection); epilogue x~' = (x~ + R)/sqrt(2) + rvec
// Every UMMA is cta_group::2, M = 256 x N = 256: each CTA supplies its own 128 activation rows and HALF of the 256 weight rows, so
// the weight bytes per SM are halved; only the leader issues MMAs, its commits are multicast to both CTAs.
//   * halo tiles: the conv's activation stage is ONE TMA box of (128 + 2*dil) frames x 64 channels; the three taps are three UMMA
//     descriptors into the same tile shifted by dil rows (K-major SWIZZLE_128B, 128-byte row pitch: a row shift is +dil*128 B on
//     the descriptor start address)
//   * grouped pipeline stages: one full/empty mbarrier pair guards a whole group of K blocks -- a conv stage = one halo tile + the
//     three tap weight blocks (12 UMMAs per hand-shake), a 1x1 stage = two activation + two weight blocks (8 UMMAs) -- because the
//     hand-shake itself (mbarrier operations cost their thread ~285 cycles each) is as expensive as four UMMAs
//   * TMEM: 2 x 256 fp32 columns, double buffered; GEMM-2 chunks of tile k are interleaved with GEMM-1 passes of tile k+1 (LayerJobs)
//   * epilogue warps stage 32 x 32 bf16 sub-tiles in shared memory and write them with TMA stores (warp_store_tile, ptx.cuh)
// Requires C % 256 == 0, H % 128 == 0 and 2^(cycle-1) <= 16; other shapes run the two-GEMM path (gemm_kernel.cuh).
// (Round 1 also carried four earlier generations of this kernel -- single CTA with z in shared memory / through L2, CTA pairs
// without halo tiles, CTA pairs with per-K-block stages; they were removed in round 2, their measurements stay in profiles/r01_*.)
#pragma once
#include <cuda.h>
#include "gemm_epilogue.cuh"

namespace dsb {

struct LayerParams {
  CUtensorMap tmX, tmCond, tmW1, tmW2;
  CUtensorMap tmZ;  // gated-activation store (dims L*C, T, B): A operand of GEMM-2 in the kZSmem = false variant
  int T, B, tiles_t, total_tiles;
  const int* len;  // optional [B] (ragged batch): x~ rows of frames t >= len[b] are written as zeros (the utterance's own conv padding)
  int kb_c, kb_h;  // C/64, H/64
  int dil;
  int p1, p2;      // GEMM-1 passes (2C/256), GEMM-2 chunks (C/BN2; 0 for the last layer)
  int w1_row0, w2_row0;
  int stages;
  const float* bias1;  // [2C] in packed (pass-permuted) order
  const float* rvec;   // folded residual vector (b2 - d_l)/sqrt(2) + d_{l+1}, [vec_bstride * b + n]
  int vec_bstride;
  const __nv_bfloat16* x_in;  // [B][T][C]
  __nv_bfloat16* x_out;       // [B][T][C]
  __nv_bfloat16* z_out;       // [B][T][ld_z], this layer's slice starts at z_c0
  int ld_z, z_c0;
  unsigned long long* trace;  // optional: CTA 0 records {role*16 + kind, local tile*16 + idx, t0, t1} (debug)
  int trace_cap;
  int dbg;  // debug: bit0 = epilogue warps skip all work, bit1 = producer skips the TMA loads (barriers only)
  // x_in map with a (64 x (128 + 2*dil) x 1) box: one halo-carrying activation tile feeds the three conv taps
  CUtensorMap tmXh;
  uint32_t a_slot;
  int n_a;
  // TMA STORE maps (box 32 channels x 32 frames, SWIZZLE_64B) of the
  // gated-activation store and of x_out -- the epilogue warps stage their 32 x 32 sub-tiles in shared memory
  CUtensorMap tmZs, tmXs;
  const uint16_t* bias1h;  // [2C] f16 copy of bias1 in the same packed order, gate entries pre-multiplied by 0.5
  // hoisted conditioner projection (grouped kernel): f16 [B][T][ld_cproj = 2C] of THIS layer, packed column order; holds W_cond * cond + b_dil + b_cond (gate half * 0.5), computed once per sampler call.  NULL: not hoisted
  // (kb_h conditioner K blocks run inside the kernel and bias1h is added instead).
  const uint16_t* cproj;
  int ld_cproj;
  // fp32-accurate ("split") instantiation: every bf16 operand is hi = bf16(v) followed, `*_lo` elements further along its channel
  // / K axis, by lo = bf16(v - hi); each K stage is issued three times (hi*hi, lo*hi, hi*lo) into the same accumulator
  int x_lo, cond_lo, z_lo, w1_lo, w2_lo;
  // ... and its hoisted conditioner projection: fp32 [B*T][2C] of THIS layer in packed column order (W_cond * cond + b_dil + b_cond,
  // computed once per sampler call by the split GEMM); NULL: the conditioner K blocks run inside the kernel and bias1 is added
  const float* cproj32;
};

constexpr uint32_t kLayerABytes = 128 * 128;  // 128 frames x 64 bf16
constexpr uint32_t kPairBBytes = 128 * 128;   // this CTA's half of a 256-row weight tile: 128 rows x 64 bf16

__device__ __forceinline__ void fence_proxy_async_global() { asm volatile("fence.proxy.async.global;" ::: "memory"); }
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

// Static job list of one CTA.  kind 0 = GEMM-1 pass `idx` of local tile `j`; kind 1 = GEMM-2 chunk `idx` of tile `j`.
struct LayerJobs {
  int n, p1, p2;
  bool interleave;
  int g, i, sub;  // group g: passes of tile g + chunks of tile g-1 (interleaved), or passes then chunks of tile g
  __device__ LayerJobs(int n_, int p1_, int p2_, bool il) : n(n_), p1(p1_), p2(p2_), interleave(il), g(0), i(0), sub(0) {}
  __device__ bool next(int& kind, int& j, int& idx) {
    if (interleave) {
      const int G = p1 > p2 ? p1 : p2;
      while (g <= n) {
        while (i < G) {
          while (sub < 2) {
            const int s = sub++;
            if (s == 0 && i < p1 && g < n) { kind = 0; j = g; idx = i; return true; }
            if (s == 1 && i < p2 && g >= 1) { kind = 1; j = g - 1; idx = i; return true; }
          }
          sub = 0;
          ++i;
        }
        i = 0;
        ++g;
      }
      return false;
    }
    while (g < n) {
      if (i < p1) { kind = 0; j = g; idx = i++; return true; }
      if (i < p1 + p2) { kind = 1; j = g; idx = i++ - p1; return true; }
      i = 0;
      ++g;
    }
    return false;
  }
};

__host__ __device__ inline uint32_t halo_a_slot_bytes(int max_dil) {
  const uint32_t b = (128u + 2u * max_dil) * 128u;
  return (b + 1023u) / 1024u * 1024u;
}

// K-major SWIZZLE_128B descriptor whose first row is `row` rows into a 1024-byte-aligned tile.  The tensor core applies
// the swizzle XOR to absolute shared-memory address bits (as TMA did when it wrote the tile), so a row shift is just a
// start-address offset; the descriptor's base-offset field must stay 0 (setting it to the row phase was measured WRONG).
__device__ __forceinline__ uint64_t make_kmajor_sw128_desc_row(uint32_t tile_addr, uint32_t row) {
  return make_kmajor_sw128_desc(tile_addr + row * 128u);
}

constexpr int kGrpMaxStages = 3;
constexpr uint32_t kGrpSmemMax = 232448;

__host__ __device__ inline uint32_t grp_stage_bytes(int max_dil) { return halo_a_slot_bytes(max_dil) + 3u * kPairBBytes; }
__host__ __device__ inline int grp_stages(int max_dil) {
  const int n = static_cast<int>((kGrpSmemMax - 1024u - 2048u - 8u * 2048u) / grp_stage_bytes(max_dil));
  return n > kGrpMaxStages ? kGrpMaxStages : n;
}
__host__ __device__ inline uint32_t grp_smem_bytes(int max_dil) {
  return 1024u + grp_stages(max_dil) * grp_stage_bytes(max_dil) + 2048u + 8u * 2048u;
}

constexpr uint32_t kGrpWarpStage = 2048;  // per epilogue warp: 32 frames x 32 channels bf16, SWIZZLE_64B

// Gate epilogue in packed half precision: z = sigmoid(g + bg) * tanh(f + bf) with sigmoid(x) = 0.5 * tanh(0.5 x) + 0.5.
// The fp32 accumulators are rounded to f16 (11-bit mantissa; pre-activations are O(1..10)), everything after that runs two
// channels per instruction (HFMA2 / MUFU.TANH.F16x2), and z is rounded to bf16 as before -- 11 instructions per channel
// pair instead of 15, half the MUFU work.  bias1h holds 0.5 * bg (gate) and bf (filter) as f16.
// `add` = the f16 row added to the pre-activations: the broadcast bias (bias1h) or, with the conditioner projection hoisted,
// this lane's row of cproj; (bg0, bf0) = its first chunk, loaded by the caller before it waits for the accumulator.
__device__ __forceinline__ void grp_load_add(const uint16_t* add, int c, uint4 (&bg)[4], uint4 (&bf)[4]) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    bg[i] = __ldg(reinterpret_cast<const uint4*>(add + c * 32) + i);
    bf[i] = __ldg(reinterpret_cast<const uint4*>(add + 128 + c * 32) + i);
  }
}
__device__ __forceinline__ void grp_epilogue_gate(const LayerParams& p, uint32_t taddr, int idx, int b, int tw0, const uint16_t* add,
                                                  const uint4 (&bg0)[4], const uint4 (&bf0)[4], int half, int lane, uint32_t sw) {
  constexpr uint32_t kHalf2 = 0x38003800u;  // (0.5h, 0.5h)
  uint4 bg1[4], bf1[4];
  grp_load_add(add, half + 2, bg1, bf1);  // second chunk's rows: issued first so that an HBM miss hides behind chunk one
#pragma unroll
  for (int k = 0; k < 2; ++k) {
    const int c = half + 2 * k;
    uint32_t g[32], f[32];
    tmem_ld_32x32(taddr + c * 32, g);
    tmem_ld_32x32(taddr + 128 + c * 32, f);
    uint4 bg[4], bf[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      bg[i] = k == 0 ? bg0[i] : bg1[i];
      bf[i] = k == 0 ? bf0[i] : bf1[i];
    }
    tmem_ld_wait();
    uint32_t zp[16];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const uint32_t bgw[4] = {bg[i].x, bg[i].y, bg[i].z, bg[i].w};
      const uint32_t bfw[4] = {bf[i].x, bf[i].y, bf[i].z, bf[i].w};
#pragma unroll
      for (int h = 0; h < 4; ++h) {
        const int q = 4 * i + h;
        const uint32_t gq = fma_f16x2(pack_f16(__uint_as_float(g[2 * q]), __uint_as_float(g[2 * q + 1])), kHalf2, bgw[h]);
        const uint32_t fq = add_f16x2(pack_f16(__uint_as_float(f[2 * q]), __uint_as_float(f[2 * q + 1])), bfw[h]);
        const uint32_t sg = fma_f16x2(tanh_f16x2(gq), kHalf2, kHalf2);
        zp[q] = f16x2_to_bf16x2(mul_f16x2(sg, tanh_f16x2(fq)));
      }
    }
    warp_store_tile(&p.tmZs, sw, lane, zp, p.z_c0 + idx * 128 + c * 32, tw0, b, tw0 < p.T);
  }
}

// Residual epilogue.  `x0` = this lane's first 32-channel chunk of x~ (prefetched by the caller BEFORE it waits for the
// accumulator: x~ does not depend on the MMAs).  The chunk loop is fully unrolled so that the one-chunk-ahead prefetch of
// x~ lives in renamed registers (a rolled loop ends in register moves that wait for the prefetch at the END of each
// iteration -- the top epilogue stall in the ncu source view).
__device__ __forceinline__ void grp_epilogue_res(const LayerParams& p, uint32_t taddr, int idx, int b, int tw0, const __nv_bfloat16* xin,
                                                 const uint4 (&x0)[4], int half, int lane, uint32_t sw, bool live) {
  const int n0 = idx * 256;
  const float* rv = p.rvec + static_cast<size_t>(b) * p.vec_bstride + n0;
  uint4 xq[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) xq[i] = x0[i];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int c = half + 2 * k;
    uint32_t a[32];
    tmem_ld_32x32(taddr + c * 32, a);
    float rr[32];
    ld_f32x32(rv + c * 32, rr);
    uint4 xn[4];
    if (k < 3) ld_x4(xin + (c + 2) * 32, xn);
    tmem_ld_wait();
    uint32_t o[16];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const uint32_t w[4] = {xq[i].x, xq[i].y, xq[i].z, xq[i].w};
#pragma unroll
      for (int h = 0; h < 4; ++h) {
        const int q = 8 * i + 2 * h;
        const float v0 = fmaf(__uint_as_float(a[q]) + __uint_as_float(w[h] << 16), 0.70710678118654752f, rr[q]);
        const float v1 = fmaf(__uint_as_float(a[q + 1]) + __uint_as_float(w[h] & 0xffff0000u), 0.70710678118654752f, rr[q + 1]);
        o[4 * i + h] = live ? pack_bf16(v0, v1) : 0u;
      }
    }
    warp_store_tile(&p.tmXs, sw, lane, o, n0 + c * 32, tw0, b, tw0 < p.T);
    if (k < 3) {
#pragma unroll
      for (int i = 0; i < 4; ++i) xq[i] = xn[i];
    }
  }
}

// fp32-accurate instantiation: gate and residual epilogues in full precision (expf / tanhf, fp32 bias), outputs as hi / lo bf16
// pairs (two staged TMA stores per 32 x 32 sub-tile: hi at its channel, lo `*_lo` channels further).
__device__ __forceinline__ void split_store_tile(const CUtensorMap* map, uint32_t sw, int lane, const float (&v)[32], int c0, int lo_off,
                                                 int tw0, int b, bool any) {
  uint32_t hi[16], lo[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    const float a = v[2 * i], c = v[2 * i + 1];
    const float ah = __bfloat162float(__float2bfloat16_rn(a)), ch = __bfloat162float(__float2bfloat16_rn(c));
    hi[i] = pack_bf16(a, c);
    lo[i] = pack_bf16(a - ah, c - ch);
  }
  warp_store_tile(map, sw, lane, hi, c0, tw0, b, any);
  warp_store_tile(map, sw, lane, lo, c0 + lo_off, tw0, b, any);
}
__device__ __forceinline__ void grp_epilogue_gate_split(const LayerParams& p, uint32_t taddr, int idx, int b, int tw0, int half, int lane,
                                                        uint32_t sw, size_t row) {
  const float* bias = (p.cproj32 ? p.cproj32 + row * (p.kb_c * 128) : p.bias1) + idx * 256;  // this frame's row, or the broadcast bias
#pragma unroll 1
  for (int k = 0; k < 2; ++k) {
    const int c = half + 2 * k;
    uint32_t g[32], f[32];
    tmem_ld_32x32(taddr + c * 32, g);
    tmem_ld_32x32(taddr + 128 + c * 32, f);
    float bg[32], bf[32], z[32];
    ld_f32x32(bias + c * 32, bg);
    ld_f32x32(bias + 128 + c * 32, bf);
    tmem_ld_wait();
#pragma unroll
    for (int j = 0; j < 32; ++j) z[j] = sigmoid_accurate(__uint_as_float(g[j]) + bg[j]) * tanhf(__uint_as_float(f[j]) + bf[j]);
    split_store_tile(&p.tmZs, sw, lane, z, p.z_c0 + idx * 128 + c * 32, p.z_lo, tw0, b, tw0 < p.T);
  }
}
__device__ __forceinline__ void grp_epilogue_res_split(const LayerParams& p, uint32_t taddr, int idx, int b, int tw0, const __nv_bfloat16* xin,
                                                       int half, int lane, uint32_t sw, bool live) {
  const int n0 = idx * 256;
  const float* rv = p.rvec + static_cast<size_t>(b) * p.vec_bstride + n0;
#pragma unroll 1
  for (int k = 0; k < 4; ++k) {
    const int c = half + 2 * k;
    uint32_t a[32];
    tmem_ld_32x32(taddr + c * 32, a);
    float rr[32], v[32];
    ld_f32x32(rv + c * 32, rr);
    uint4 xh[4], xl[4];
    ld_x4(xin + c * 32, xh);
    ld_x4(xin + p.x_lo + c * 32, xl);
    tmem_ld_wait();
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const uint32_t wh[4] = {xh[i].x, xh[i].y, xh[i].z, xh[i].w};
      const uint32_t wl[4] = {xl[i].x, xl[i].y, xl[i].z, xl[i].w};
#pragma unroll
      for (int h = 0; h < 4; ++h) {
        const int q = 8 * i + 2 * h;
        const float x0 = __uint_as_float(wh[h] << 16) + __uint_as_float(wl[h] << 16);
        const float x1 = __uint_as_float(wh[h] & 0xffff0000u) + __uint_as_float(wl[h] & 0xffff0000u);
        v[q] = live ? fmaf(__uint_as_float(a[q]) + x0, 0.70710678118654752f, rr[q]) : 0.0f;
        v[q + 1] = live ? fmaf(__uint_as_float(a[q + 1]) + x1, 0.70710678118654752f, rr[q + 1]) : 0.0f;
      }
    }
    split_store_tile(&p.tmXs, sw, lane, v, n0 + c * 32, p.x_lo, tw0, b, tw0 < p.T);
  }
}

// Requires C % 256 == 0, H % 128 == 0 (two-block 1x1 stages) and 2^(cycle-1) <= 16.  LayerParams as for the halo kernel;
// a_slot = halo_a_slot_bytes(max_dil), n_a = grp_stages(max_dil).
template <bool kSplit>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kGemmThreads, 1)
ds_layer_grp_kernel(const __grid_constant__ LayerParams p) {
  constexpr int kSub = kSplit ? 3 : 1;  // split: (A hi, W hi), (A lo, W hi), (A hi, W lo)
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024u - (raw_addr & 1023u)) & 1023u);
  const int NS = p.n_a;
  const uint32_t stage_bytes = p.a_slot + 3u * kPairBBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + NS * stage_bytes);
  uint64_t* full = bars;                      // leader: 2 producer arrivals + the bytes of both CTAs
  uint64_t* empty = full + kGrpMaxStages;     // per CTA: released by the leader's multicast commit
  uint64_t* tmem_full = empty + kGrpMaxStages;
  uint64_t* tmem_empty = tmem_full + 2;       // leader: 16 warp arrivals (8 per CTA)
  uint64_t* z_full = tmem_empty + 2;          // per CTA
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(z_full + 1);
  uint8_t* warp_stage = reinterpret_cast<uint8_t*>(bars) + 2048;  // 8 x kGrpWarpStage, 2 KB aligned (stage_bytes % 2048 == 0)

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);  // warp-uniform to the compiler: converged single-issuer roles (gemm_kernel.cuh)
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&p.tmXh);
    tma_prefetch_desc(&p.tmCond);
    tma_prefetch_desc(&p.tmW1);
    tma_prefetch_desc(&p.tmW2);
    tma_prefetch_desc(&p.tmZ);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < NS; ++i) {
      mbar_init(&full[i], 2);
      mbar_init(&empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tmem_full[i], 1);
      mbar_init(&tmem_empty[i], 16);
    }
    mbar_init(z_full, kEpiThreads);
    fence_barrier_init();
  }
  cluster_sync_all();  // peer barriers are initialised before any remote arrive / multicast commit
  if (warp == 2) tmem_alloc_pair<512>(tmem_slot);
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int pair_id = blockIdx.x >> 1;
  const int n_pairs = gridDim.x >> 1;
  const int n_local = (p.total_tiles > pair_id) ? (p.total_tiles - 1 - pair_id) / n_pairs + 1 : 0;
  const uint32_t halo_bytes = (128u + 2u * p.dil) * 128u;
  // stage layout: conv  [halo A tile (a_slot)] [B tap0] [B tap1] [B tap2]
  //               1x1   [A blk0 16K] [A blk1 16K] [B blk0 16K] [B blk1 16K]      (fits: a_slot + 48K >= 64K)
  constexpr uint32_t kOne = kLayerABytes;  // 16 KB

  if (warp == 0) {
    {
      // ------------------------------------------------------------ TMA producer (both CTAs)
      const bool lead = elect_one();
      const bool load = lead && !(p.dbg & 2);
      int st = 0;
      uint32_t ph = 0;
      auto begin_stage = [&](uint32_t bytes_per_cta) -> uint8_t* {
        mbar_wait(&empty[st], ph ^ 1);
        if (lead) {
          if (leader) {
            if (p.dbg & 2) mbar_arrive(&full[st]);  // debug: barriers only, no bytes
            else mbar_arrive_expect_tx(&full[st], 2 * bytes_per_cta);
          } else {
            mbar_arrive_cluster(&full[st], 0);
          }
        }
        return smem + st * stage_bytes;
      };
      auto end_stage = [&]() { if (++st == NS) { st = 0; ph ^= 1; } };
      // 1x1 stages: blocks [k, k + n) of activation map `am` (channel offset ac0) against weight columns wk0 + ...
      auto one_by_one = [&](const CUtensorMap* am, int ac0, int a_lo, const CUtensorMap* wm, int wk0, int w_lo, int nblk, int row0, int t0,
                            int b) {
        for (int k = 0; k < nblk; k += 2) {
          const int n = nblk - k < 2 ? nblk - k : 2;
#pragma unroll 1
          for (int sub = 0; sub < kSub; ++sub) {
            uint8_t* s = begin_stage(n * (kOne + kPairBBytes));
            for (int j = 0; j < n && load; ++j) {
              tma_load_3d_pair(am, &full[st], s + j * kOne, ac0 + (k + j) * 64 + (sub == 1 ? a_lo : 0), t0, b);
              tma_load_2d_pair(wm, &full[st], s + 2 * kOne + j * kPairBBytes, wk0 + (k + j) * 64 + (sub == 2 ? w_lo : 0), row0);
            }
            end_stage();
          }
        }
      };
      LayerJobs jobs(n_local, p.p1, p.p2, true);
      int kind, j, idx;
      while (jobs.next(kind, j, idx)) {
        const int tile = pair_id + j * n_pairs;
        const int t0 = (tile % p.tiles_t) * 256 + rank * 128;
        const int b = tile / p.tiles_t;
        if (kind == 0) {
          const int row0 = p.w1_row0 + idx * 256 + rank * 128;
          for (int cb = 0; cb < p.kb_c; ++cb) {
#pragma unroll 1
            for (int sub = 0; sub < kSub; ++sub) {
              uint8_t* s = begin_stage(halo_bytes + 3 * kPairBBytes);
              if (load) tma_load_3d_pair(&p.tmXh, &full[st], s, cb * 64 + (sub == 1 ? p.x_lo : 0), t0 - p.dil, b);
#pragma unroll
              for (int tap = 0; tap < 3 && load; ++tap)
                tma_load_2d_pair(&p.tmW1, &full[st], s + p.a_slot + tap * kPairBBytes, (tap * p.kb_c + cb) * 64 + (sub == 2 ? p.w1_lo : 0),
                                 row0);
              end_stage();
            }
          }
          one_by_one(&p.tmCond, 0, p.cond_lo, &p.tmW1, 3 * p.kb_c * 64, p.w1_lo, p.kb_h, row0, t0, b);
        } else {
          if (idx == 0) mbar_wait(z_full, j & 1);  // this CTA's z rows of tile j are in global memory
          const int row0 = p.w2_row0 + idx * 256 + rank * 128;
          one_by_one(&p.tmZ, p.z_c0, p.z_lo, &p.tmW2, 0, p.w2_lo, p.kb_c, row0, t0, b);
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
      uint32_t acc_it = 0;
      auto mma4 = [&](uint32_t tmem_d, uint64_t da, uint64_t db, bool fresh) {
        if ((p.dbg & 4) || !lead) return;  // debug: commits only
#pragma unroll
        for (int k = 0; k < 4; ++k) umma_bf16_pair(tmem_d, da + 2 * k, db + 2 * k, idesc, (!fresh) || k != 0);
      };
      auto release = [&](bool last_of_job, uint32_t as) {
        if (lead) {
          umma_commit_pair(&empty[st]);
          if (last_of_job) umma_commit_pair(&tmem_full[as]);
        }
        if (++st == NS) { st = 0; ph ^= 1; }
      };
      auto one_by_one = [&](uint32_t tmem_d, int nblk, bool fresh_first, uint32_t as) {
        for (int k = 0; k < nblk; k += 2) {
          const int n = nblk - k < 2 ? nblk - k : 2;
#pragma unroll 1
          for (int sub = 0; sub < kSub; ++sub) {
            mbar_wait(&full[st], ph);
            tc_fence_after();
            const uint32_t s = smem_u32(smem + st * stage_bytes);
            for (int jj = 0; jj < n; ++jj)
              mma4(tmem_d, make_kmajor_sw128_desc(s + jj * kOne), make_kmajor_sw128_desc(s + 2 * kOne + jj * kPairBBytes),
                   fresh_first && k == 0 && jj == 0 && sub == 0);
            release(k + n >= nblk && sub == kSub - 1, as);
          }
        }
      };
      LayerJobs jobs(n_local, p.p1, p.p2, true);
      int kind, j, idx;
      while (jobs.next(kind, j, idx)) {
        const uint32_t as = acc_it & 1;
        mbar_wait(&tmem_empty[as], ((acc_it >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + as * 256;
        if (kind == 0) {
          for (int cb = 0; cb < p.kb_c; ++cb) {
#pragma unroll 1
            for (int sub = 0; sub < kSub; ++sub) {
              mbar_wait(&full[st], ph);
              tc_fence_after();
              const uint32_t s = smem_u32(smem + st * stage_bytes);
#pragma unroll
              for (int tap = 0; tap < 3; ++tap)
                mma4(tmem_d, make_kmajor_sw128_desc_row(s, tap * p.dil), make_kmajor_sw128_desc(s + p.a_slot + tap * kPairBBytes),
                     cb == 0 && tap == 0 && sub == 0);
              release(p.kb_h == 0 && cb == p.kb_c - 1 && sub == kSub - 1, as);
            }
          }
          one_by_one(tmem_d, p.kb_h, false, as);
        } else {
          one_by_one(tmem_d, p.kb_c, true, as);
        }
        ++acc_it;
      }
    }
  } else if (warp >= 4) {
    // -------------------------------------------------------------- epilogue (both CTAs; 256 threads each)
    const int quad = warp & 3;
    const int half = (warp - 4) >> 2;
    const int rr = quad * 32 + lane;
    const uint32_t lane_addr = tmem_base + (static_cast<uint32_t>(quad * 32) << 16);
    uint32_t acc_it = 0;
    LayerJobs jobs(n_local, p.p1, p.p2, true);
    int kind, j, idx;
    while (jobs.next(kind, j, idx)) {
      const int tile = pair_id + j * n_pairs;
      const int b = tile / p.tiles_t;
      const int t = (tile % p.tiles_t) * 256 + rank * 128 + rr;
      const bool valid = t < p.T;
      const size_t row = static_cast<size_t>(b) * p.T + (valid ? t : 0);
      const uint32_t as = acc_it & 1;
      const __nv_bfloat16* xin = p.x_in + row * (p.kb_c * 64 * (kSplit ? 2 : 1)) + idx * 256;
      uint4 x0[4], bg0[4], bf0[4];
      const uint16_t* add = p.cproj ? p.cproj + row * p.ld_cproj + idx * 256 : p.bias1h + idx * 256;
      if (!kSplit && !(p.dbg & 1)) {  // in flight while the job's MMAs finish
        if (kind == 1) ld_x4(xin + half * 32, x0);
        else grp_load_add(add, half, bg0, bf0);
      }
      mbar_wait(&tmem_full[as], (acc_it >> 1) & 1);
      tc_fence_after();
      const uint32_t taddr = lane_addr + as * 256;
      const int tw0 = (tile % p.tiles_t) * 256 + rank * 128 + quad * 32;  // first frame of this warp's 32 rows
      const uint32_t sw = smem_u32(warp_stage + (warp - 4) * kGrpWarpStage);
      const bool live = p.len == nullptr || t < __ldg(p.len + b);
      if (p.dbg & 1) {  // debug: epilogue warps only hand the accumulator back
      } else if constexpr (kSplit) {
        if (kind == 0) grp_epilogue_gate_split(p, taddr, idx, b, tw0, half, lane, sw, row);
        else grp_epilogue_res_split(p, taddr, idx, b, tw0, xin, half, lane, sw, live);
      } else if (kind == 0) {
        grp_epilogue_gate(p, taddr, idx, b, tw0, add, bg0, bf0, half, lane, sw);
      } else {
        grp_epilogue_res(p, taddr, idx, b, tw0, xin, x0, half, lane, sw, live);
      }
      // release the accumulator on the LEADER's barrier: one arrival per warp
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (leader) mbar_arrive(&tmem_empty[as]);
        else mbar_arrive_cluster(&tmem_empty[as], 0);
      }
      if (kind == 0 && idx == p.p1 - 1 && p.p2 > 0) {
        bulk_wait_all();  // this warp's TMA stores of z are complete before the producer's TMA re-reads them
        mbar_arrive(z_full);
      }
      ++acc_it;
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
