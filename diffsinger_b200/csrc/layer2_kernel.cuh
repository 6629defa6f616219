// CTA-pair (cta_group::2) version of the fused residual-layer kernel (layer_kernel.cuh, kZSmem = false flavour).
//
// Two CTAs of a cluster -- the two SMs of a TPC -- work on 256 consecutive frames (128 each).  Every UMMA is
// M = 256 x N = 256: each CTA supplies its own 128 activation rows (A) and HALF of the 256 weight rows (B), and
// receives its 128 x 256 accumulator slice in its own TMEM.  Per SM this halves the weight bytes that cross
// L2 -> SMEM and the shared-memory write/read traffic per FLOP, and the 16 KB + 16 KB stages allow a 6-deep ring.
//   * both CTAs run a TMA producer; all transaction bytes are credited to the LEADER CTA's `full` barrier
//   * only the leader issues tcgen05.mma; its commits are multicast to both CTAs' `empty` / `tmem_full` barriers
//   * both CTAs run epilogue warps; they release the accumulator on the leader's `tmem_empty` barrier
// Job order and epilogue math are identical to layer_kernel.cuh.
#pragma once
#include "layer_kernel.cuh"

namespace dsb {

constexpr uint32_t kPairBBytes = 128 * 128;  // this CTA's half of the weight tile: 128 rows x 64 bf16
constexpr int kPairStages = 6;
constexpr uint32_t kPairSmemBytes = 1024u + kPairStages * (kLayerABytes + kPairBBytes) + 256u;

// Requires BN2 == 256 (C % 256 == 0).  LayerParams.tmW1 / tmW2 must be encoded with a 128-row box;
// tiles_t = number of 256-frame PAIR tiles per utterance, total_tiles = B * tiles_t.
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kGemmThreads, 1)
ds_layer_pair_kernel(const __grid_constant__ LayerParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024u - (raw_addr & 1023u)) & 1023u);
  constexpr int S = kPairStages;
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem_a + S * kLayerABytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_b + S * kPairBBytes);
  uint64_t* full = bars;            // used on the leader: 2 producer arrivals + all bytes of both CTAs
  uint64_t* empty = bars + S;       // per CTA: released by the leader's multicast commit
  uint64_t* tmem_full = bars + 2 * S;
  uint64_t* tmem_empty = tmem_full + 2;  // used on the leader: 16 warp arrivals (8 per CTA)
  uint64_t* z_full = tmem_empty + 2;     // per CTA
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(z_full + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&p.tmX);
    tma_prefetch_desc(&p.tmCond);
    tma_prefetch_desc(&p.tmW1);
    tma_prefetch_desc(&p.tmW2);
    tma_prefetch_desc(&p.tmZ);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < S; ++i) {
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

  const int kb_taps = 3 * p.kb_c;
  const int kb1 = kb_taps + p.kb_h;
  const int pair_id = blockIdx.x >> 1;
  const int n_pairs = gridDim.x >> 1;
  const int n_local = (p.total_tiles > pair_id) ? (p.total_tiles - 1 - pair_id) / n_pairs + 1 : 0;
  const bool tracing = p.trace != nullptr && blockIdx.x == 0 && p.p2 > 0;
  // De-synchronise the pairs: every pair walks the K blocks (and the passes) in its own rotated order, so that at any
  // instant the chip reads many different weight blocks instead of all SMs hitting the same few L2 slices.
  const int prot = (p.dbg & 8) ? 0 : pair_id % p.p1;
  const int krot1 = (p.dbg & 4) ? 0 : (pair_id / p.p1 * 5) % kb1;
  const int krot2 = (p.dbg & 4) ? 0 : pair_id % p.kb_c;

  if (warp == 0) {
    if (lane == 0) {
      // ------------------------------------------------------------ TMA producer (both CTAs)
      int stage = 0;
      uint32_t phase = 0;
      LayerJobs jobs(n_local, p.p1, p.p2, true);
      int kind, j, idx;
      while (jobs.next(kind, j, idx)) {
        const int tile = pair_id + j * n_pairs;
        const int t0 = (tile % p.tiles_t) * 256 + rank * 128;
        const int b = tile / p.tiles_t;
        const int nkb = kind == 0 ? kb1 : p.kb_c;
        const int pidx = kind == 0 ? (idx + prot) % p.p1 : idx;
        const int row0 = (kind == 0 ? p.w1_row0 + pidx * 256 : p.w2_row0 + idx * 256) + rank * 128;
        if (kind == 1 && idx == 0) mbar_wait(z_full, j & 1);  // this CTA's z rows of tile j are in global memory
        for (int ki = 0; ki < nkb; ++ki) {
          int kb = ki + (kind == 0 ? krot1 : krot2);
          if (kb >= nkb) kb -= nkb;
          mbar_wait(&empty[stage], phase ^ 1);
          if (p.dbg & 2) {
            if (leader) mbar_arrive(&full[stage]); else mbar_arrive_cluster(&full[stage], 0);
            if (++stage == S) { stage = 0; phase ^= 1; }
            continue;
          }
          const bool ldA = !(p.dbg & 16), ldB = !(p.dbg & 32);
          if (leader) mbar_arrive_expect_tx(&full[stage], 2 * ((ldA ? kLayerABytes : 0u) + (ldB ? kPairBBytes : 0u)));
          else mbar_arrive_cluster(&full[stage], 0);
          uint8_t* sa = smem_a + stage * kLayerABytes;
          uint8_t* sb = smem_b + stage * kPairBBytes;
          if (kind == 0) {
            if (ldA) {
              if (kb < kb_taps) {
                const int tap = kb / p.kb_c;
                const int cb = kb - tap * p.kb_c;
                tma_load_3d_pair(&p.tmX, &full[stage], sa, cb * 64, t0 + (tap - 1) * p.dil, b);
              } else {
                tma_load_3d_pair(&p.tmCond, &full[stage], sa, (kb - kb_taps) * 64, t0, b);
              }
            }
            if (ldB) tma_load_2d_pair(&p.tmW1, &full[stage], sb, kb * 64, row0);
          } else {
            if (ldA) tma_load_3d_pair(&p.tmZ, &full[stage], sa, p.z_c0 + kb * 64, t0, b);
            if (ldB) tma_load_2d_pair(&p.tmW2, &full[stage], sb, kb * 64, row0);
          }
          if (++stage == S) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0 && leader) {
      // ------------------------------------------------------------ MMA issuer (leader CTA only)
      constexpr uint32_t idesc = make_idesc_bf16_f32_m(256, 256);
      int stage = 0;
      uint32_t phase = 0;
      uint32_t acc_it = 0;
      int ntrace = 0;
      LayerJobs jobs(n_local, p.p1, p.p2, true);
      int kind, j, idx;
      while (jobs.next(kind, j, idx)) {
        const uint32_t as = acc_it & 1;
        const unsigned long long ts0 = tracing ? globaltimer_ns() : 0ull;
        mbar_wait(&tmem_empty[as], ((acc_it >> 1) & 1) ^ 1);
        tc_fence_after();
        const unsigned long long ts1 = tracing ? globaltimer_ns() : 0ull;
        if (tracing && ntrace < p.trace_cap / 2) {
          unsigned long long* e = p.trace + 4 * ntrace++;
          e[0] = 16 + kind; e[1] = j * 16 + idx; e[2] = ts0; e[3] = ts1;
        }
        const uint32_t tmem_d = tmem_base + as * 256;
        const int nkb = kind == 0 ? kb1 : p.kb_c;
        for (int kb = 0; kb < nkb; ++kb) {
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          const uint64_t da = make_kmajor_sw128_desc(smem_u32(smem_a + stage * kLayerABytes));
          const uint64_t db = make_kmajor_sw128_desc(smem_u32(smem_b + stage * kPairBBytes));
#pragma unroll
          for (int k = 0; k < 4; ++k) umma_bf16_pair(tmem_d, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0);
          umma_commit_pair(&empty[stage]);
          if (kb == nkb - 1) umma_commit_pair(&tmem_full[as]);
          if (++stage == S) { stage = 0; phase ^= 1; }
        }
        ++acc_it;
      }
    }
  } else if (warp >= 4) {
    // -------------------------------------------------------------- epilogue (both CTAs; 256 threads each)
    const int quad = warp & 3;
    const int half = (warp - 4) >> 2;
    const int r = quad * 32 + lane;
    const uint32_t lane_addr = tmem_base + (static_cast<uint32_t>(quad * 32) << 16);
    uint32_t acc_it = 0;
    int ntrace = 0;
    LayerJobs jobs(n_local, p.p1, p.p2, true);
    int kind, j, idx;
    while (jobs.next(kind, j, idx)) {
      const int tile = pair_id + j * n_pairs;
      const int b = tile / p.tiles_t;
      const int t = (tile % p.tiles_t) * 256 + rank * 128 + r;
      const bool valid = t < p.T;
      const size_t row = static_cast<size_t>(b) * p.T + (valid ? t : 0);
      const uint32_t as = acc_it & 1;
      mbar_wait(&tmem_full[as], (acc_it >> 1) & 1);
      tc_fence_after();
      const unsigned long long ts0 = (tracing && threadIdx.x == 128) ? globaltimer_ns() : 0ull;
      const uint32_t taddr = lane_addr + as * 256;
      if (p.dbg & 1) {
      } else if (kind == 0) {
        const int pidx = (idx + prot) % p.p1;
        const float* bias = p.bias1 + pidx * 256;
        __nv_bfloat16* zg = p.z_out + row * p.ld_z + p.z_c0 + pidx * 128;
#pragma unroll 1
        for (int c = half; c < 4; c += 2) {
          uint32_t g[32], f[32];
          tmem_ld_32x32(taddr + c * 32, g);
          tmem_ld_32x32(taddr + 128 + c * 32, f);
          float bg[32], bf[32];
          ld_f32x32(bias + c * 32, bg);
          ld_f32x32(bias + 128 + c * 32, bf);
          tmem_ld_wait();
          uint32_t zp[16];
#pragma unroll
          for (int q = 0; q < 16; ++q) {
            float z2[2];
#pragma unroll
            for (int h = 0; h < 2; ++h) {
              const float gj = __uint_as_float(g[2 * q + h]) + bg[2 * q + h];
              const float fj = __uint_as_float(f[2 * q + h]) + bf[2 * q + h];
              z2[h] = fmaf(0.5f, tanh_approx(0.5f * gj), 0.5f) * tanh_approx(fj);
            }
            zp[q] = pack_bf16(z2[0], z2[1]);
          }
          if (valid) {
#pragma unroll
            for (int i = 0; i < 4; ++i)
              reinterpret_cast<uint4*>(zg + c * 32)[i] = make_uint4(zp[4 * i], zp[4 * i + 1], zp[4 * i + 2], zp[4 * i + 3]);
          }
        }
      } else {
        const int n0 = idx * 256;
        const float* rv = p.rvec + static_cast<size_t>(b) * p.vec_bstride + n0;
        const __nv_bfloat16* xin = p.x_in + row * (p.kb_c * 64) + n0;
        __nv_bfloat16* xout = p.x_out + row * (p.kb_c * 64) + n0;
        uint4 xq[4];
        ld_x4(xin + half * 32, xq);
#pragma unroll 1
        for (int c = half; c < 8; c += 2) {
          uint32_t a[32];
          tmem_ld_32x32(taddr + c * 32, a);
          float v[32], rr[32];
          ld_f32x32(rv + c * 32, rr);
          uint4 xn[4];
          if (c + 2 < 8) ld_x4(xin + (c + 2) * 32, xn);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const uint32_t w[4] = {xq[i].x, xq[i].y, xq[i].z, xq[i].w};
#pragma unroll
            for (int h = 0; h < 4; ++h) {
              const int q = 8 * i + 2 * h;
              v[q] = fmaf(__uint_as_float(a[q]) + __uint_as_float(w[h] << 16), 0.70710678118654752f, rr[q]);
              v[q + 1] = fmaf(__uint_as_float(a[q + 1]) + __uint_as_float(w[h] & 0xffff0000u), 0.70710678118654752f, rr[q + 1]);
            }
          }
          if (valid) st_bf16x32(xout + c * 32, v);
#pragma unroll
          for (int i = 0; i < 4; ++i) xq[i] = xn[i];
        }
      }
      if (tracing && threadIdx.x == 128 && ntrace < p.trace_cap / 2) {
        unsigned long long* e = p.trace + 4 * (p.trace_cap / 2 + ntrace++);
        e[0] = kind; e[1] = j * 16 + idx; e[2] = ts0; e[3] = globaltimer_ns();
      }
      // release the accumulator on the LEADER's barrier: one arrival per warp
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (leader) mbar_arrive(&tmem_empty[as]);
        else mbar_arrive_cluster(&tmem_empty[as], 0);
      }
      if (kind == 0 && idx == p.p1 - 1 && p.p2 > 0) {
        if (!(p.dbg & 1)) __threadfence();
        fence_proxy_async_global();
        mbar_arrive(z_full);
      }
      ++acc_it;
    }
  }
  tc_fence_before();
  cluster_sync_all();  // the leader's MMAs read this CTA's shared memory and write its TMEM until the very end
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc_pair<512>(tmem_base);
  }
}

}  // namespace dsb
