// CTA-pair (cta_group::2) fused residual-layer kernel with a TIME-TILED activation stage that carries the dilation
// halo (reference: ResidualBlock.forward, modules/backbones/wavenet.py:33-48; dilated_conv = Conv1d(C, 2C, 3,
// padding=dilation, dilation=dilation), wavenet.py:22-28).
//
// layer2_kernel.cuh stages every (tap, 64-channel block) of the dilated conv as its own 128-frame TMA tile, i.e. it
// moves x three times per pass.  Measured on B200 the kernel is bound by L2 -> SM operand traffic, and the
// activation (A) tiles are the expensive half of it (the weight tiles are read by every SM at once and dedup in L2).
// Here one TMA box of 128 + 2*dil frames x 64 channels is staged ONCE and the three taps are three UMMA descriptors
// into the same shared-memory tile, shifted by dil rows (the tile is K-major SWIZZLE_128B with a 128-byte row pitch,
// so a row shift is a plain +dil*128 B on the descriptor start address).
// A traffic per pass drops from 28 to (8 * (1 + 2*dil/128) + 4) sixteen-KB blocks at C = 512, H = 256.
//
// Operands move through two rings: A slots (halo tile / conditioner tile / gated-activation tile) and B slots (this
// CTA's 128-row half of a 256-row weight block).  One A item serves 3 (conv) or 1 (conditioner, GEMM-2) "units";
// one unit = one B slot = 4 UMMAs of M = 256 (pair) x N = 256 x K = 16.
// Roles, job order, accumulator double buffering and epilogue math are those of layer2_kernel.cuh.
#pragma once
#include "layer2_kernel.cuh"

namespace dsb {

constexpr int kHaloNB = 6;          // B ring depth
constexpr int kHaloMaxNA = 8;       // A ring depth limit (actual depth is LayerParams::n_a)
constexpr uint32_t kHaloSmemMax = 232448;

__host__ __device__ inline uint32_t halo_a_slot_bytes(int max_dil) {
  const uint32_t b = (128u + 2u * max_dil) * 128u;
  return (b + 1023u) / 1024u * 1024u;
}
__host__ __device__ inline int halo_n_a(int max_dil) {
  const int n = static_cast<int>((kHaloSmemMax - 1024u - 512u - kHaloNB * kPairBBytes) / halo_a_slot_bytes(max_dil));
  return n > kHaloMaxNA ? kHaloMaxNA : n;
}
__host__ __device__ inline uint32_t halo_smem_bytes(int max_dil) {
  return 1024u + halo_n_a(max_dil) * halo_a_slot_bytes(max_dil) + kHaloNB * kPairBBytes + 512u;
}

// K-major SWIZZLE_128B descriptor whose first row is `row` rows into a 1024-byte-aligned tile.  The tensor core applies
// the swizzle XOR to absolute shared-memory address bits (as TMA did when it wrote the tile), so a row shift is just a
// start-address offset; the descriptor's base-offset field must stay 0 (setting it to the row phase was measured WRONG).
__device__ __forceinline__ uint64_t make_kmajor_sw128_desc_row(uint32_t tile_addr, uint32_t row) {
  return make_kmajor_sw128_desc(tile_addr + row * 128u);
}

struct HaloRing {
  int a, b;
  uint32_t pa, pb;
  __device__ HaloRing() : a(0), b(0), pa(0), pb(0) {}
  __device__ void next_a(int n_a) { if (++a == n_a) { a = 0; pa ^= 1; } }
  __device__ void next_b() { if (++b == kHaloNB) { b = 0; pb ^= 1; } }
};

__device__ __forceinline__ void pair_epilogue_gate(const LayerParams& p, uint32_t taddr, int idx, size_t row, bool valid, int half) {
  const float* bias = p.bias1 + idx * 256;
  __nv_bfloat16* zg = p.z_out + row * p.ld_z + p.z_c0 + idx * 128;
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
}

__device__ __forceinline__ void pair_epilogue_res(const LayerParams& p, uint32_t taddr, int idx, int b, size_t row, bool valid, int half) {
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

// Requires C % 256 == 0 and 2^(cycle-1) <= 16.  LayerParams: tmXh = halo map of x_in (box 64 x (128 + 2*dil) x 1),
// tmW1 / tmW2 with 128-row boxes, tiles_t = 256-frame pair tiles per utterance, a_slot / n_a from halo_*().
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kGemmThreads, 1)
ds_layer_halo_kernel(const __grid_constant__ LayerParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024u - (raw_addr & 1023u)) & 1023u);
  const int NA = p.n_a;
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem_a + NA * p.a_slot;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_b + kHaloNB * kPairBBytes);
  uint64_t* a_full = bars;                    // leader: 2 producer arrivals + the bytes of both CTAs
  uint64_t* a_empty = a_full + kHaloMaxNA;    // per CTA: released by the leader's multicast commit
  uint64_t* b_full = a_empty + kHaloMaxNA;
  uint64_t* b_empty = b_full + kHaloNB;
  uint64_t* tmem_full = b_empty + kHaloNB;
  uint64_t* tmem_empty = tmem_full + 2;       // leader: 16 warp arrivals (8 per CTA)
  uint64_t* z_full = tmem_empty + 2;          // per CTA
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(z_full + 1);

  const int warp = threadIdx.x >> 5;
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
    for (int i = 0; i < NA; ++i) {
      mbar_init(&a_full[i], 2);
      mbar_init(&a_empty[i], 1);
    }
    for (int i = 0; i < kHaloNB; ++i) {
      mbar_init(&b_full[i], 2);
      mbar_init(&b_empty[i], 1);
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
  const bool tracing = p.trace != nullptr && blockIdx.x == 0 && p.p2 > 0;

  if (warp == 0) {
    if (lane == 0) {
      // ------------------------------------------------------------ TMA producer (both CTAs)
      HaloRing r;
      auto a_item = [&](const CUtensorMap* map, uint32_t bytes, int c0, int t, int b) {
        mbar_wait(&a_empty[r.a], r.pa ^ 1);
        if (p.dbg & 2) {  // debug: barriers only, no bytes
          if (leader) mbar_arrive(&a_full[r.a]);
          else mbar_arrive_cluster(&a_full[r.a], 0);
        } else {
          if (leader) mbar_arrive_expect_tx(&a_full[r.a], 2 * bytes);
          else mbar_arrive_cluster(&a_full[r.a], 0);
          tma_load_3d_pair(map, &a_full[r.a], smem_a + r.a * p.a_slot, c0, t, b);
        }
        r.next_a(NA);
      };
      auto b_item = [&](const CUtensorMap* map, int k0, int row0) {
        mbar_wait(&b_empty[r.b], r.pb ^ 1);
        if (p.dbg & 2) {
          if (leader) mbar_arrive(&b_full[r.b]);
          else mbar_arrive_cluster(&b_full[r.b], 0);
        } else {
          if (leader) mbar_arrive_expect_tx(&b_full[r.b], 2 * kPairBBytes);
          else mbar_arrive_cluster(&b_full[r.b], 0);
          tma_load_2d_pair(map, &b_full[r.b], smem_b + r.b * kPairBBytes, k0, row0);
        }
        r.next_b();
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
            a_item(&p.tmXh, halo_bytes, cb * 64, t0 - p.dil, b);
            for (int tap = 0; tap < 3; ++tap) b_item(&p.tmW1, (tap * p.kb_c + cb) * 64, row0);
          }
          for (int h = 0; h < p.kb_h; ++h) {
            a_item(&p.tmCond, kLayerABytes, h * 64, t0, b);
            b_item(&p.tmW1, (3 * p.kb_c + h) * 64, row0);
          }
        } else {
          if (idx == 0) mbar_wait(z_full, j & 1);  // this CTA's z rows of tile j are in global memory
          const int row0 = p.w2_row0 + idx * 256 + rank * 128;
          for (int kb = 0; kb < p.kb_c; ++kb) {
            a_item(&p.tmZ, kLayerABytes, p.z_c0 + kb * 64, t0, b);
            b_item(&p.tmW2, kb * 64, row0);
          }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0 && leader) {
      // ------------------------------------------------------------ MMA issuer (leader CTA only)
      constexpr uint32_t idesc = make_idesc_bf16_f32_m(256, 256);
      HaloRing r;
      uint32_t acc_it = 0;
      // one unit: 4 UMMAs (K = 64) of A rows [row, row + 128) of slot r.a against B slot r.b
      auto unit = [&](uint32_t tmem_d, uint32_t row, bool first_of_item, bool last_of_item, bool first_of_job, bool last_of_job,
                      uint32_t as) {
        if (first_of_item) mbar_wait(&a_full[r.a], r.pa);
        mbar_wait(&b_full[r.b], r.pb);
        tc_fence_after();
        const uint64_t da = make_kmajor_sw128_desc_row(smem_u32(smem_a + r.a * p.a_slot), row);
        const uint64_t db = make_kmajor_sw128_desc(smem_u32(smem_b + r.b * kPairBBytes));
        if (!(p.dbg & 4)) {  // debug bit 2: commits only, no MMAs (operand-delivery rate alone)
#pragma unroll
          for (int k = 0; k < 4; ++k) umma_bf16_pair(tmem_d, da + 2 * k, db + 2 * k, idesc, (!first_of_job) || k != 0);
        }
        umma_commit_pair(&b_empty[r.b]);
        r.next_b();
        if (last_of_item) {
          umma_commit_pair(&a_empty[r.a]);
          r.next_a(NA);
        }
        if (last_of_job) umma_commit_pair(&tmem_full[as]);
      };
      LayerJobs jobs(n_local, p.p1, p.p2, true);
      int kind, j, idx;
      int ntrace = 0;
      while (jobs.next(kind, j, idx)) {
        const uint32_t as = acc_it & 1;
        const unsigned long long ts0 = tracing ? globaltimer_ns() : 0ull;
        mbar_wait(&tmem_empty[as], ((acc_it >> 1) & 1) ^ 1);
        tc_fence_after();
        if (tracing && ntrace < p.trace_cap / 2) {
          unsigned long long* e = p.trace + 4 * ntrace++;
          e[0] = 16 + kind; e[1] = j * 16 + idx; e[2] = ts0; e[3] = globaltimer_ns();
        }
        const uint32_t tmem_d = tmem_base + as * 256;
        if (kind == 0) {
          for (int cb = 0; cb < p.kb_c; ++cb)
            for (int tap = 0; tap < 3; ++tap)
              unit(tmem_d, tap * p.dil, tap == 0, tap == 2, cb == 0 && tap == 0, false, as);
          for (int h = 0; h < p.kb_h; ++h) unit(tmem_d, 0, true, true, false, h == p.kb_h - 1, as);
        } else {
          for (int kb = 0; kb < p.kb_c; ++kb) unit(tmem_d, 0, true, true, kb == 0, kb == p.kb_c - 1, as);
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
    int ntrace = 0;
    LayerJobs jobs(n_local, p.p1, p.p2, true);
    int kind, j, idx;
    while (jobs.next(kind, j, idx)) {
      const int tile = pair_id + j * n_pairs;
      const int b = tile / p.tiles_t;
      const int t = (tile % p.tiles_t) * 256 + rank * 128 + rr;
      const bool valid = t < p.T;
      const size_t row = static_cast<size_t>(b) * p.T + (valid ? t : 0);
      const uint32_t as = acc_it & 1;
      mbar_wait(&tmem_full[as], (acc_it >> 1) & 1);
      tc_fence_after();
      const uint32_t taddr = lane_addr + as * 256;
      const unsigned long long ts0 = (tracing && threadIdx.x == 128) ? globaltimer_ns() : 0ull;
      if (p.dbg & 1) {
      } else if (kind == 0) {
        pair_epilogue_gate(p, taddr, idx, row, valid, half);
      } else {
        pair_epilogue_res(p, taddr, idx, b, row, valid, half);
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
        if (!(p.dbg & 1)) __threadfence();  // z stores reach L2 ...
        fence_proxy_async_global();         // ... and are ordered before the TMA (async proxy) re-read
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
