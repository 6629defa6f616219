// One WaveNet residual layer as ONE persistent tcgen05 kernel (reference: ResidualBlock.forward,
// modules/backbones/wavenet.py:33-48), for C a multiple of 128.
//
// Per 128-frame tile a CTA runs two back-to-back GEMMs without leaving the SM:
//   GEMM-1  Y[128 x 2C] = [x~(t-d) | x~(t) | x~(t+d) | cond] (K = 3C+H)  x  W1^T      (C/128 passes "P" of N = 256:
//           128 gate + 128 filter columns of the same channels, TMEM double buffered)
//           epilogue: z = sigmoid(gate) * tanh(filter) -> bf16 -> gated-activation store in HBM (also the operand
//           of the deferred skip GEMM), and in the kZSmem variant a K-major SWIZZLE_128B copy in shared memory
//   GEMM-2  R[128 x C] = z (K = C)  x  W2^T   (residual half of output_projection; C/BN2 chunks "C")
//           epilogue: x~_next = (x~ + R)/sqrt(2) + [(b - d_l)/sqrt(2) + d_{l+1}]  -> bf16 -> HBM (ping-pong buffer)
// TMA feeds both GEMMs through one mbarrier ring.  All three roles walk the same static job list; in the default
// (kZSmem = false) variant the chunks of tile k are interleaved with the passes of tile k+1
//        P0(k+1) C0(k) P1(k+1) C1(k) P2(k+1) P3(k+1) ...
// so the z round trip (epilogue -> L2 -> TMA) and the long residual epilogues hide behind GEMM-1 work.
#pragma once
#include <cuda.h>
#include "gemm_epilogue.cuh"

namespace dsb {

struct LayerParams {
  CUtensorMap tmX, tmCond, tmW1, tmW2;
  CUtensorMap tmZ;  // gated-activation store (dims L*C, T, B): A operand of GEMM-2 in the kZSmem = false variant
  int T, B, tiles_t, total_tiles;
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
  // halo variant (layer3_kernel.cuh): x_in map with a (64 x (128 + 2*dil) x 1) box, A ring geometry
  CUtensorMap tmXh;
  uint32_t a_slot;
  int n_a;
  // grouped-stage variant (layer4_kernel.cuh): TMA STORE maps (box 32 channels x 32 frames, SWIZZLE_64B) of the
  // gated-activation store and of x_out -- the epilogue warps stage their 32 x 32 sub-tiles in shared memory
  CUtensorMap tmZs, tmXs;
  const uint16_t* bias1h;  // [2C] f16 copy of bias1 in the same packed order, gate entries pre-multiplied by 0.5
  // hoisted conditioner projection (grouped kernel): f16 [B][T][ld_cproj = 2C] of THIS layer, packed column order; holds W_cond * cond + b_dil + b_cond (gate half * 0.5), computed once per sampler call.  NULL: not hoisted
  // (kb_h conditioner K blocks run inside the kernel and bias1h is added instead).
  const uint16_t* cproj;
  int ld_cproj;
};

constexpr uint32_t kLayerABytes = 128 * 128;  // 128 frames x 64 bf16
constexpr uint32_t kLayerBBytes = 256 * 128;  // up to 256 weight rows x 64 bf16
constexpr int kLayerMaxStages = 4;

__host__ __device__ inline uint32_t layer_smem_bytes(int kb_c, int stages, bool z_smem) {
  return 1024u + (z_smem ? kb_c : 0) * kLayerABytes + stages * (kLayerABytes + kLayerBBytes) + 256u;
}

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

// kZSmem = true : z stays in shared memory between the two GEMMs (needs C*256 B of smem -> few TMA stages at C=512;
//                 jobs run tile by tile)
// kZSmem = false: GEMM-2 re-reads the CTA's own z tile from the gated-activation store through TMA (L2 hit), which
//                 frees the shared memory for a 4-deep operand ring and allows the interleaved job order.
template <int BN2, bool kZSmem>
__global__ void __launch_bounds__(kGemmThreads, 1) ds_layer_kernel(const __grid_constant__ LayerParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024u - (raw_addr & 1023u)) & 1023u);
  const int S = p.stages;
  uint8_t* smem_z = smem;  // kZSmem: kb_c blocks of [128 x 64] bf16, SW128 K-major
  uint8_t* smem_a = smem_z + (kZSmem ? p.kb_c : 0) * kLayerABytes;
  uint8_t* smem_b = smem_a + S * kLayerABytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_b + S * kLayerBBytes);
  uint64_t* full = bars;
  uint64_t* empty = bars + kLayerMaxStages;
  uint64_t* tmem_full = bars + 2 * kLayerMaxStages;
  uint64_t* tmem_empty = tmem_full + 2;
  uint64_t* z_full = tmem_empty + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(z_full + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&p.tmX);
    tma_prefetch_desc(&p.tmCond);
    tma_prefetch_desc(&p.tmW1);
    tma_prefetch_desc(&p.tmW2);
    if (!kZSmem) tma_prefetch_desc(&p.tmZ);
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
    mbar_init(z_full, kEpiThreads);
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc<512>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int kb_taps = 3 * p.kb_c;
  const int kb1 = kb_taps + p.kb_h;
  const int n_local = (p.total_tiles > static_cast<int>(blockIdx.x))
                          ? (p.total_tiles - 1 - static_cast<int>(blockIdx.x)) / static_cast<int>(gridDim.x) + 1 : 0;
  const bool tracing = p.trace != nullptr && blockIdx.x == 0 && p.p2 > 0;
  // De-synchronise the CTAs: each walks the K blocks (and the passes) in its own rotated order, so that at any instant
  // the chip reads many different weight blocks instead of all SMs hitting the same few L2 slices.
  const int prot = (p.dbg & 8) ? 0 : static_cast<int>(blockIdx.x) % p.p1;
  const int krot1 = (p.dbg & 4) ? 0 : (static_cast<int>(blockIdx.x) / p.p1 * 5) % kb1;
  const int krot2 = (p.dbg & 4) ? 0 : static_cast<int>(blockIdx.x) % p.kb_c;

  if (warp == 0) {
    if (lane == 0) {
      // ------------------------------------------------------------ TMA producer
      int stage = 0;
      uint32_t phase = 0;
      LayerJobs jobs(n_local, p.p1, p.p2, !kZSmem);
      int kind, j, idx;
      while (jobs.next(kind, j, idx)) {
        const int tile = blockIdx.x + j * gridDim.x;
        const int t0 = (tile % p.tiles_t) * 128;
        const int b = tile / p.tiles_t;
        if (kind == 0) {
          const int row0 = p.w1_row0 + ((idx + prot) % p.p1) * 256;
          for (int ki = 0; ki < kb1; ++ki) {
            int kb = ki + krot1;
            if (kb >= kb1) kb -= kb1;
            mbar_wait(&empty[stage], phase ^ 1);
            if (p.dbg & 2) { mbar_arrive(&full[stage]); if (++stage == S) { stage = 0; phase ^= 1; } continue; }
            mbar_arrive_expect_tx(&full[stage], kLayerABytes + kLayerBBytes);
            uint8_t* sa = smem_a + stage * kLayerABytes;
            if (kb < kb_taps) {
              const int tap = kb / p.kb_c;
              const int cb = kb - tap * p.kb_c;
              tma_load_3d(&p.tmX, &full[stage], sa, cb * 64, t0 + (tap - 1) * p.dil, b);
            } else {
              tma_load_3d(&p.tmCond, &full[stage], sa, (kb - kb_taps) * 64, t0, b);
            }
            tma_load_2d(&p.tmW1, &full[stage], smem_b + stage * kLayerBBytes, kb * 64, row0);
            if (++stage == S) { stage = 0; phase ^= 1; }
          }
        } else {
          if (!kZSmem && idx == 0) mbar_wait(z_full, j & 1);  // tile j's z has been published to global memory
          const int row0 = p.w2_row0 + idx * BN2;
          for (int ki = 0; ki < p.kb_c; ++ki) {
            int kb = ki + krot2;
            if (kb >= p.kb_c) kb -= p.kb_c;
            mbar_wait(&empty[stage], phase ^ 1);
            if (p.dbg & 2) { mbar_arrive(&full[stage]); if (++stage == S) { stage = 0; phase ^= 1; } continue; }
            mbar_arrive_expect_tx(&full[stage], BN2 * 128 + (kZSmem ? 0u : kLayerABytes));
            if (!kZSmem) tma_load_3d(&p.tmZ, &full[stage], smem_a + stage * kLayerABytes, p.z_c0 + kb * 64, t0, b);
            tma_load_2d(&p.tmW2, &full[stage], smem_b + stage * kLayerBBytes, kb * 64, row0);
            if (++stage == S) { stage = 0; phase ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      // ------------------------------------------------------------ MMA issuer
      constexpr uint32_t idesc1 = make_idesc_bf16_f32(256);
      constexpr uint32_t idesc2 = make_idesc_bf16_f32(BN2);
      int stage = 0;
      uint32_t phase = 0;
      uint32_t acc_it = 0;
      int ntrace = 0;
      LayerJobs jobs(n_local, p.p1, p.p2, !kZSmem);
      int kind, j, idx;
      while (jobs.next(kind, j, idx)) {
        const uint32_t as = acc_it & 1;
        const unsigned long long ts0 = tracing ? globaltimer_ns() : 0ull;
        if (kZSmem && kind == 1 && idx == 0) {
          mbar_wait(z_full, j & 1);  // all epilogue threads have published tile j's z to shared memory
          tc_fence_after();
        }
        mbar_wait(&tmem_empty[as], ((acc_it >> 1) & 1) ^ 1);
        tc_fence_after();
        const unsigned long long ts1 = tracing ? globaltimer_ns() : 0ull;
        const uint32_t tmem_d = tmem_base + as * 256;
        const int nkb = kind == 0 ? kb1 : p.kb_c;
        for (int kb = 0; kb < nkb; ++kb) {
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          const uint8_t* a_src = (kZSmem && kind == 1) ? smem_z + ((kb + krot2) % p.kb_c) * kLayerABytes : smem_a + stage * kLayerABytes;
          const uint64_t da = make_kmajor_sw128_desc(smem_u32(a_src));
          const uint64_t db = make_kmajor_sw128_desc(smem_u32(smem_b + stage * kLayerBBytes));
          if (kind == 0) {
#pragma unroll
            for (int k = 0; k < 4; ++k) umma_bf16(tmem_d, da + 2 * k, db + 2 * k, idesc1, (kb | k) != 0);
          } else {
#pragma unroll
            for (int k = 0; k < 4; ++k) umma_bf16(tmem_d, da + 2 * k, db + 2 * k, idesc2, (kb | k) != 0);
          }
          umma_commit(&empty[stage]);
          if (kb == nkb - 1) umma_commit(&tmem_full[as]);
          if (++stage == S) { stage = 0; phase ^= 1; }
        }
        if (tracing && ntrace < p.trace_cap / 2) {
          unsigned long long* e = p.trace + 4 * ntrace++;
          e[0] = 16 + kind; e[1] = j * 16 + idx; e[2] = ts0; e[3] = ts1;
        }
        ++acc_it;
      }
    }
  } else if (warp >= 4) {
    // -------------------------------------------------------------- epilogue (256 threads, thread <-> frame;
    // the two warps of a lane quadrant interleave 32-column chunks)
    const int quad = warp & 3;
    const int half = (warp - 4) >> 2;
    const int r = quad * 32 + lane;  // row inside the tile == TMEM lane
    const uint32_t lane_addr = tmem_base + (static_cast<uint32_t>(quad * 32) << 16);
    uint32_t acc_it = 0;
    int ntrace = 0;
    LayerJobs jobs(n_local, p.p1, p.p2, !kZSmem);
    int kind, j, idx;
    while (jobs.next(kind, j, idx)) {
      const int tile = blockIdx.x + j * gridDim.x;
      const int b = tile / p.tiles_t;
      const int t = (tile % p.tiles_t) * 128 + r;
      const bool valid = t < p.T;
      const size_t row = static_cast<size_t>(b) * p.T + (valid ? t : 0);
      const uint32_t as = acc_it & 1;
      mbar_wait(&tmem_full[as], (acc_it >> 1) & 1);
      tc_fence_after();
      const unsigned long long ts0 = (tracing && threadIdx.x == 128) ? globaltimer_ns() : 0ull;
      const uint32_t taddr = lane_addr + as * 256;
      if (p.dbg & 1) {
        tc_fence_before();
        mbar_arrive(&tmem_empty[as]);
        if (kind == 0 && idx == p.p1 - 1 && p.p2 > 0) mbar_arrive(z_full);
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
          if (kZSmem && p.p2 > 0) {
            // K-major SWIZZLE_128B: block = 64 channels, row pitch 128 B, 16-B piece q -> q ^ (row & 7)
            uint8_t* blk = smem_z + (pidx * 2 + (c >> 1)) * kLayerABytes + r * 128;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const int q = (c & 1) * 4 + i;
              *reinterpret_cast<uint4*>(blk + ((q ^ (r & 7)) << 4)) = make_uint4(zp[4 * i], zp[4 * i + 1], zp[4 * i + 2], zp[4 * i + 3]);
            }
          }
          if (valid) {
#pragma unroll
            for (int i = 0; i < 4; ++i)
              reinterpret_cast<uint4*>(zg + c * 32)[i] = make_uint4(zp[4 * i], zp[4 * i + 1], zp[4 * i + 2], zp[4 * i + 3]);
          }
        }
        tc_fence_before();
        mbar_arrive(&tmem_empty[as]);
        if (idx == p.p1 - 1 && p.p2 > 0) {
          if (kZSmem) {
            fence_proxy_async_smem();  // generic-proxy z writes -> visible to the tensor core's async-proxy reads
          } else {
            __threadfence();             // z stores reach L2 ...
            fence_proxy_async_global();  // ... and are ordered before the TMA (async proxy) re-read
          }
          mbar_arrive(z_full);
        }
      } else {
        const int n0 = idx * BN2;
        const float* rv = p.rvec + static_cast<size_t>(b) * p.vec_bstride + n0;
        const __nv_bfloat16* xin = p.x_in + row * (p.kb_c * 64) + n0;
        __nv_bfloat16* xout = p.x_out + row * (p.kb_c * 64) + n0;
        uint4 xq[4];
        ld_x4(xin + half * 32, xq);
#pragma unroll 1
        for (int c = half; c < BN2 / 32; c += 2) {
          uint32_t a[32];
          tmem_ld_32x32(taddr + c * 32, a);
          float v[32], rr[32];
          ld_f32x32(rv + c * 32, rr);
          uint4 xn[4];
          if (c + 2 < BN2 / 32) ld_x4(xin + (c + 2) * 32, xn);
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
        tc_fence_before();
        mbar_arrive(&tmem_empty[as]);
      }
      if (tracing && threadIdx.x == 128 && ntrace < p.trace_cap / 2) {
        unsigned long long* e = p.trace + 4 * (p.trace_cap / 2 + ntrace++);
        e[0] = kind; e[1] = j * 16 + idx; e[2] = ts0; e[3] = globaltimer_ns();
      }
      ++acc_it;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc<512>(tmem_base);
  }
}

}  // namespace dsb
