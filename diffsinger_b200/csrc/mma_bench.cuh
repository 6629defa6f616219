// Micro-benchmark of the raw tcgen05.mma issue/execute rate (debug hook ds_debug_mma_rate): one thread issues
// `n_mma` UMMAs (M = 128 per CTA, N = 256, K = 16) on fixed shared-memory operands, committing to one of 8
// mbarriers every `commit_every` MMAs (never waited on) and once more at the end (waited on), single-CTA or
// CTA-pair.  Reports SM cycles for the whole train; nothing is loaded, so this is the tensor pipe alone.
#pragma once
#include "ptx.cuh"

namespace dsb {

template <bool kPair>
__global__ void __launch_bounds__(128, 1) mma_rate_kernel(int n_mma, int commit_every, long long* out) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + ((1024u - (raw_addr & 1023u)) & 1023u);
  uint8_t* sa = smem;          // 128 x 64 bf16
  uint8_t* sb = smem + 16384;  // 256 x 64 bf16 (pair: 128 x 64 used)
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 16384 + 32768);  // [0..7] ring, [8] done
  uint32_t* slot = reinterpret_cast<uint32_t*>(bar + 9);
  for (int i = threadIdx.x; i < (16384 + 32768) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = kPair ? cluster_ctarank() : 0;
  if (threadIdx.x == 0) {
    for (int i = 0; i < 9; ++i) mbar_init(&bar[i], 1);
    fence_barrier_init();
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  if (kPair) cluster_sync_all(); else __syncthreads();
  if (warp == 0) {
    if (kPair) tmem_alloc_pair<512>(slot); else tmem_alloc<512>(slot);
  }
  tc_fence_before();
  if (kPair) cluster_sync_all(); else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *slot;
  if (warp == 1 && lane == 0 && rank == 0) {
    const uint64_t da = make_kmajor_sw128_desc(smem_u32(sa));
    const uint64_t db = make_kmajor_sw128_desc(smem_u32(sb));
    constexpr uint32_t idesc = kPair ? make_idesc_bf16_f32_m(256, 256) : make_idesc_bf16_f32(256);
    // mode bits (passed in commit_every): 1 = commit after every group of 4 MMAs, 2 = (already satisfied) mbarrier
    // wait + tcgen05 fence before every group, 4 = rebuild the descriptors per group like the GEMM kernels do
    const int mode = commit_every;
    const int groups = n_mma / 4;
    const long long c0 = clock64();
    int stage = 0;
    for (int g = 0; g < groups; ++g) {
      if (mode & 2) {
        mbar_wait(&bar[stage], 1);  // fresh barrier: the "previous" phase counts as complete -> returns at once
        tc_fence_after();
      }
      uint64_t xa = da, xb = db;
      if (mode & 4) {
        xa = make_kmajor_sw128_desc(smem_u32(sa + (stage & 1) * 0));
        xb = make_kmajor_sw128_desc(smem_u32(sb + (stage & 1) * 0));
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        if (kPair) umma_bf16_pair(tmem_base, xa + 2 * k, xb + 2 * k, idesc, (g | k) != 0);
        else umma_bf16(tmem_base, xa + 2 * k, xb + 2 * k, idesc, (g | k) != 0);
      }
      if (mode & 1) {
        if (kPair) umma_commit_pair(&bar[4 + (stage & 3)]); else umma_commit(&bar[4 + (stage & 3)]);
      }
      if (++stage == 4) stage = 0;
    }
    const long long c1 = clock64();
    if (kPair) umma_commit_pair(&bar[8]); else umma_commit(&bar[8]);
    mbar_wait(&bar[8], 0);
    const long long c2 = clock64();
    out[2 * blockIdx.x] = c1 - c0;      // issue time
    out[2 * blockIdx.x + 1] = c2 - c0;  // issue + drain
  }
  tc_fence_before();
  if (kPair) cluster_sync_all(); else __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    if (kPair) tmem_dealloc_pair<512>(tmem_base); else tmem_dealloc<512>(tmem_base);
  }
}

}  // namespace dsb
