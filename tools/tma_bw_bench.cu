// Micro-benchmark: how fast can ONE SM (a few SMs) pull L2-resident data into shared memory?
//   mode 0: 2-D TMA tensor boxes (64 bf16 x rows, SWIZZLE_128B) as the weight stages of stack_kernel.cuh
//   mode 1: 1-D cp.async.bulk of `chunk` contiguous bytes
// with `depth` copies in flight (ring of mbarriers), `grid` CTAs (1 per SM), data set of `mb` MB read `reps` times.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/bin/tma_bw_bench tools/tma_bw_bench.cu
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include "../diffsinger_b200/csrc/ptx.cuh"
using namespace dsb;

__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(dst), "l"(reinterpret_cast<uint64_t>(src)), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

__global__ void __launch_bounds__(128, 1) bw_kernel(const __grid_constant__ CUtensorMap tm, const uint8_t* data, size_t bytes_per_cta,
                                                    int mode, int chunk, int rows, int depth, long long* out, int nthr) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  __shared__ uint64_t full[32];
  if (threadIdx.x == 0) {
    for (int i = 0; i < depth; ++i) mbar_init(&full[i], 1);
    fence_barrier_init();
  }
  __syncthreads();
  // `nthr` issuing threads (one per warp), each with its own slice of the ring and of the data
  const bool same_warp = nthr < 0;  // negative: the issuers are lanes 0..n-1 of warp 0 instead of lane 0 of n warps
  if (same_warp) nthr = -nthr;
  if (same_warp ? ((int)threadIdx.x < nthr) : ((threadIdx.x & 31) == 0 && (int)(threadIdx.x >> 5) < nthr)) {
    const int w = same_warp ? threadIdx.x : threadIdx.x >> 5;
    depth /= nthr;
    smem += (size_t)w * depth * chunk;
    uint64_t* full_w = full + w * depth;
    bytes_per_cta /= nthr;
    data += (size_t)w * bytes_per_cta;
    const int n = (int)(bytes_per_cta / chunk);
    const long long t0 = clock64();
    unsigned long long g0;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(g0));
    int issued = 0, done = 0;
    uint32_t ph[32] = {0};
    auto issue = [&](int i) {
      const int s = i % depth;
      mbar_arrive_expect_tx(&full_w[s], chunk);
      if (mode == 0) {
        const int row0 = (int)(((size_t)(blockIdx.x * nthr + w) * bytes_per_cta + (size_t)i * chunk) / 128);  // rows of 128 B in a [rows_total][64] bf16 matrix
        tma_load_2d(&tm, &full_w[s], smem + (size_t)s * chunk, 0, row0);
      } else {
        bulk_g2s(smem_u32(smem + (size_t)s * chunk), data + (size_t)blockIdx.x * bytes_per_cta * nthr + (size_t)i * chunk, chunk, &full_w[s]);
      }
    };
    for (; issued < depth && issued < n; ++issued) issue(issued);
    for (; done < n; ++done) {
      const int s = done % depth;
      mbar_wait(&full_w[s], ph[s]);
      ph[s] ^= 1;
      if (issued < n) issue(issued++);
    }
    unsigned long long g1;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(g1));
    if (w == 0) {
      out[blockIdx.x * 2] = clock64() - t0;
      out[blockIdx.x * 2 + 1] = (long long)(g1 - g0);
    }
  }
}

int main(int argc, char** argv) {
  const int grid = argc > 1 ? atoi(argv[1]) : 4;
  const size_t per_cta = (size_t)(argc > 2 ? atoi(argv[2]) : 4) << 20;  // MB per CTA
  typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                               const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
  EncodeFn enc = (EncodeFn)fn;
  uint8_t* data;
  const size_t total = per_cta * grid;
  cudaMalloc(&data, total);
  cudaMemset(data, 1, total);
  long long* out;
  cudaMalloc(&out, grid * 16);
  cudaFuncSetAttribute(bw_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  const int chunks[] = {8192, 16384, 32768, 49152, 65536};
  const int depths[] = {2, 4, 8};
  const int nthrs[] = {1, 2, 4, -2, -4};
  for (int mode = 0; mode < 2; ++mode)
    for (int chunk : chunks)
      for (int depth : depths)
       for (int nthr : nthrs) {
        const int an = nthr < 0 ? -nthr : nthr;
        if (depth < an * 2 && depth != an) continue;
        if (chunk < 16384 || chunk > 32768) continue;
        if ((size_t)chunk * depth > 190 * 1024) continue;
        const int rows = chunk / 128;
        if (mode == 0 && rows > 256) continue;
        CUtensorMap tm;
        cuuint64_t dims[2] = {64, total / 128};
        cuuint64_t strides[1] = {128};
        cuuint32_t box[2] = {64, (cuuint32_t)rows};
        cuuint32_t es[2] = {1, 1};
        enc(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, data, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
            CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        long long h[2 * 148];
        double best = 1e30, clk = 0;
        for (int rep = 0; rep < 4; ++rep) {  // first rep warms L2
          bw_kernel<<<grid, 128, (size_t)chunk * depth + 1024>>>(tm, data, per_cta, mode, chunk, rows, depth, out, nthr);
          cudaError_t e = cudaDeviceSynchronize();
          if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
          cudaMemcpy(h, out, grid * 16, cudaMemcpyDeviceToHost);
          double mx = 0, c = 0;
          for (int i = 0; i < grid; ++i) { if (h[2 * i + 1] > mx) { mx = (double)h[2 * i + 1]; c = (double)h[2 * i]; } }
          if (rep > 0 && mx < best) { best = mx; clk = c; }
        }
        printf("mode %d (%s) chunk %6d depth %2d issuers %d grid %3d: %7.1f GB/s per SM, %5.1f B/clk, clock %.2f GHz\n", mode, mode ? "1-D bulk" : "2-D box ",
               chunk, depth, nthr, grid, per_cta / best, per_cta / clk, clk / best);
      }
  return 0;
}
