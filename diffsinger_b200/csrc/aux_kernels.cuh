// SIMT helper kernels around the tensor-core GEMMs: weight packing (once per checkpoint load),
// step-embedding tables (once per sampler configuration), layout/precision conversion and the
// vectorised sampler update (HBM-bound elementwise work).
#pragma once
#include <cstdint>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include "ptx.cuh"

namespace dsb {

// ------------------------------------------------------------------ weight packing
// dst[r*dst_ld + k] (bf16) = src[r*src_rs + k*src_ks] for r < rows, k < cols; zero for the padding
// (r < dst_rows, k < dst_cols).  lo_off > 0 (fp32-accurate mode): also dst[.. + lo_off] = bf16(v - bf16(v)).
__global__ void pack_matrix_kernel(__nv_bfloat16* __restrict__ dst, int dst_rows, int dst_cols, long dst_ld,
                                   const float* __restrict__ src, int rows, int cols, long src_rs, long src_ks, long lo_off) {
  const long n = static_cast<long>(dst_rows) * dst_cols;
  for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < n; i += static_cast<long>(gridDim.x) * blockDim.x) {
    const int r = static_cast<int>(i / dst_cols), k = static_cast<int>(i % dst_cols);
    float v = 0.f;
    if (r < rows && k < cols) v = src[r * src_rs + k * src_ks];
    const __nv_bfloat16 hi = __float2bfloat16_rn(v);
    dst[r * dst_ld + k] = hi;
    if (lo_off > 0) dst[r * dst_ld + k + lo_off] = __float2bfloat16_rn(v - __bfloat162float(hi));
  }
}

__global__ void accum_kernel(float* __restrict__ dst, const float* __restrict__ src, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dst[i] += src[i];
}

// Gate GEMM weights of one layer (reference: ResidualBlock.dilated_conv [2C,C,3] + conditioner_projection
// [2C,H,1], wavenet.py:22-30).  Rows are permuted so each BN-row pass holds BN/2 gate rows followed by
// the BN/2 filter rows of the SAME channels; K = [tap0 | tap1 | tap2 | cond].
__global__ void pack_gate_kernel(__nv_bfloat16* __restrict__ dst, float* __restrict__ bias_dst,
                                 const float* __restrict__ w_dil, const float* __restrict__ b_dil,
                                 const float* __restrict__ w_cond, const float* __restrict__ b_cond, int C, int H, int BN,
                                 long ld, long lo_off) {
  const int K = 3 * C + H;
  const int half = BN / 2;
  const long n = static_cast<long>(2 * C) * K;
  for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < n; i += static_cast<long>(gridDim.x) * blockDim.x) {
    const int r = static_cast<int>(i / K), k = static_cast<int>(i % K);
    const int pass = r / BN, rr = r % BN;
    const int src_row = (rr < half) ? pass * half + rr : C + pass * half + (rr - half);
    float v;
    if (k < 3 * C) {
      const int tap = k / C, c = k % C;
      v = w_dil[(static_cast<long>(src_row) * C + c) * 3 + tap];
    } else {
      v = w_cond[static_cast<long>(src_row) * H + (k - 3 * C)];
    }
    const __nv_bfloat16 hi = __float2bfloat16_rn(v);
    dst[r * ld + k] = hi;
    if (lo_off > 0) dst[r * ld + k + lo_off] = __float2bfloat16_rn(v - __bfloat162float(hi));
    if (k == 0) bias_dst[r] = b_dil[src_row] + b_cond[src_row];
  }
}

// f16 copy of a packed gate bias (BN-row passes: BN/2 gate rows then BN/2 filter rows); gate entries are pre-multiplied by
// 0.5 for the sigmoid-by-tanh identity used in the packed-half gate epilogue.
__global__ void bias_half_kernel(uint16_t* __restrict__ dst, const float* __restrict__ src, int n, int BN) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    const float v = (i % BN < BN / 2) ? 0.5f * src[i] : src[i];
    const __half h = __float2half_rn(v);
    dst[i] = *reinterpret_cast<const uint16_t*>(&h);
  }
}

// ------------------------------------------------------------------ step embedding
// g[j][:] = mlp(SinusoidalPosEmb(t_j))   (common_layers.py:273-280, wavenet.py:58-62,89-90)
// one block per time value; blockDim = 256; dynamic smem = 5*C floats.
__global__ void step_mlp_kernel(float* __restrict__ g, const float* __restrict__ t, const float* __restrict__ w0,
                                const float* __restrict__ b0, const float* __restrict__ w2, const float* __restrict__ b2, int C) {
  extern __shared__ float sm[];
  float* e = sm;       // [C]
  float* h = sm + C;   // [4C]
  const int j = blockIdx.x;
  const float tj = t[j];
  const int half = C / 2;
  const float sc = logf(10000.0f) / static_cast<float>(half - 1);
  for (int i = threadIdx.x; i < half; i += blockDim.x) {
    const float a = tj * expf(static_cast<float>(i) * -sc);
    e[i] = sinf(a);
    e[half + i] = cosf(a);
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  for (int o = warp; o < 4 * C; o += nw) {
    float acc = 0.f;
    for (int k = lane; k < C; k += 32) acc = fmaf(w0[static_cast<long>(o) * C + k], e[k], acc);
    for (int s = 16; s > 0; s >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, s);
    if (lane == 0) {
      const float x = acc + b0[o];
      const float sp = (x > 20.f) ? x : log1pf(expf(x));  // softplus, torch threshold 20
      h[o] = x * tanhf(sp);                              // Mish
    }
  }
  __syncthreads();
  for (int o = warp; o < C; o += nw) {
    float acc = 0.f;
    for (int k = lane; k < 4 * C; k += 32) acc = fmaf(w2[static_cast<long>(o) * 4 * C + k], h[k], acc);
    for (int s = 16; s > 0; s >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, s);
    if (lane == 0) g[static_cast<long>(j) * C + o] = acc + b2[o];
  }
}

// d[j][l][:] = diffusion_projection_l(g[j])  (wavenet.py:29,34).  grid = (L, n); wd [L][C][C], bd [L][C].
__global__ void step_proj_kernel(float* __restrict__ d, const float* __restrict__ g, const float* __restrict__ wd,
                                 const float* __restrict__ bd, int C, int L) {
  extern __shared__ float sg[];
  const int l = blockIdx.x, j = blockIdx.y;
  for (int i = threadIdx.x; i < C; i += blockDim.x) sg[i] = g[static_cast<long>(j) * C + i];
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  const float* w = wd + static_cast<long>(l) * C * C;
  for (int o = warp; o < C; o += nw) {
    float acc = 0.f;
    for (int k = lane; k < C; k += 32) acc = fmaf(w[static_cast<long>(o) * C + k], sg[k], acc);
    for (int s = 16; s > 0; s >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, s);
    if (lane == 0) d[(static_cast<long>(j) * L + l) * C + o] = acc + bd[l * C + o];
  }
}

// r[j][l][:] = (b2_l - d[j][l]) / sqrt(2) + d[j][l+1]   for l < L-1: everything the residual epilogue adds to
// (x~ + R)/sqrt(2), folded once per time value (wavenet.py:36,44-48).
__global__ void res_vec_kernel(float* __restrict__ r, const float* __restrict__ d, const float* __restrict__ b2, int C, int L, int n) {
  const long total = static_cast<long>(n) * L * C;
  for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < total; i += static_cast<long>(gridDim.x) * blockDim.x) {
    const int o = static_cast<int>(i % C);
    const int l = static_cast<int>((i / C) % L);
    r[i] = (l + 1 < L) ? (b2[l * C + o] - d[i]) * 0.70710678118654752f + d[i + C] : 0.f;
  }
}

// ------------------------------------------------------------------ conversions
// fp32 [rows][cols] -> bf16 [rows][ld] (zero padded columns).  cols, ld multiples of 4.  Row pitch `pitch` >= ld;
// lo_off > 0 (fp32-accurate mode): the residual bf16(v - bf16(v)) goes lo_off elements further along the row.
__global__ void f32_to_bf16_pad_kernel(__nv_bfloat16* __restrict__ dst, const float* __restrict__ src, long rows, int cols, int ld,
                                       long pitch, int lo_off) {
  grid_dep_launch();
  grid_dep_wait();
  const int q = ld / 4;
  const long n = rows * q;
  for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < n; i += static_cast<long>(gridDim.x) * blockDim.x) {
    const long r = i / q;
    const int c = static_cast<int>(i % q) * 4;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (c < cols) v = *reinterpret_cast<const float4*>(src + r * cols + c);
    __nv_bfloat162 p0 = __floats2bfloat162_rn(v.x, v.y), p1 = __floats2bfloat162_rn(v.z, v.w);
    uint2 o;
    o.x = *reinterpret_cast<uint32_t*>(&p0);
    o.y = *reinterpret_cast<uint32_t*>(&p1);
    *reinterpret_cast<uint2*>(dst + r * pitch + c) = o;
    if (lo_off > 0) {
      const float2 h0 = __bfloat1622float2(p0), h1 = __bfloat1622float2(p1);
      p0 = __floats2bfloat162_rn(v.x - h0.x, v.y - h0.y);
      p1 = __floats2bfloat162_rn(v.z - h1.x, v.w - h1.y);
      o.x = *reinterpret_cast<uint32_t*>(&p0);
      o.y = *reinterpret_cast<uint32_t*>(&p1);
      *reinterpret_cast<uint2*>(dst + r * pitch + c + lo_off) = o;
    }
  }
}

// [B][R][Cc] <-> [B][Cc][R] fp32 transpose through a 32x33 tile (coalesced both ways).
__global__ void transpose_kernel(float* __restrict__ dst, const float* __restrict__ src, int R, int Cc) {
  __shared__ float tile[32][33];
  const int b = blockIdx.z;
  const float* s = src + static_cast<long>(b) * R * Cc;
  float* d = dst + static_cast<long>(b) * R * Cc;
  const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int r = r0 + i, c = c0 + threadIdx.x;
    if (r < R && c < Cc) tile[i][threadIdx.x] = s[static_cast<long>(r) * Cc + c];
  }
  __syncthreads();
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int c = c0 + i, r = r0 + threadIdx.x;
    if (r < R && c < Cc) d[static_cast<long>(c) * R + r] = tile[threadIdx.x][i];
  }
}

// Shallow-diffusion start state in ONE pass (ddpm.py:235-239,379-380 and reflow.py:112-118: norm_spec, then q_sample / the linear mix
// with the noise, then the transpose to the kernels' time-major layout):
//   x[b][t][m] = ca * ((src[b][t][m] - min[m]) / (max[m] - min[m]) * 2 - 1) + cb * noise[b][m][t]
// Every operation is rounded separately (no FMA contraction), in the reference's order, so the result equals the eager sequence bit
// for bit.  The noise arrives [B][M][T] and is transposed through a 32 x 33 tile.
__global__ void shallow_start_kernel(float* __restrict__ x, const float* __restrict__ src, const float* __restrict__ noise,
                                     const float* __restrict__ smin, const float* __restrict__ smax, float ca, float cb, int T, int M) {
  __shared__ float tile[32][33];
  const int b = blockIdx.z;
  const int m0 = blockIdx.x * 32, t0 = blockIdx.y * 32;
  const float* nz = noise + static_cast<long>(b) * M * T;
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {  // rows = mel bins, columns = frames (contiguous in the noise)
    const int m = m0 + i, t = t0 + threadIdx.x;
    if (m < M && t < T) tile[i][threadIdx.x] = nz[static_cast<long>(m) * T + t];
  }
  __syncthreads();
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {  // rows = frames, columns = mel bins (contiguous in src / x)
    const int t = t0 + i, m = m0 + threadIdx.x;
    if (t < T && m < M) {
      const long at = (static_cast<long>(b) * T + t) * M + m;
      const float lo = __ldg(smin + m), hi = __ldg(smax + m);
      const float nrm = __fsub_rn(__fmul_rn(__fdiv_rn(__fsub_rn(src[at], lo), __fsub_rn(hi, lo)), 2.0f), 1.0f);
      x[at] = __fadd_rn(__fmul_rn(ca, nrm), __fmul_rn(cb, tile[threadIdx.x][i]));
    }
  }
}

// ------------------------------------------------------------------ sampler update
// dst = sum_i coef[i] * src[i]  (n <= 8 terms), fp32, float4-vectorised, grid-stride.
// Covers every sampler family's tensor update (DDIM / PLMS / ancestral / DPM-Solver++ / UniPC /
// Euler / RK stages) because all of them are linear in (x, model outputs, history) with scalar
// coefficients that depend only on the schedule.  Optionally also writes a
// bf16 copy (row-padded) that the next backbone evaluation's head GEMM reads.
struct LincombArgs {
  const float* src[8];
  float coef[8];
  int n;
};
__global__ void lincomb_kernel(float* __restrict__ dst, LincombArgs a, long n4) {
  grid_dep_launch();
  grid_dep_wait();
  for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < n4; i += static_cast<long>(gridDim.x) * blockDim.x) {
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      if (k < a.n) {
        const float4 v = reinterpret_cast<const float4*>(a.src[k])[i];
        acc.x = fmaf(a.coef[k], v.x, acc.x);
        acc.y = fmaf(a.coef[k], v.y, acc.y);
        acc.z = fmaf(a.coef[k], v.z, acc.z);
        acc.w = fmaf(a.coef[k], v.w, acc.w);
      }
    }
    reinterpret_cast<float4*>(dst)[i] = acc;
  }
}

}  // namespace dsb

namespace dsb {

// ------------------------------------------------------------------ primitives for the other denoisers (LYNXNet, aux decoder)
// LayerNorm over the channel axis of time-major bf16 activations: y[r][:] = (x'[r][:] - mean) * rstd * gamma + beta with
// x'[r] = x[r] + add[r] + vec[r / rows_per_vec] (both optional: a second bf16 operand and an fp32 per-utterance vector -- the
// conditioner projection and the diffusion-step projection of lynxnet.py:82-89).
// One warp per row, the row lives in registers (C <= 2048, C % 8 == 0), two-pass statistics in fp32 (biased variance,
// like torch.nn.LayerNorm; lynxnet.py:52,123).
template <int kMaxChunks>
constexpr int kLnRows = kMaxChunks <= 2 ? 2 : 1;
// Persistent warps (the host sizes the grid to the resident warps): gamma / beta of the lane's columns are loaded once
// (kMaxChunks <= 2; re-read from L1 per row they cost 4x the row's own bytes in LSU wavefronts), and the raw 16-byte loads of
// the next row group are issued before the current one is reduced, so every warp keeps two row groups in flight.
template <int kMaxChunks>  // ceil(C / 256): a row lives in kMaxChunks x 8 registers per lane
__global__ void __launch_bounds__(256) layernorm_kernel(__nv_bfloat16* __restrict__ y, const __nv_bfloat16* __restrict__ x,
                                 const float* __restrict__ gamma, const float* __restrict__ beta, long rows, int C, float eps,
                                 const float* __restrict__ vec, long rows_per_vec, const __nv_bfloat16* __restrict__ add) {
  grid_dep_launch();
  constexpr int kRows = kLnRows<kMaxChunks>;  // rows per warp iteration: independent rows keep more bytes in flight
  constexpr bool kHoist = kMaxChunks <= 2;
  const int lane = threadIdx.x & 31;
  const long warp0 = (static_cast<long>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const long nwarps = (static_cast<long>(gridDim.x) * blockDim.x) >> 5;
  const int nch = (C + 255) / 256;
  const float inv_c = 1.f / static_cast<float>(C);
  float gg[kHoist ? kMaxChunks : 1][8], bb[kHoist ? kMaxChunks : 1][8];
  auto load_gb = [&](int i, float (&g8)[8], float (&b8)[8]) {
    const int e = (i * 32 + lane) * 8;
    const float4 g0 = __ldg(reinterpret_cast<const float4*>(gamma + e)), g1 = __ldg(reinterpret_cast<const float4*>(gamma + e) + 1);
    const float4 b0 = __ldg(reinterpret_cast<const float4*>(beta + e)), b1 = __ldg(reinterpret_cast<const float4*>(beta + e) + 1);
    g8[0] = g0.x; g8[1] = g0.y; g8[2] = g0.z; g8[3] = g0.w; g8[4] = g1.x; g8[5] = g1.y; g8[6] = g1.z; g8[7] = g1.w;
    b8[0] = b0.x; b8[1] = b0.y; b8[2] = b0.z; b8[3] = b0.w; b8[4] = b1.x; b8[5] = b1.y; b8[6] = b1.z; b8[7] = b1.w;
  };
  if (kHoist) {
#pragma unroll
    for (int i = 0; i < kMaxChunks; ++i)
      if (i < nch && (i * 32 + lane) * 8 < C) load_gb(i, gg[kHoist ? i : 0], bb[kHoist ? i : 0]);
  }
  grid_dep_wait();
  uint4 raw[kRows][kMaxChunks], rawa[kRows][kMaxChunks];
  auto fetch = [&](long r0) {
#pragma unroll
    for (int q = 0; q < kRows; ++q) {
      const long r = r0 + q < rows ? r0 + q : rows - 1;  // the tail warp recomputes the last row (same values rewritten)
#pragma unroll
      for (int i = 0; i < kMaxChunks; ++i) {
        const int e = (i * 32 + lane) * 8;
        if (i < nch && e < C) {
          raw[q][i] = *reinterpret_cast<const uint4*>(x + r * C + e);
          if (add) rawa[q][i] = *reinterpret_cast<const uint4*>(add + r * C + e);
        }
      }
    }
  };
  long r0 = warp0 * kRows;
  if (r0 < rows) fetch(r0);
  for (; r0 < rows; r0 += nwarps * kRows) {
    float v[kRows][kMaxChunks][8];
    float sum[kRows];
#pragma unroll
    for (int q = 0; q < kRows; ++q) {
      const long r = r0 + q < rows ? r0 + q : rows - 1;
      sum[q] = 0.f;
#pragma unroll
      for (int i = 0; i < kMaxChunks; ++i) {
        const int e = (i * 32 + lane) * 8;
        if (i < nch && e < C) {
          const uint32_t w[4] = {raw[q][i].x, raw[q][i].y, raw[q][i].z, raw[q][i].w};
#pragma unroll
          for (int h = 0; h < 4; ++h) {
            v[q][i][2 * h] = __uint_as_float(w[h] << 16);
            v[q][i][2 * h + 1] = __uint_as_float(w[h] & 0xffff0000u);
          }
          if (add) {  // second bf16 operand (LYNXNet: the step-invariant conditioner projection, hoisted out of the loop)
            const uint32_t wa[4] = {rawa[q][i].x, rawa[q][i].y, rawa[q][i].z, rawa[q][i].w};
#pragma unroll
            for (int h = 0; h < 4; ++h) {
              v[q][i][2 * h] += __uint_as_float(wa[h] << 16);
              v[q][i][2 * h + 1] += __uint_as_float(wa[h] & 0xffff0000u);
            }
          }
          if (vec) {
            const float* vr = vec + (r / rows_per_vec) * C + e;
#pragma unroll
            for (int h = 0; h < 8; ++h) v[q][i][h] += __ldg(vr + h);
          }
#pragma unroll
          for (int h = 0; h < 8; ++h) sum[q] += v[q][i][h];
        } else {
#pragma unroll
          for (int h = 0; h < 8; ++h) v[q][i][h] = 0.f;
        }
      }
    }
    if (r0 + nwarps * kRows < rows) fetch(r0 + nwarps * kRows);  // next row group in flight during the reductions below
#pragma unroll
    for (int q = 0; q < kRows; ++q)
      for (int s = 16; s > 0; s >>= 1) sum[q] += __shfl_xor_sync(0xffffffffu, sum[q], s);
    float mean[kRows], sq[kRows];
#pragma unroll
    for (int q = 0; q < kRows; ++q) {
      mean[q] = sum[q] * inv_c;
      sq[q] = 0.f;
#pragma unroll
      for (int i = 0; i < kMaxChunks; ++i) {
        const int e = (i * 32 + lane) * 8;
        if (i < nch && e < C) {
#pragma unroll
          for (int h = 0; h < 8; ++h) {
            v[q][i][h] -= mean[q];
            sq[q] = fmaf(v[q][i][h], v[q][i][h], sq[q]);
          }
        }
      }
    }
#pragma unroll
    for (int q = 0; q < kRows; ++q)
      for (int s = 16; s > 0; s >>= 1) sq[q] += __shfl_xor_sync(0xffffffffu, sq[q], s);
#pragma unroll
    for (int q = 0; q < kRows; ++q) {
      const long r = r0 + q < rows ? r0 + q : rows - 1;
      const float rstd = rsqrtf(sq[q] * inv_c + eps);
#pragma unroll
      for (int i = 0; i < kMaxChunks; ++i) {
        const int e = (i * 32 + lane) * 8;
        if (i < nch && e < C) {
          float g8[8], b8[8];
          if (!kHoist) load_gb(i, g8, b8);
          float o[8];
#pragma unroll
          for (int h = 0; h < 8; ++h)
            o[h] = fmaf(v[q][i][h] * rstd, kHoist ? gg[kHoist ? i : 0][h] : g8[h], kHoist ? bb[kHoist ? i : 0][h] : b8[h]);
          uint4 qo;
          qo.x = pack_bf16(o[0], o[1]); qo.y = pack_bf16(o[2], o[3]); qo.z = pack_bf16(o[4], o[5]); qo.w = pack_bf16(o[6], o[7]);
          *reinterpret_cast<uint4*>(y + r * C + e) = qo;
        }
      }
    }
  }
}

// The same LayerNorm for rows of up to 512 channels with the loads taken off the registers: every warp owns a ring of kLnSlots
// shared-memory slots, each filled by one elected lane with 1-D bulk copies (cp.async.bulk, a row is C * 2 contiguous bytes) of
// the next two rows (and their second operand) and guarded by an mbarrier -- three row groups per warp are in flight at no
// register cost, where the register-prefetching kernel above stalls on its loads (long scoreboard 5 per issue,
// profiles/r01_layernorm_kernel.txt).  The warp that consumes a slot refills it itself (after __syncwarp: all lanes have their
// values in registers), so no empty barrier is needed.  Statistics and normalisation are unchanged.
constexpr int kLnSlots = 3;
__device__ __forceinline__ void bulk_load_1d(uint32_t dst_smem, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(dst_smem), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
// dynamic smem = warps * kLnSlots * 2 rows * (kAdd ? 2 : 1) * 1024 B; mbarriers in static shared memory
template <bool kAdd>
__global__ void __launch_bounds__(256) layernorm_bulk_kernel(__nv_bfloat16* __restrict__ y, const __nv_bfloat16* __restrict__ x,
                                 const float* __restrict__ gamma, const float* __restrict__ beta, long rows, int C, float eps,
                                 const float* __restrict__ vec, long rows_per_vec, const __nv_bfloat16* __restrict__ add) {
  constexpr int kRows = 2, kOps = kAdd ? 2 : 1;
  constexpr uint32_t kSlotBytes = kRows * kOps * 1024;
  extern __shared__ __align__(128) uint8_t ln_smem[];
  __shared__ uint64_t ln_bars[8 * kLnSlots];
  grid_dep_launch();
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const long warp0 = (static_cast<long>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const long nwarps = (static_cast<long>(gridDim.x) * blockDim.x) >> 5;
  const int nch = (C + 255) / 256;
  const float inv_c = 1.f / static_cast<float>(C);
  const uint32_t row_bytes = static_cast<uint32_t>(C) * 2;
  uint64_t* bars = ln_bars + wib * kLnSlots;
  const uint32_t ring = smem_u32(ln_smem) + wib * kLnSlots * kSlotBytes;
  if (lane == 0) {
    for (int i = 0; i < kLnSlots; ++i) mbar_init(&bars[i], 1);
    fence_barrier_init();
  }
  float gg[2][8], bb[2][8];
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const int e = (i * 32 + lane) * 8;
    if (i < nch && e < C) {
      const float4 g0 = __ldg(reinterpret_cast<const float4*>(gamma + e)), g1 = __ldg(reinterpret_cast<const float4*>(gamma + e) + 1);
      const float4 b0 = __ldg(reinterpret_cast<const float4*>(beta + e)), b1 = __ldg(reinterpret_cast<const float4*>(beta + e) + 1);
      gg[i][0] = g0.x; gg[i][1] = g0.y; gg[i][2] = g0.z; gg[i][3] = g0.w; gg[i][4] = g1.x; gg[i][5] = g1.y; gg[i][6] = g1.z; gg[i][7] = g1.w;
      bb[i][0] = b0.x; bb[i][1] = b0.y; bb[i][2] = b0.z; bb[i][3] = b0.w; bb[i][4] = b1.x; bb[i][5] = b1.y; bb[i][6] = b1.z; bb[i][7] = b1.w;
    }
  }
  __syncwarp();
  grid_dep_wait();
  const long stride = nwarps * kRows;
  auto issue = [&](long r0, int slot) {  // lane 0 only
    if (r0 >= rows) return;
    mbar_arrive_expect_tx(&bars[slot], kRows * kOps * row_bytes);
#pragma unroll
    for (int q = 0; q < kRows; ++q) {
      const long r = r0 + q < rows ? r0 + q : rows - 1;  // the tail warp recomputes the last row (same values rewritten)
      bulk_load_1d(ring + slot * kSlotBytes + q * kOps * 1024, x + r * C, row_bytes, &bars[slot]);
      if (kAdd) bulk_load_1d(ring + slot * kSlotBytes + q * kOps * 1024 + 1024, add + r * C, row_bytes, &bars[slot]);
    }
  };
  long r0 = warp0 * kRows;
  if (lane == 0)
    for (int i = 0; i < kLnSlots; ++i) issue(r0 + i * stride, i);
  int slot = 0;
  uint32_t phase = 0;
  for (; r0 < rows; r0 += stride) {
    mbar_wait(&bars[slot], phase);
    float v[kRows][2][8];
    float sum[kRows];
    const uint8_t* sl = ln_smem + (wib * kLnSlots + slot) * kSlotBytes;
#pragma unroll
    for (int q = 0; q < kRows; ++q) {
      const long r = r0 + q < rows ? r0 + q : rows - 1;
      sum[q] = 0.f;
#pragma unroll
      for (int i = 0; i < 2; ++i) {
        const int e = (i * 32 + lane) * 8;
        if (i < nch && e < C) {
          const uint4 qv = *reinterpret_cast<const uint4*>(sl + q * kOps * 1024 + e * 2);
          const uint32_t w[4] = {qv.x, qv.y, qv.z, qv.w};
#pragma unroll
          for (int h = 0; h < 4; ++h) {
            v[q][i][2 * h] = __uint_as_float(w[h] << 16);
            v[q][i][2 * h + 1] = __uint_as_float(w[h] & 0xffff0000u);
          }
          if (kAdd) {
            const uint4 qa = *reinterpret_cast<const uint4*>(sl + q * kOps * 1024 + 1024 + e * 2);
            const uint32_t wa[4] = {qa.x, qa.y, qa.z, qa.w};
#pragma unroll
            for (int h = 0; h < 4; ++h) {
              v[q][i][2 * h] += __uint_as_float(wa[h] << 16);
              v[q][i][2 * h + 1] += __uint_as_float(wa[h] & 0xffff0000u);
            }
          }
          if (vec) {
            const float* vr = vec + (r / rows_per_vec) * C + e;
#pragma unroll
            for (int h = 0; h < 8; ++h) v[q][i][h] += __ldg(vr + h);
          }
#pragma unroll
          for (int h = 0; h < 8; ++h) sum[q] += v[q][i][h];
        } else {
#pragma unroll
          for (int h = 0; h < 8; ++h) v[q][i][h] = 0.f;
        }
      }
    }
#pragma unroll
    for (int q = 0; q < kRows; ++q)  // the butterfly also orders every lane's slot reads before the refill below
      for (int s = 16; s > 0; s >>= 1) sum[q] += __shfl_xor_sync(0xffffffffu, sum[q], s);
    if (lane == 0) {
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy reads of the slot before its async-proxy refill
      issue(r0 + kLnSlots * stride, slot);
    }
    float mean[kRows], sq[kRows];
#pragma unroll
    for (int q = 0; q < kRows; ++q) {
      mean[q] = sum[q] * inv_c;
      sq[q] = 0.f;
#pragma unroll
      for (int i = 0; i < 2; ++i) {
        const int e = (i * 32 + lane) * 8;
        if (i < nch && e < C) {
#pragma unroll
          for (int h = 0; h < 8; ++h) {
            v[q][i][h] -= mean[q];
            sq[q] = fmaf(v[q][i][h], v[q][i][h], sq[q]);
          }
        }
      }
    }
#pragma unroll
    for (int q = 0; q < kRows; ++q)
      for (int s = 16; s > 0; s >>= 1) sq[q] += __shfl_xor_sync(0xffffffffu, sq[q], s);
#pragma unroll
    for (int q = 0; q < kRows; ++q) {
      const long r = r0 + q < rows ? r0 + q : rows - 1;
      const float rstd = rsqrtf(sq[q] * inv_c + eps);
#pragma unroll
      for (int i = 0; i < 2; ++i) {
        const int e = (i * 32 + lane) * 8;
        if (i < nch && e < C) {
          float o[8];
#pragma unroll
          for (int h = 0; h < 8; ++h) o[h] = fmaf(v[q][i][h] * rstd, gg[i][h], bb[i][h]);
          uint4 qo;
          qo.x = pack_bf16(o[0], o[1]); qo.y = pack_bf16(o[2], o[3]); qo.z = pack_bf16(o[4], o[5]); qo.w = pack_bf16(o[6], o[7]);
          *reinterpret_cast<uint4*>(y + r * C + e) = qo;
        }
      }
    }
    if (++slot == kLnSlots) { slot = 0; phase ^= 1; }
  }
}

// Depthwise Conv1d along time on time-major bf16 activations [B][T][CH] with zero padding, fused bias and activation
// (0 none, 1 ReLU, 2 PReLU with per-channel slope, 3 SiLU): y[b][t][c] = act(bias[c] + sum_j w[c][j] * x[b][t + j - pad][c]).
// One thread owns a channel PAIR and kFrames = K * kGroups consecutive frames.  All kFrames + K - 1 input rows of its pair
// are fetched up front with 4-byte cp.async (zero fill outside [0, T)) into the thread's own column of a shared-memory
// tile -- one exposed memory latency per block instead of one per output -- and then walked with a K-deep register
// window: per output one LDS and K packed fp32 FMAs (fma.rn.f32x2, both channels per instruction, two chains).
// grid = (CH / 2 / 128, ceil(T / kFrames), B), block = 128, dynamic smem = (kFrames + K - 1) * 512 B.   lynxnet.py:59
__device__ __forceinline__ void cp_async_4_zfill(uint32_t smem_addr, const void* gptr, bool valid) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(smem_addr), "l"(gptr), "r"(valid ? 4 : 0) : "memory");
}
template <int K, int kGroups>
__global__ void __launch_bounds__(128) dwconv_kernel(__nv_bfloat16* __restrict__ y, const __nv_bfloat16* __restrict__ x,
                                                     const float* __restrict__ w, const float* __restrict__ bias,
                                                     const float* __restrict__ slope, int T, int CH, int pad, int act) {
  extern __shared__ uint32_t dw_tile[];  // [kRows][128 channel pairs]
  grid_dep_launch();
  grid_dep_wait();
  constexpr int kFrames = K * kGroups;
  constexpr int kRows = kFrames + K - 1;
  const int cp = blockIdx.x * 128 + threadIdx.x;  // channel pair
  if (2 * cp >= CH) return;
  const int b = blockIdx.z;
  const int t0 = blockIdx.y * kFrames;
  const __nv_bfloat16* xb = x + static_cast<size_t>(b) * T * CH + 2 * cp;
  __nv_bfloat16* yb = y + static_cast<size_t>(b) * T * CH + 2 * cp;
  const uint32_t col = smem_u32(dw_tile + threadIdx.x);
#pragma unroll 4
  for (int r = 0; r < kRows; ++r) {
    const int t = t0 - pad + r;
    const bool valid = t >= 0 && t < T;
    cp_async_4_zfill(col + r * 512, xb + static_cast<size_t>(valid ? t : 0) * CH, valid);
  }
  unsigned long long wv[K];
#pragma unroll
  for (int j = 0; j < K; ++j)
    wv[j] = pack_f32x2(__ldg(w + static_cast<size_t>(2 * cp) * K + j), __ldg(w + static_cast<size_t>(2 * cp + 1) * K + j));
  const unsigned long long bv = pack_f32x2(bias ? __ldg(bias + 2 * cp) : 0.f, bias ? __ldg(bias + 2 * cp + 1) : 0.f);
  const float s0 = slope ? __ldg(slope + 2 * cp) : 0.f, s1 = slope ? __ldg(slope + 2 * cp + 1) : 0.f;
  asm volatile("cp.async.wait_all;" ::: "memory");  // each thread reads back only the column it filled itself
  auto row = [&](int r) -> unsigned long long {
    const uint32_t q = dw_tile[r * 128 + threadIdx.x];
    return pack_f32x2(__uint_as_float(q << 16), __uint_as_float(q & 0xffff0000u));
  };
  unsigned long long win[K];  // slot i % K holds input row (t0 - pad + i)
#pragma unroll
  for (int i = 0; i < K - 1; ++i) win[i] = row(i);
#pragma unroll 1
  for (int g = 0; g < kGroups; ++g) {
    if (t0 + g * K >= T) break;
#pragma unroll
    for (int s = 0; s < K; ++s) {
      const int o = g * K + s;  // output frame t0 + o needs rows o .. o + K - 1; row o + K - 1 goes to slot (s + K - 1) % K
      win[(s + K - 1) % K] = row(o + K - 1);
      unsigned long long acc[4] = {bv, 0ull, 0ull, 0ull};  // four independent chains over the taps
#pragma unroll
      for (int j = 0; j < K; ++j) acc[j & 3] = fma_f32x2(wv[j], win[(s + j) % K], acc[j & 3]);
      float a0 = 0.f, a1 = 0.f;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        float p0, p1;
        unpack_f32x2(acc[c], p0, p1);
        a0 += p0;
        a1 += p1;
      }
      if (act == 1) { a0 = fmaxf(a0, 0.f); a1 = fmaxf(a1, 0.f); }
      else if (act == 2) { a0 = a0 > 0.f ? a0 : s0 * a0; a1 = a1 > 0.f ? a1 : s1 * a1; }
      else if (act == 3) { a0 = a0 / (1.f + expf(-a0)); a1 = a1 / (1.f + expf(-a1)); }
      if (t0 + o < T) *reinterpret_cast<uint32_t*>(yb + static_cast<size_t>(t0 + o) * CH) = pack_bf16(a0, a1);
    }
  }
}

}  // namespace dsb

// ------------------------------------------------------------------------------------ training: gated activation, fwd / bwd
// The training path (ddpm.py:212-219 under autograd) keeps the pre-activations z = [gate | filter]
// (wavenet.py:40, plain channel order, bf16) of every layer instead of fusing the gate into the GEMM epilogue, so the
// backward pass needs no recomputation.  Both kernels are HBM-bound streaming passes: 8 channels (16 B) per thread per operand.
namespace dsb {

__device__ __forceinline__ void unpack_bf16x8(const uint4& q, float (&v)[8]) {
  const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
  for (int h = 0; h < 4; ++h) {
    v[2 * h] = __uint_as_float(w[h] << 16);
    v[2 * h + 1] = __uint_as_float(w[h] & 0xffff0000u);
  }
}
__device__ __forceinline__ uint4 pack_bf16x8(const float (&v)[8]) {
  uint4 q;
  uint32_t* w = reinterpret_cast<uint32_t*>(&q);
#pragma unroll
  for (int h = 0; h < 4; ++h) {
    const __nv_bfloat162 b = __floats2bfloat162_rn(v[2 * h], v[2 * h + 1]);
    w[h] = *reinterpret_cast<const uint32_t*>(&b);
  }
  return q;
}

// y[r][c] = sigmoid(z[r][c]) * tanh(z[r][C + c])                                        wavenet.py:40-42
__global__ void __launch_bounds__(256) gate_fwd_kernel(__nv_bfloat16* __restrict__ y, long ld_y, const __nv_bfloat16* __restrict__ z,
                                                       long ld_z, long rows, int C) {
  const int cg = C / 8;
  const long n = rows * cg;
  for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
    const long r = i / cg;
    const int c = (int)(i - r * cg) * 8;
    float g[8], f[8], o[8];
    unpack_bf16x8(__ldg(reinterpret_cast<const uint4*>(z + r * ld_z + c)), g);
    unpack_bf16x8(__ldg(reinterpret_cast<const uint4*>(z + r * ld_z + C + c)), f);
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = tanhf(f[j]) / (1.0f + __expf(-g[j]));
    *reinterpret_cast<uint4*>(y + r * ld_y + c) = pack_bf16x8(o);
  }
}

// dz_gate = dy * tanh(f) * s (1 - s),  dz_filter = dy * s * (1 - tanh(f)^2),  s = sigmoid(g)
__global__ void __launch_bounds__(256) gate_bwd_kernel(__nv_bfloat16* __restrict__ dz, long ld_dz, const __nv_bfloat16* __restrict__ dy,
                                                       long ld_dy, const __nv_bfloat16* __restrict__ z, long ld_z, long rows, int C) {
  const int cg = C / 8;
  const long n = rows * cg;
  for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
    const long r = i / cg;
    const int c = (int)(i - r * cg) * 8;
    float g[8], f[8], d[8], og[8], of[8];
    unpack_bf16x8(__ldg(reinterpret_cast<const uint4*>(z + r * ld_z + c)), g);
    unpack_bf16x8(__ldg(reinterpret_cast<const uint4*>(z + r * ld_z + C + c)), f);
    unpack_bf16x8(__ldg(reinterpret_cast<const uint4*>(dy + r * ld_dy + c)), d);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float s = 1.0f / (1.0f + __expf(-g[j]));
      const float th = tanhf(f[j]);
      og[j] = d[j] * th * s * (1.0f - s);
      of[j] = d[j] * s * (1.0f - th * th);
    }
    *reinterpret_cast<uint4*>(dz + r * ld_dz + c) = pack_bf16x8(og);
    *reinterpret_cast<uint4*>(dz + r * ld_dz + C + c) = pack_bf16x8(of);
  }
}

}  // namespace dsb

// ------------------------------------------------------------------------------------ NSF-HiFiGAN helpers
namespace dsb {

// Conv1d(1 -> channels, k, stride, pad) over the single-channel harmonic source (nsf_hifigan/models.py:237-241, 281):
// y[b][n][c] = bias[c] + sum_k w[k][c] * src[b][n * stride + k - pad], w tap-major [k][ld] with zero columns >= channels.
// One thread per (b, n, 8 channels): a warp covers consecutive channel groups of one output sample, so the source sample
// is a broadcast load and the weight row a coalesced one.
__global__ void __launch_bounds__(256) source_conv_kernel(__nv_bfloat16* __restrict__ y, const float* __restrict__ src,
                                                          const float* __restrict__ w, const float* __restrict__ bias, int B, long n_src,
                                                          long n_out, int ld, int K, int stride, int pad) {
  const int cg = ld / 8;
  const long total = (long)B * n_out * cg;
  for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int g = (int)(i % cg);
    const long bn = i / cg;
    const long n = bn % n_out;
    const long b = bn / n_out;
    const int c0 = g * 8;
    float acc[8];
    const float4 b0 = __ldg(reinterpret_cast<const float4*>(bias + c0)), b1 = __ldg(reinterpret_cast<const float4*>(bias + c0) + 1);
    acc[0] = b0.x; acc[1] = b0.y; acc[2] = b0.z; acc[3] = b0.w; acc[4] = b1.x; acc[5] = b1.y; acc[6] = b1.z; acc[7] = b1.w;
    const float* sb = src + b * n_src;
    const long s0 = n * stride - pad;
    const int k_lo = s0 < 0 ? (int)(-s0) : 0;
    const int k_hi = (s0 + K > n_src) ? (int)(n_src - s0) : K;
#pragma unroll 4
    for (int k = k_lo; k < k_hi; ++k) {
      const float sv = __ldg(sb + s0 + k);
      const float4 w0 = __ldg(reinterpret_cast<const float4*>(w + (long)k * ld + c0));
      const float4 w1 = __ldg(reinterpret_cast<const float4*>(w + (long)k * ld + c0) + 1);
      acc[0] = fmaf(w0.x, sv, acc[0]); acc[1] = fmaf(w0.y, sv, acc[1]); acc[2] = fmaf(w0.z, sv, acc[2]); acc[3] = fmaf(w0.w, sv, acc[3]);
      acc[4] = fmaf(w1.x, sv, acc[4]); acc[5] = fmaf(w1.y, sv, acc[5]); acc[6] = fmaf(w1.z, sv, acc[6]); acc[7] = fmaf(w1.w, sv, acc[7]);
    }
    *reinterpret_cast<uint4*>(y + bn * ld + c0) = pack_bf16x8(acc);
  }
}

// y = LeakyReLU((a + b + c) / 3, slope): the ResBlock average and the next stage's activation (models.py:286-294).
__global__ void __launch_bounds__(256) mean3_lrelu_kernel(__nv_bfloat16* __restrict__ y, const __nv_bfloat16* __restrict__ a,
                                                          const __nv_bfloat16* __restrict__ b, const __nv_bfloat16* __restrict__ c,
                                                          long n8, float slope) {
  for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < n8; i += (long)gridDim.x * blockDim.x) {
    float va[8], vb[8], vc[8], o[8];
    unpack_bf16x8(__ldg(reinterpret_cast<const uint4*>(a) + i), va);
    unpack_bf16x8(__ldg(reinterpret_cast<const uint4*>(b) + i), vb);
    unpack_bf16x8(__ldg(reinterpret_cast<const uint4*>(c) + i), vc);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float m = (va[j] + vb[j] + vc[j]) * (1.0f / 3.0f);
      o[j] = m > 0.0f ? m : m * slope;
    }
    reinterpret_cast<uint4*>(y)[i] = pack_bf16x8(o);
  }
}

}  // namespace dsb
