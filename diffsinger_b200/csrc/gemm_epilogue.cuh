// Fused epilogues of ds_gemm_kernel: TMEM -> registers -> math -> global memory.
// One thread owns one frame (TMEM lane); columns are output channels, so every thread writes
// contiguous channel runs of its own [B][T][C] row (64-byte runs per 32-column chunk).
#pragma once
#include "gemm_kernel.cuh"

namespace dsb {

__device__ __forceinline__ void ld_f32x32(const float* __restrict__ ptr, float (&v)[32]) {
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const float4 q = __ldg(reinterpret_cast<const float4*>(ptr) + i);
    v[4 * i + 0] = q.x; v[4 * i + 1] = q.y; v[4 * i + 2] = q.z; v[4 * i + 3] = q.w;
  }
}

__device__ __forceinline__ void st_bf16x32(__nv_bfloat16* ptr, const float (&v)[32]) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    uint4 q;
    q.x = pack_bf16(v[8 * i + 0], v[8 * i + 1]);
    q.y = pack_bf16(v[8 * i + 2], v[8 * i + 3]);
    q.z = pack_bf16(v[8 * i + 4], v[8 * i + 5]);
    q.w = pack_bf16(v[8 * i + 6], v[8 * i + 7]);
    reinterpret_cast<uint4*>(ptr)[i] = q;
  }
}

// fp32-accurate mode: v = hi + lo with hi = bf16(v), lo = bf16(v - hi); hi at ptr, lo at ptr + lo_off.
__device__ __forceinline__ void st_split_bf16x32(__nv_bfloat16* ptr, int lo_off, const float (&v)[32]) {
  float r[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) r[j] = v[j] - __bfloat162float(__float2bfloat16_rn(v[j]));
  st_bf16x32(ptr, v);
  st_bf16x32(ptr + lo_off, r);
}
template <bool kSplit>
__device__ __forceinline__ void st_act32(__nv_bfloat16* ptr, int lo_off, const float (&v)[32]) {
  if constexpr (kSplit) st_split_bf16x32(ptr, lo_off, v);
  else st_bf16x32(ptr, v);
}
// 32 consecutive output channels of this lane's frame, starting at channel `ch` of p.out: through the warp's staging tile and
// one TMA store when the kernel provides one (all lanes call this; rows beyond T are clipped by the map), else per-lane stores.
template <bool kSplit>
__device__ __forceinline__ void epi_store32(const GemmParams& p, const EpiStage& es, size_t row, int ch, int b, bool valid,
                                            const float (&v)[32]) {
  if constexpr (!kSplit) {
    if (es.on) {
      uint32_t o[16];
#pragma unroll
      for (int i = 0; i < 16; ++i) o[i] = pack_bf16(v[2 * i], v[2 * i + 1]);
      warp_store_tile(&p.tmOut, es.sw, es.lane, o, ch, es.tw0, b, es.tw0 < p.T);
      return;
    }
  }
  if (valid) st_act32<kSplit>(reinterpret_cast<__nv_bfloat16*>(p.out) + row * p.ld_out + ch, p.out_lo, v);
}
__device__ __forceinline__ float sigmoid_accurate(float x) { return 1.0f / (1.0f + expf(-x)); }

__device__ __forceinline__ void ld_x4(const __nv_bfloat16* ptr, uint4 (&q)[4]) {
#pragma unroll
  for (int i = 0; i < 4; ++i) q[i] = __ldg(reinterpret_cast<const uint4*>(ptr) + i);
}

// partial sums of a folded LayerNorm's input (GemmParams::stats_out): r = the bf16 value that is stored, d = the step vector chunk
__device__ __forceinline__ void ln_stats_acc(const float (&v)[32], const float* d, float& s1, float& s2, float& s3) {
  float dv[32];
  if (d != nullptr) ld_f32x32(d, dv);
#pragma unroll
  for (int j = 0; j < 32; ++j) {
    const float r = __bfloat162float(__float2bfloat16_rn(v[j]));
    s1 += r;
    s2 = fmaf(r, r, s2);
    if (d != nullptr) s3 = fmaf(r, dv[j], s3);
  }
}

// `half` in {0,1}: the two epilogue warps that share a TMEM lane quadrant interleave the 32-column chunks.
template <int BN, int EPI, bool kSplit>
__device__ __forceinline__ void epilogue_tile(const GemmParams& p, uint32_t taddr, int b, int t, bool valid, int nt, int half,
                                              const EpiStage& es) {
  const size_t row = static_cast<size_t>(b) * p.T + (valid ? t : 0);
  const bool live = p.len == nullptr || t < __ldg(p.len + b);  // ragged batch: frames beyond the utterance's end are stored as zeros
  if constexpr (EPI == EPI_GATE) {
    // columns [0, BN/2) = gate pre-activations, [BN/2, BN) = filter pre-activations of the SAME channels
    constexpr int H = BN / 2;
    const float* bias = p.bias + nt * BN;
#pragma unroll 1
    for (int c = half; c < H / 32; c += 2) {
      uint32_t g[32], f[32];
      tmem_ld_32x32(taddr + c * 32, g);
      tmem_ld_32x32(taddr + H + c * 32, f);
      float bg[32], bf[32], z[32];
      ld_f32x32(bias + c * 32, bg);
      ld_f32x32(bias + H + c * 32, bf);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        const float gj = __uint_as_float(g[j]) + bg[j];
        const float fj = __uint_as_float(f[j]) + bf[j];
        if constexpr (kSplit) {
          z[j] = sigmoid_accurate(gj) * tanhf(fj);
        } else {
          const float sg = fmaf(0.5f, tanh_approx(0.5f * gj), 0.5f);  // sigmoid
          z[j] = sg * tanh_approx(fj);
        }
      }
      epi_store32<kSplit>(p, es, row, p.out_c0 + nt * H + c * 32, b, valid, z);
    }
  } else if constexpr (EPI == EPI_SWIGLU) {
    // columns [0, BN/2) = "out" half, [BN/2, BN) = "gate" half of the SAME channels (pack-time row permutation):
    // z = out * silu(gate)   (SwiGLU, modules/commons/common_layers.py; LYNXConvModule, lynxnet.py:57)
    constexpr int H = BN / 2;
    const float* bias = p.bias + nt * BN;
    // folded LayerNorm (GemmParams::ln_*): this row's mean / rstd from the producer's partial sums
    float mu = 0.0f, rstd = 1.0f;
    const float* ls = nullptr;
    const float* le = nullptr;
    if (p.ln_stats != nullptr) {
      const float* erow = p.ln_e + static_cast<size_t>(b) * p.ln_e_bstride;
      float s1 = __ldg(erow + p.ln_n), s2 = __ldg(erow + p.ln_n + 1);  // sum d, sum d^2
      const float4* st = p.ln_stats + row * p.ln_parts;
      for (int k = 0; k < p.ln_parts; ++k) {
        const float4 q = __ldg(st + k);
        s1 += q.x;
        s2 += q.y + 2.0f * q.z;
      }
      const float inv = 1.0f / static_cast<float>(p.ln_channels);
      mu = s1 * inv;
      rstd = rsqrtf(fmaxf(s2 * inv - mu * mu, 0.0f) + p.ln_eps);
      ls = p.ln_s + nt * BN;
      le = erow + nt * BN;
    }
#pragma unroll 1
    for (int c = half; c < H / 32; c += 2) {
      uint32_t a[32], g[32];
      tmem_ld_32x32(taddr + c * 32, a);
      tmem_ld_32x32(taddr + H + c * 32, g);
      float ba[32], bgt[32], z[32];
      ld_f32x32(bias + c * 32, ba);
      ld_f32x32(bias + H + c * 32, bgt);
      tmem_ld_wait();
      if (ls != nullptr) {  // acc <- rstd * (acc + e - mean * s)
        float sv[32], ev[32];
        ld_f32x32(ls + c * 32, sv);
        ld_f32x32(le + c * 32, ev);
#pragma unroll
        for (int j = 0; j < 32; ++j) a[j] = __float_as_uint(rstd * (__uint_as_float(a[j]) + ev[j] - mu * sv[j]));
        ld_f32x32(ls + H + c * 32, sv);
        ld_f32x32(le + H + c * 32, ev);
#pragma unroll
        for (int j = 0; j < 32; ++j) g[j] = __float_as_uint(rstd * (__uint_as_float(g[j]) + ev[j] - mu * sv[j]));
      }
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        const float aj = __uint_as_float(a[j]) + ba[j];
        const float gj = __uint_as_float(g[j]) + bgt[j];
        if constexpr (kSplit) z[j] = aj * gj * sigmoid_accurate(gj);
        else z[j] = aj * gj * fmaf(0.5f, tanh_approx(0.5f * gj), 0.5f);  // MUFU path: the tile's MMAs take only 4 k cycles
      }
      if (!live) {  // ragged batch: the depthwise conv that follows reads zeros beyond the utterance's end
#pragma unroll
        for (int j = 0; j < 32; ++j) z[j] = 0.0f;
      }
      epi_store32<kSplit>(p, es, row, p.out_c0 + nt * H + c * 32, b, valid, z);
    }
  } else if constexpr (EPI == EPI_GATEBWD) {
    // training backward of the gated activation, fused behind dy = W2^T [h | dskip] (wavenet.py:41-44 under autograd):
    // z = [gate | filter] pre-activations of this layer (res_in, plain channel order, filter half n_real columns on),
    // dz_gate = dy * tanh(f) * s (1 - s),  dz_filter = dy * s * (1 - tanh(f)^2),  s = sigmoid(g); both halves written.
    const int n0 = nt * BN;
    const int Ch = p.n_real;
    const __nv_bfloat16* zr = p.res_in + row * p.ld_res + n0;
#pragma unroll 1
    for (int c = half; c < BN / 32; c += 2) {
      uint32_t a[32];
      tmem_ld_32x32(taddr + c * 32, a);
      uint4 gq[4], fq[4];
      ld_x4(zr + c * 32, gq);
      ld_x4(zr + Ch + c * 32, fq);
      tmem_ld_wait();
      float og[32], of[32];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const uint32_t gw[4] = {gq[i].x, gq[i].y, gq[i].z, gq[i].w};
        const uint32_t fw[4] = {fq[i].x, fq[i].y, fq[i].z, fq[i].w};
#pragma unroll
        for (int h = 0; h < 4; ++h) {
#pragma unroll
          for (int e = 0; e < 2; ++e) {
            const int j = 8 * i + 2 * h + e;
            const float g = __uint_as_float(e ? (gw[h] & 0xffff0000u) : (gw[h] << 16));
            const float f = __uint_as_float(e ? (fw[h] & 0xffff0000u) : (fw[h] << 16));
            const float sg = fmaf(0.5f, tanh_approx(0.5f * g), 0.5f);  // MUFU: 2^-11 relative, below the bf16 rounding of dz
            const float th = tanh_approx(f);
            const float dy = __uint_as_float(a[j]);
            og[j] = dy * th * sg * (1.0f - sg);
            of[j] = dy * sg * (1.0f - th * th);
          }
        }
      }
      epi_store32<false>(p, es, row, p.out_c0 + n0 + c * 32, b, valid, og);
      epi_store32<false>(p, es, row, p.out_c0 + n0 + Ch + c * 32, b, valid, of);
    }
  } else if constexpr (EPI == EPI_CPROJ) {
    // hoisted conditioner projection (wavenet.py:35, step-invariant): f16 out = (acc + bias) in the gate kernel's packed
    // column order, gate columns (first half of every BN-column pass) pre-multiplied by 0.5 for sigmoid-by-tanh
    const int n0 = nt * BN;
    const float* bias = p.bias + n0;
    const int grp = p.grp_n > 0 ? n0 / p.grp_n : 0;  // a BN-column pass never straddles two groups (grp_n % BN == 0)
    uint16_t* out = reinterpret_cast<uint16_t*>(p.out) + grp * p.grp_stride + row * p.ld_out + p.out_c0 + (n0 - grp * p.grp_n);
#pragma unroll 1
    for (int c = half; c < BN / 32; c += 2) {
      uint32_t a[32];
      tmem_ld_32x32(taddr + c * 32, a);
      float bb[32];
      ld_f32x32(bias + c * 32, bb);
      tmem_ld_wait();
      const float sc = (c * 32 < BN / 2) ? 0.5f : 1.0f;
      if (valid) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          uint4 q;
          q.x = pack_f16(sc * (__uint_as_float(a[8 * i + 0]) + bb[8 * i + 0]), sc * (__uint_as_float(a[8 * i + 1]) + bb[8 * i + 1]));
          q.y = pack_f16(sc * (__uint_as_float(a[8 * i + 2]) + bb[8 * i + 2]), sc * (__uint_as_float(a[8 * i + 3]) + bb[8 * i + 3]));
          q.z = pack_f16(sc * (__uint_as_float(a[8 * i + 4]) + bb[8 * i + 4]), sc * (__uint_as_float(a[8 * i + 5]) + bb[8 * i + 5]));
          q.w = pack_f16(sc * (__uint_as_float(a[8 * i + 6]) + bb[8 * i + 6]), sc * (__uint_as_float(a[8 * i + 7]) + bb[8 * i + 7]));
          reinterpret_cast<uint4*>(out + c * 32)[i] = q;
        }
      }
    }
  } else if constexpr (EPI == EPI_RES) {
    // x~_next = (x~ + R) / sqrt(2) + rvec,  rvec = (b - d_l)/sqrt(2) + d_{l+1}  (folded on the host side of the
    // step-embedding table; wavenet.py:36,44-48).  x~ is prefetched one chunk ahead.
    const int n0 = nt * BN;
    const float* rv = p.vec_add + static_cast<size_t>(b) * p.vec_bstride + n0;
    const __nv_bfloat16* xin = p.res_in + row * p.ld_res + n0;
    uint4 xq[4];
    ld_x4(xin + half * 32, xq);
    float s1 = 0.0f, s2 = 0.0f, s3 = 0.0f;
    const float* sd = p.stats_d ? p.stats_d + static_cast<size_t>(b) * p.stats_d_bstride + n0 : nullptr;
#pragma unroll 1
    for (int c = half; c < BN / 32; c += 2) {
      uint32_t a[32];
      tmem_ld_32x32(taddr + c * 32, a);
      float v[32], rr[32];
      ld_f32x32(rv + c * 32, rr);
      uint4 xn[4];
      if (c + 2 < BN / 32) ld_x4(xin + (c + 2) * 32, xn);
      uint4 xl[4];
      if constexpr (kSplit) ld_x4(xin + p.res_lo + c * 32, xl);
      tmem_ld_wait();
      if constexpr (kSplit) {  // x = hi + lo: fold the lo half into the accumulator values
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const uint32_t w[4] = {xl[i].x, xl[i].y, xl[i].z, xl[i].w};
#pragma unroll
          for (int h = 0; h < 4; ++h) {
            const int j = 8 * i + 2 * h;
            a[j] = __float_as_uint(__uint_as_float(a[j]) + __uint_as_float(w[h] << 16));
            a[j + 1] = __float_as_uint(__uint_as_float(a[j + 1]) + __uint_as_float(w[h] & 0xffff0000u));
          }
        }
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const uint32_t w[4] = {xq[i].x, xq[i].y, xq[i].z, xq[i].w};
#pragma unroll
        for (int h = 0; h < 4; ++h) {
          const int j = 8 * i + 2 * h;
          v[j] = fmaf(__uint_as_float(a[j]) + __uint_as_float(w[h] << 16), p.scale, rr[j]);
          v[j + 1] = fmaf(__uint_as_float(a[j + 1]) + __uint_as_float(w[h] & 0xffff0000u), p.scale, rr[j + 1]);
        }
      }
      if (!live) {
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = 0.0f;
      }
      if (p.stats_out != nullptr) ln_stats_acc(v, sd ? sd + c * 32 : nullptr, s1, s2, s3);
      epi_store32<kSplit>(p, es, row, p.out_c0 + n0 + c * 32, b, valid, v);
      if (p.out2) {  // pre-activated copy for the next convolution (HiFi-GAN ResBlock, nsf_hifigan/models.py:60-67)
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = v[j] > 0.0f ? v[j] : v[j] * p.slope2;
        if (valid) st_act32<false>(p.out2 + row * p.ld_out + p.out_c0 + n0 + c * 32, 0, v);
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) xq[i] = xn[i];
    }
    if (p.stats_out != nullptr && valid) p.stats_out[row * (2 * p.n_tiles) + 2 * nt + half] = make_float4(s1, s2, s3, 0.0f);
  } else {
    const int n0 = nt * BN;
    const float* bias = p.bias + n0;
    const float* vadd = p.vec_add ? p.vec_add + static_cast<size_t>(b) * p.vec_bstride + n0 : nullptr;
    float s1 = 0.0f, s2 = 0.0f, s3 = 0.0f;
    const float* sd = p.stats_d ? p.stats_d + static_cast<size_t>(b) * p.stats_d_bstride + n0 : nullptr;
#pragma unroll 1
    for (int c = half; c < BN / 32; c += 2) {
      uint32_t a[32];
      tmem_ld_32x32(taddr + c * 32, a);
      float v[32], bb[32];
      ld_f32x32(bias + c * 32, bb);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(a[j]) + bb[j];
      if (p.relu == 1) {
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j], 0.0f);
      } else if (p.relu == 2) {
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = 0.5f * v[j] * (1.0f + erff(v[j] * 0.70710678118654752f));
      } else if (p.relu == 3) {
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = v[j] > 0.0f ? v[j] : v[j] * p.slope2;
      } else if (p.relu == 4) {
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = tanhf(v[j]);
      }
#pragma unroll
      for (int j = 0; j < 32; ++j) v[j] *= p.scale;
      if constexpr (EPI == EPI_OUTF32) {
        if (valid) {
          float* out = reinterpret_cast<float*>(p.out) + row * p.ld_out + n0 + c * 32;
#pragma unroll
          for (int i = 0; i < 8; ++i)
            if (n0 + c * 32 + 4 * i < p.n_real)
              reinterpret_cast<float4*>(out)[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
        }
      } else {
        if (vadd) {
          float da[32];
          ld_f32x32(vadd + c * 32, da);
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] += da[j];
        }
        if (!live) {
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = 0.0f;
        }
        if (p.stats_out != nullptr) ln_stats_acc(v, sd ? sd + c * 32 : nullptr, s1, s2, s3);
        epi_store32<kSplit>(p, es, row, p.out_c0 + n0 + c * 32, b, valid, v);
      }
    }
    if constexpr (EPI != EPI_OUTF32) {
      if (p.stats_out != nullptr && valid) p.stats_out[row * (2 * p.n_tiles) + 2 * nt + half] = make_float4(s1, s2, s3, 0.0f);
    }
  }
}

}  // namespace dsb
