// Kernels of the CONDITION PRODUCER: the pieces of the FastSpeech2 acoustic encoder that are not GEMMs /
// LayerNorms -- rotary position embedding + self-attention over the phoneme tokens, row masking, and the length regulator that
// expands token states to frames and adds the per-frame embeddings.
//   reference: modules/commons/common_layers.py:152-213 (MultiheadSelfAttentionWithRoPE), modules/commons/rotary_embedding_torch.py
//   (rotate_half :35-39, freqs :119, rotate_queries_or_keys :174-188), modules/fastspeech/acoustic_encoder.py:82-118.
// Token sequences are short (tens to a few hundred phonemes per `.ds` segment): these kernels are latency-, not throughput-bound;
// one warp owns one query row, fp32 math, bf16 storage between the GEMMs.
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cstdint>

namespace dsb {

// In-place rotary embedding of the q and k thirds of qkv [B*T][3H] (bf16): for every head and pair i < D/2
//   (x[2i], x[2i+1]) <- (x[2i] cos a - x[2i+1] sin a,  x[2i+1] cos a + x[2i] sin a),  a = t * theta^(-2i / D)
// (interleaved pairs: rotary_embedding_torch's rotate_half; positions 0..T-1, no offset / xpos / interpolation).
__global__ void __launch_bounds__(256) rope_kernel(__nv_bfloat16* __restrict__ qkv, int B, int T, int H, int heads, float theta) {
  const int D = H / heads;
  const long pairs = static_cast<long>(B) * T * H;  // q and k: 2 * (H / 2) pairs per row
  for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < pairs; i += static_cast<long>(gridDim.x) * blockDim.x) {
    const long row = i / H;
    const int rem = static_cast<int>(i % H);
    const int which = rem / (H / 2);        // 0 = q, 1 = k
    const int pr = rem % (H / 2);           // pair index inside the H-wide block
    const int pi = pr % (D / 2);            // pair index inside the head
    const int t = static_cast<int>(row % T);
    const float freq = exp2f(-log2f(theta) * (2.0f * pi) / D);
    float sn, cs;
    sincosf(static_cast<float>(t) * freq, &sn, &cs);
    __nv_bfloat162* p = reinterpret_cast<__nv_bfloat162*>(qkv + row * 3 * H + which * H) + pr;
    const float2 v = __bfloat1622float2(*p);
    *p = __floats2bfloat162_rn(v.x * cs - v.y * sn, v.y * cs + v.x * sn);
  }
}

// out[b, t, h*D + d] = softmax_j(q_t . k_j / sqrt(D) | key j not padded) v_j, one warp per (b, head, query), online softmax.
// qkv bf16 [B*T][3H] (q | k | v, heads contiguous inside each third), pad: optional [B][T], nonzero = padded key.
template <int DPL>  // head dims per lane: D = 32 * DPL
__global__ void __launch_bounds__(256) attention_kernel(const __nv_bfloat16* __restrict__ qkv, __nv_bfloat16* __restrict__ out,
                                                        const uint8_t* __restrict__ pad, int T, int H) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int t = blockIdx.x * 8 + warp;
  const int h = blockIdx.y, b = blockIdx.z;
  if (t >= T) return;
  constexpr int D = 32 * DPL;
  const float scale = rsqrtf(static_cast<float>(D));
  const __nv_bfloat16* base = qkv + static_cast<long>(b) * T * 3 * H + h * D + lane * DPL;
  float q[DPL], acc[DPL];
#pragma unroll
  for (int d = 0; d < DPL; ++d) {
    q[d] = __bfloat162float(base[static_cast<long>(t) * 3 * H + d]) * scale;
    acc[d] = 0.0f;
  }
  float m = -INFINITY, l = 0.0f;
  for (int j = 0; j < T; ++j) {
    if (pad != nullptr && pad[static_cast<long>(b) * T + j]) continue;
    const __nv_bfloat16* kj = base + static_cast<long>(j) * 3 * H + H;
    float s = 0.0f;
#pragma unroll
    for (int d = 0; d < DPL; ++d) s = fmaf(q[d], __bfloat162float(kj[d]), s);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    const float mn = fmaxf(m, s);
    const float c = __expf(m - mn), p = __expf(s - mn);
    l = l * c + p;
    const __nv_bfloat16* vj = kj + H;
#pragma unroll
    for (int d = 0; d < DPL; ++d) acc[d] = fmaf(p, __bfloat162float(vj[d]), acc[d] * c);
    m = mn;
  }
  const float inv = 1.0f / l;
  __nv_bfloat16* o = out + (static_cast<long>(b) * T + t) * H + h * D + lane * DPL;
#pragma unroll
  for (int d = 0; d < DPL; ++d) o[d] = __float2bfloat16_rn(acc[d] * inv);
}

// x[row][:] = 0 where mask[row] != 0 (the `x * (1 - padding_mask)` after every sub-layer, common_layers.py:246,253).
__global__ void __launch_bounds__(256) mask_rows_kernel(__nv_bfloat16* __restrict__ x, const uint8_t* __restrict__ mask, long rows, int ch8) {
  const long n = rows * ch8;
  for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < n; i += static_cast<long>(gridDim.x) * blockDim.x)
    if (mask[i / ch8]) reinterpret_cast<uint4*>(x)[i] = make_uint4(0, 0, 0, 0);
}

// Length regulator + per-frame embeddings (acoustic_encoder.py:97-116):
//   cond[b, t, :] = enc[b, mel2ph[b, t] - 1, :]  (zeros where mel2ph == 0: the F.pad row)  + sum_j (w_j * curve_j[b, t] + bias_j) + add[b, t, :]
struct LrCurve {
  const float* v;  // value at [b * sb + t * st]
  long sb, st;
  const float* w;  // [H]  (nn.Linear(1, H).weight[:, 0])
  const float* bias;
};
constexpr int kLrMaxCurves = 8;
struct LrParams {
  LrCurve c[kLrMaxCurves];
  int n;
  const float* add;  // optional [.. H], strides below (speaker embedding / mix)
  long add_sb, add_st;
};
__global__ void __launch_bounds__(256) length_regulate_kernel(float* __restrict__ cond, const __nv_bfloat16* __restrict__ enc,
                                                              const long long* __restrict__ mel2ph, int B, int T, int T_txt, int H,
                                                              const LrParams p) {
  const int h4 = H / 4;
  const long n = static_cast<long>(B) * T * h4;
  for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < n; i += static_cast<long>(gridDim.x) * blockDim.x) {
    const long bt = i / h4;
    const int c0 = static_cast<int>(i % h4) * 4;
    const int b = static_cast<int>(bt / T), t = static_cast<int>(bt % T);
    const long long ph = mel2ph[bt];
    float o[4] = {0.0f, 0.0f, 0.0f, 0.0f};
    if (ph > 0 && ph <= T_txt) {
      const __nv_bfloat162* e = reinterpret_cast<const __nv_bfloat162*>(enc + (static_cast<long>(b) * T_txt + (ph - 1)) * H + c0);
      const float2 e0 = __bfloat1622float2(e[0]), e1 = __bfloat1622float2(e[1]);
      o[0] = e0.x; o[1] = e0.y; o[2] = e1.x; o[3] = e1.y;
    }
    for (int j = 0; j < p.n; ++j) {
      const float v = __ldg(p.c[j].v + b * p.c[j].sb + t * p.c[j].st);
      const float4 w = __ldg(reinterpret_cast<const float4*>(p.c[j].w + c0));
      const float4 bb = __ldg(reinterpret_cast<const float4*>(p.c[j].bias + c0));
      o[0] += fmaf(w.x, v, bb.x); o[1] += fmaf(w.y, v, bb.y); o[2] += fmaf(w.z, v, bb.z); o[3] += fmaf(w.w, v, bb.w);
    }
    if (p.add != nullptr) {
      const float4 a = __ldg(reinterpret_cast<const float4*>(p.add + b * p.add_sb + t * p.add_st + c0));
      o[0] += a.x; o[1] += a.y; o[2] += a.z; o[3] += a.w;
    }
    *reinterpret_cast<float4*>(cond + bt * H + c0) = make_float4(o[0], o[1], o[2], o[3]);
  }
}

}  // namespace dsb
