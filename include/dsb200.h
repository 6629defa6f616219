/* dsb200 -- C ABI of the B200-native DiffSinger denoiser sampling path (libdsb200.so).
 *
 * The reference (hrukalive/DiffSinger) is pure Python/PyTorch and has NO native interface; the
 * "FFI" a maintainer binds is therefore new.  Each entry point below names the reference code it
 * replaces (paths relative to the reference root).  Conventions:
 *   - every function returns 0 on success, a negative ds_status on failure; ds_last_error() gives the
 *     message (thread-local).  No exceptions cross the boundary.
 *   - all tensors are BORROWED device pointers owned by the caller (PyTorch) for the duration of
 *     the call / the object's life; no ownership is transferred.  `stream` is a cudaStream_t
 *     passed as void* (0 = legacy default stream).  Nothing here synchronises the device.
 *   - activations cross the ABI "time-major": x[B][T][F*M] fp32, cond[B][T][H] fp32 (the layout
 *     GaussianDiffusion.forward receives `condition` in, modules/core/ddpm.py:353-357).
 */
#ifndef DSB200_H_
#define DSB200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DSB200_ABI_VERSION 1

typedef enum {
  DS_OK = 0,
  DS_ERR_INVALID = -1,   /* bad argument / unsupported shape */
  DS_ERR_CUDA = -2,      /* a CUDA runtime / driver call failed */
  DS_ERR_NO_DEVICE = -3, /* no sm_100 device / driver entry point missing */
  DS_ERR_STATE = -4      /* call order violated (weights not packed, times not set, ...) */
} ds_status;

typedef struct ds_plan ds_plan;       /* one WaveNet: configuration + packed weights */
typedef struct ds_session ds_session; /* one (plan, B, T): workspace, tensor maps, captured graphs */

/* WaveNet.__init__(in_dims, n_feats, *, num_layers, num_channels, dilation_cycle_length) and
 * hparams['hidden_size']  -- modules/backbones/wavenet.py:52-73 */
typedef struct {
  int32_t in_dims;               /* M: mel bins / repeat bins */
  int32_t n_feats;               /* F */
  int32_t num_layers;            /* L */
  int32_t num_channels;          /* C, multiple of 64 */
  int32_t dilation_cycle_length; /* dilation of layer i = 2 ** (i % cycle) */
  int32_t hidden_size;           /* H, multiple of 64 */
  int32_t precision;             /* 0: bf16 operands, fp32 accumulate (final mel rel-L2 <= 1e-2)                    */
                                 /* 1: fp32-accurate, 3 x bf16 split products (per-evaluation rel-L2 <= 1e-4)       */
} ds_wavenet_config;

/* fp32 device pointers in the reference's state_dict layout (wavenet.py:22-31,56-73).
 * The per-layer members are HOST arrays of num_layers device pointers. */
typedef struct {
  const float* input_projection_weight;  /* [C, F*M, 1] */
  const float* input_projection_bias;    /* [C] */
  const float* mlp0_weight;              /* [4C, C] */
  const float* mlp0_bias;                /* [4C] */
  const float* mlp2_weight;              /* [C, 4C] */
  const float* mlp2_bias;                /* [C] */
  const float* skip_projection_weight;   /* [C, C, 1] */
  const float* skip_projection_bias;     /* [C] */
  const float* output_projection_weight; /* [F*M, C, 1] */
  const float* output_projection_bias;   /* [F*M] */
  const float* const* dilated_conv_weight;           /* L x [2C, C, 3] */
  const float* const* dilated_conv_bias;             /* L x [2C] */
  const float* const* diffusion_projection_weight;   /* L x [C, C] */
  const float* const* diffusion_projection_bias;     /* L x [C] */
  const float* const* conditioner_projection_weight; /* L x [2C, H, 1] */
  const float* const* conditioner_projection_bias;   /* L x [2C] */
  const float* const* layer_output_projection_weight; /* L x [2C, C, 1] */
  const float* const* layer_output_projection_bias;   /* L x [2C] */
} ds_wavenet_weights;

/* One instruction of a sampler program (see ds_session_run). */
enum { DS_OP_EVAL = 0, DS_OP_LINCOMB = 1 };
typedef struct {
  int32_t kind;    /* DS_OP_EVAL: bufs[dst] = backbone(bufs[src[0]], time index `time`)          */
                   /* DS_OP_LINCOMB: bufs[dst] = sum_{i<n} coef[i] * bufs[src[i]]  (dst may alias) */
  int32_t dst;
  int32_t time;    /* row of the step-embedding table installed by ds_session_set_times */
  int32_t n;       /* number of terms (LINCOMB), <= 8 */
  int32_t src[8];
  float coef[8];
} ds_op;

const char* ds_last_error(void);
int ds_abi_version(void);
/* 0 when a compute-capability-10.x device and the driver's tensor-map encoder are usable. */
int ds_device_check(void);

/* ---- plan: replaces WaveNet.__init__ + load_state_dict (wavenet.py:52-73, utils/__init__.py:166-222) */
int ds_plan_create(const ds_wavenet_config* cfg, ds_plan** out);
int ds_plan_destroy(ds_plan* plan);
size_t ds_plan_packed_bytes(const ds_plan* plan);
/* Packs/permutes/casts the fp32 reference weights into `packed` (caller-allocated device memory of
 * ds_plan_packed_bytes, 256-byte aligned, alive as long as the plan).  Re-call after a weight update. */
int ds_plan_pack_weights(ds_plan* plan, const ds_wavenet_weights* w, void* packed, size_t packed_bytes, void* stream);

/* ---- session: fixed (B, T) execution context */
size_t ds_session_workspace_bytes(const ds_plan* plan, int32_t B, int32_t T, int32_t max_times, int32_t per_sample_times);
int ds_session_create(ds_plan* plan, int32_t B, int32_t T, int32_t max_times, int32_t per_sample_times, void* workspace,
                      size_t workspace_bytes, ds_session** out);
int ds_session_destroy(ds_session* s);
/* cond[B][T][H] fp32 -> internal bf16 copy.  Replaces the per-layer conditioner input of
 * ResidualBlock.forward (wavenet.py:35); cond is step-invariant so it is staged once per call. */
int ds_session_set_cond(ds_session* s, const float* cond_bth, void* stream);
/* Step-embedding tables for n_times diffusion-step values (device fp32 array of n_times values; of
 * n_times*B values ordered [time][batch] when the session was created with per_sample_times = 1).
 * Replaces SinusoidalPosEmb + mlp + every layer's diffusion_projection
 * (common_layers.py:273-280, wavenet.py:89-90, :34). */
int ds_session_set_times(ds_session* s, const float* t_dev, int32_t n_times, void* stream);
/* One backbone evaluation, out = WaveNet.forward(x, t[time_index], cond) (wavenet.py:75-107) with
 * x/out time-major [B][T][F*M] fp32. */
int ds_session_forward(ds_session* s, const float* x_btc, int32_t time_index, float* out_btc, void* stream);
/* Runs a whole sampler program (the loop bodies of GaussianDiffusion.inference, ddpm.py:221-351,
 * and RectifiedFlow.inference, reflow.py:104-138, flattened by the host into EVAL/LINCOMB ops).
 * bufs: n_bufs fp32 device buffers of B*T*F*M elements.  use_graph != 0 captures the program into a
 * CUDA graph on first use (keyed by ops + buffer addresses) and replays it afterwards. */
int ds_session_run(ds_session* s, const ds_op* ops, int32_t n_ops, float* const* bufs, int32_t n_bufs, int32_t use_graph,
                   void* stream);
/* Number of kernel launches the last ds_session_forward / ds_session_run issued (or replayed). */
int64_t ds_session_last_launches(const ds_session* s);
/* Which kernels the session's residual layers run on, chosen at creation from (B, T) and the plan: 0 = one fused kernel per
 * layer (ds_plan_layer_mode), 1 = unfused GEMMs with narrow N tiles (few frame tiles), 2 = unfused GEMMs with the plan's 256-wide
 * tiles (between the two), 3 = the whole network in ONE launch per evaluation (csrc/stack_kernel.cuh: launch-bound sessions of a
 * few 128-frame tiles at C = 256, e.g. batch-1 .ds segments and the variance predictors); negative on a null session. */
int ds_session_path(const ds_session* s);
/* Sessions of this plan with at most this many 128-frame tiles (B * ceil(T / 128)) run on the whole-network kernel (path 3); 0 when
 * the plan cannot (C not 192 / 256, dilations > 16, ...) or DSB200_STACK=0.  It is also the number of tile clusters the device can
 * hold at once: neighbouring tiles of an utterance wait for each other's edge frames, so all clusters of a launch must be resident,
 * and callers that run several such sessions CONCURRENTLY (different streams) should keep the tiles in flight below this number
 * (diffsinger_b200.sample_segments does).  No counterpart in the reference (it launches ~240 library kernels per evaluation). */
int ds_plan_stack_capacity(const ds_plan* plan);
/* Ragged batch: utterance b of the session's [B, T] layout has only lengths[b] (1 .. T) frames.  Frames at and beyond it are the
 * dilated convolutions' zero padding, exactly as if the utterance ran alone: every kernel that writes the residual stream x~ stores
 * zeros there (whole-network kernel: tiles entirely beyond the end also exit at once), so batching utterances of different lengths
 * does not change any utterance's result.  Output rows beyond lengths[b] are undefined.  lengths = NULL (or n = 0) restores B
 * full-length utterances.  The values are copied into session memory on `stream` (graphs captured by ds_session_run read them at
 * replay).  This is how diffsinger_b200.sample_segments runs many batch-1 `.ds` segments of different lengths
 * (inference/ds_acoustic.py:234-244 loops over them one by one) as one batch.  No counterpart in the reference (its batches are
 * padded and unmasked, SURVEY.md trap T4). */
int ds_session_set_lengths(ds_session* s, const int32_t* lengths_host, int32_t n, void* stream);

/* ---- no-allocation entry point.  NO COUNTERPART IN THE REFERENCE: added for callers that
 * cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that forbid it).
 * out = WaveNet.forward(x, t, cond) exactly as ds_session_set_cond + ds_session_set_times(1 value) + ds_session_forward compute it,
 * but the session workspace is a file-scope `static __device__` array of DS_STATIC_SCRATCH_BYTES inside the library (held by every
 * GPU that loads the library, until exit) instead of caller memory; the call performs no cudaMalloc / cudaFree / allocating
 * library call.  cond_bth [B][T][H], x_btc / out_btc [B][T][F*M] fp32 device pointers; t_dev: 1 fp32 device value, or B values
 * when per_sample_times != 0.  Shapes whose workspace (ds_session_workspace_bytes(plan, B, T, 1, per_sample_times)) exceeds the
 * array are rejected with DS_ERR_INVALID.  Every call on a device shares the one array, which keeps the previous call's
 * contents (nothing in it needs to start at zero); calls that overlap on the GPU (different streams / threads) RACE on it. */
#define DS_STATIC_SCRATCH_BYTES (32u << 20)
size_t ds_static_scratch_bytes(void);
int ds_forward_static(ds_plan* plan, const float* cond_bth, const float* t_dev, int32_t per_sample_times, const float* x_btc,
                      float* out_btc, int32_t B, int32_t T, void* stream);

/* ---- stand-alone helpers */
/* dst = sum coef[i]*src[i] over numel fp32 elements (numel % 4 == 0): the sampler tensor update
 * (ddpm.py:164-166,181-201; dpm_solver_pytorch.py:573-576,827-831; uni_pc.py:548-567; reflow.py:68-100). */
int ds_lincomb(float* dst, const float* const* srcs, const float* coefs, int32_t n, size_t numel, void* stream);
/* [B][R][C] -> [B][C][R] fp32 (spec [B,F*M,T] <-> time-major; ddpm.py:350, reflow.py:137). */
int ds_transpose(float* dst, const float* src, int32_t B, int32_t R, int32_t C, void* stream);
/* Start state of shallow diffusion in one pass (the step immediately before the loop): replaces norm_spec
 * (ddpm.py:379-380) + q_sample (ddpm.py:206-210,235-239) -- or RectifiedFlow's `t_start * x_end + (1 - t_start) * noise`
 * (reflow.py:112-118) -- + the transpose to time-major:  x[b][t][m] = ca * ((src[b][t][m] - min[m]) / (max[m] - min[m]) * 2 - 1) +
 * cb * noise[b][m][t].  src: the aux decoder's mel [B][T][M] fp32, noise: [B][M][T] fp32 as torch.randn(B, 1, M, T) draws it;
 * every operation rounded separately in the reference's order (bit-identical to the eager sequence). */
int ds_shallow_start(float* x_btm, const float* src_btm, const float* noise_bmt, const float* spec_min, const float* spec_max, float ca,
                     float cb, int32_t B, int32_t T, int32_t M, void* stream);

/* ---- primitive ops: one tcgen05 GEMM launch with an explicit epilogue, LayerNorm, depthwise conv.  They are the unit-test
 * hooks of the kernels above and the building blocks of the path's neighbours behind the same registry (LYNXNet,
 * modules/backbones/lynxnet.py:90-163).  Activations are time-major bf16 [B][T][channels]; weights bf16 [rows][K]. */
typedef struct {
  int32_t epi;  /* 0 linear: bf16 out = act(acc + bias) * scale + vec_add[b]      1 gate: sigmoid(first half) * tanh(second half)
                   2 residual: bf16 out = (acc + res_in) * scale + vec_add[b]     3 fp32 out = act(acc + bias) * scale
                   5 SwiGLU: first half * silu(second half)   (halves = BN/2-row blocks of the same channels, see pack order)
                   6 gate backward (training): acc = dy; res_in = z = [gate | filter] of the layer (filter half n_real columns
                     on); out[n] = dy * tanh(f) * s (1 - s), out[n_real + n] = dy * s * (1 - tanh(f)^2), s = sigmoid(g) */
  int32_t BN;   /* 64 / 128 / 256 */
  int32_t B, T;
  const void* a0; int32_t a0_channels; int32_t a0_c0; int32_t kb_a0; int32_t n_taps; int32_t dil;
  const void* a1; int32_t a1_channels; int32_t kb_a1;
  const void* w; int32_t w_rows; int32_t w_k; int32_t b_row0; int32_t n_tiles;
  const float* bias; const float* vec_sub; const float* vec_add; int32_t vec_bstride;
  const void* res_in; int32_t ld_res;
  void* out; int32_t ld_out; int32_t out_c0; int32_t n_real;
  int32_t relu;  /* activation of epilogues 0 / 3: 0 none, 1 ReLU, 2 GELU (erf), 3 LeakyReLU(slope2), 4 tanh */
  float scale;
  void* out2;    /* epilogue 2 (residual), optional: second bf16 output LeakyReLU(out, slope2), same ld / offset as out --
                  * the pre-activated copy the next convolution of a HiFi-GAN ResBlock reads (nsf_hifigan/models.py:60-67) */
  float slope2;
  /* LayerNorm folded into the GEMMs around it (the reference normalises between them: lynxnet.py:52-56 LayerNorm -> Conv1d(dim,
   * 2*inner, 1) -> SwiGLU, with the step vector added to the norm's input, :82-89).  Producer side (epilogues 0 / 2): per frame and
   * per (N tile, warp half) the partial sums {sum r, sum r^2, sum r*d} over the bf16 outputs r and the NEXT norm's step vector d
   * (stats_d, [B or 1][N]) are written to stats_out[frame][2 * n_tiles] (float4).  Consumer side (epilogue 5): the frame's mean /
   * rstd of (r + d) over ln_channels come from those sums, and y[n] = rstd * (acc[n] + e[n] - mean * s[n]) + bias[n] with the
   * host-folded s = W~ 1, e = W~ d, W~ = W diag(gamma), bias = W beta + b; ln_e rows are [N + 4]: e, then sum d, sum d^2.  All NULL /
   * zero: plain epilogues. */
  void* stats_out; const float* stats_d; int32_t stats_d_bstride;
  const void* ln_stats; int32_t ln_parts; int32_t ln_channels; float ln_eps;
  const float* ln_s; const float* ln_e; int32_t ln_e_bstride;
  /* Ragged batch (see ds_session_set_lengths): optional device array [B] of frame counts; the bf16 outputs of epilogues 0 / 2 / 5 for
   * frames t >= lengths[b] are written as ZEROS, so that a convolution over time that reads them next (LYNXNet's depthwise k = 31
   * after the SwiGLU GEMM, lynxnet.py:57-59) sees the utterance's own zero padding. */
  const int32_t* lengths;
} ds_gemm_desc;
int ds_debug_gemm(const ds_gemm_desc* d, void* stream);
int ds_gemm(const ds_gemm_desc* d, void* stream);
/* y = LayerNorm(x + add + vec[row / rows_per_vec]) over `channels` (multiple of 8, <= 2048) for `rows` frames;
 * torch.nn.LayerNorm semantics (lynxnet.py:52,123).  add (bf16 [rows][channels]) and vec (fp32 [rows / rows_per_vec][channels])
 * may be NULL: the conditioner projection and the per-utterance diffusion-step projection added before the norm
 * (lynxnet.py:82-89). */
int ds_layernorm(void* y_bf16, const void* x_bf16, const float* gamma, const float* beta, int64_t rows, int32_t channels, float eps,
                 const float* vec, int64_t rows_per_vec, const void* add_bf16, void* stream);
/* Depthwise Conv1d along time with zero padding (nn.Conv1d(ch, ch, k, padding=pad_left, groups=ch), lynxnet.py:59;
 * convnext.py dwconv), w fp32 [channels][k], fused bias and activation: 0 none, 1 ReLU, 2 PReLU(slope[channels]), 3 SiLU. */
int ds_dwconv(void* y_bf16, const void* x_bf16, const float* w, const float* bias, const float* slope, int32_t B, int32_t T,
              int32_t channels, int32_t kernel_size, int32_t pad_left, int32_t act, void* stream);
/* fp32 [rows][cols] -> bf16 [rows][ld] (columns >= cols zero filled); cols, ld multiples of 4. */
int ds_cast_bf16(void* dst_bf16, const float* src, int64_t rows, int32_t cols, int32_t ld, void* stream);
/* NSF-HiFiGAN vocoder, streaming helpers around ds_gemm:
 *   ds_source_conv: y[b][n][c] = bias[c] + sum_k w[k][c] * src[b][n * stride + k - pad] (zero padded), c < ld; w fp32 tap-major
 *     [kernel_size][ld], bias fp32 [ld] (pad columns zero) -- the noise_convs Conv1d(1, ch, k, stride) over the harmonic source
 *     (nsf_hifigan/models.py:237-241, 281).
 *   ds_mean3_lrelu: y = LeakyReLU((a + b + c) / 3, slope) -- the average of the three ResBlocks followed by the next
 *     stage's activation (models.py:286-294); n multiple of 8. */
int ds_source_conv(void* y_bf16, const float* src, const float* w, const float* bias, int32_t B, int64_t n_src, int64_t n_out,
                   int32_t ld, int32_t kernel_size, int32_t stride, int32_t pad, void* stream);
/* k-tap dilated Conv1d along time for narrow channel counts (c_in = c_out in {16, 32}; taps odd <= 11, 128 + (taps - 1) * dil
 * <= 192): the 32- / 16-channel ResBlock convolutions, the last transposed conv in polyphase form and conv_post of the
 * generator (models.py:35-67, 285-298).  x bf16 [B][T][c_in]; w_img bf16 [taps][c_in / 8][c_out][8] (w_img[k][p][n][j] =
 * weight[n][8 p + j][k]); out = act(conv + bias + res) bf16 [B][T][c_out], out2 (optional) = LeakyReLU(out, slope);
 * act: 0 none, 3 LeakyReLU(slope), 4 tanh with out = fp32 [B][T] holding output channel 0 only. */
int ds_conv_small(const void* x_bf16, int32_t c_in, const void* w_img_bf16, int32_t c_out, const float* bias, const void* res_bf16,
                  void* out, void* out2_bf16, int32_t B, int32_t T, int32_t taps, int32_t dil, int32_t act, float slope, void* stream);
int ds_mean3_lrelu(void* y_bf16, const void* a_bf16, const void* b_bf16, const void* c_bf16, int64_t n, float slope, void* stream);
/* Condition producer: the FastSpeech2 acoustic encoder in front of the sampler (modules/fastspeech/
 * acoustic_encoder.py:82-118, tts_modules.py:353-440, commons/common_layers.py:120-258).  Its Linear / Conv1d / LayerNorm calls are
 * ds_gemm / ds_layernorm; these three are the rest:
 *   ds_attention: multi-head self-attention over the token axis with optional rotary position embedding
 *     (MultiheadSelfAttentionWithRoPE.forward, common_layers.py:175-213; rotary_embedding_torch.py:35-39,119,174-188).  qkv bf16
 *     [B][T][3H] = in_proj output (q | k | v, heads contiguous; q and k are rotated IN PLACE when rope != 0), out bf16 [B][T][H] =
 *     softmax(q k^T / sqrt(D) | key not padded) v with heads concatenated; key_pad_mask [B][T] (nonzero = padded key) or NULL.
 *   ds_mask_rows: x[row][:] = 0 where mask[row] != 0 (`x * (1 - padding_mask)`, common_layers.py:246,253).
 *   ds_length_regulate: cond[b][t][:] = enc[b][mel2ph[b][t] - 1][:] (zeros for mel2ph == 0) + sum_j (weight_j * value_j[b, t] + bias_j)
 *     + add[b][t][:] -- the torch.gather length regulation plus pitch / variance / key-shift / speed embeddings (each nn.Linear(1, H))
 *     and the speaker embedding of acoustic_encoder.py:97-116, one pass, fp32 out [B][T][H] (what GaussianDiffusion.forward takes). */
typedef struct {
  const float* value;          /* curve sample of (b, t) at value[b * stride_b + t * stride_t] (a stride of 0 broadcasts) */
  int64_t stride_b, stride_t;
  const float* weight;         /* [H], 16-byte aligned */
  const float* bias;           /* [H], 16-byte aligned */
} ds_curve_embed;
int ds_attention(void* qkv_bf16, void* out_bf16, const uint8_t* key_pad_mask, int32_t B, int32_t T, int32_t H, int32_t heads, int32_t rope,
                 float theta, void* stream);
int ds_mask_rows(void* x_bf16, const uint8_t* mask, int64_t rows, int32_t channels, void* stream);
int ds_length_regulate(float* cond_bth, const void* enc_bf16, const int64_t* mel2ph, int32_t B, int32_t T, int32_t T_txt, int32_t H,
                       const ds_curve_embed* curves, int32_t n_curves, const float* add, int64_t add_stride_b, int64_t add_stride_t,
                       void* stream);
/* Training path (ddpm.py:212-219 / reflow.py:36-41 under autograd).  z = [gate | filter] pre-activations of
 * one residual layer (wavenet.py:40), bf16 [rows][ld_z], plain channel order:
 *   ds_gate_fwd:  y[r][c] = sigmoid(z[r][c]) * tanh(z[r][channels + c])                                       (wavenet.py:41-42)
 *   ds_gate_bwd:  dz[r][c] = dy * tanh(f) * s (1 - s),  dz[r][channels + c] = dy * s * (1 - tanh(f)^2),  s = sigmoid(g)
 * channels and every ld multiple of 8, pointers 16-byte aligned (column slices of wider buffers are fine). */
int ds_gate_fwd(void* y_bf16, int64_t ld_y, const void* z_bf16, int64_t ld_z, int64_t rows, int32_t channels, void* stream);
int ds_gate_bwd(void* dz_bf16, int64_t ld_dz, const void* dy_bf16, int64_t ld_dy, const void* z_bf16, int64_t ld_z, int64_t rows,
                int32_t channels, void* stream);
/* Weight-gradient GEMMs of the training path: for every job (one 128 x BN output tile, BN in {64, 128, 256})
 *   out[out_off + m * ld_out + n] = sum_{b < B, t < T} G[g_b0 + b][t][g_c0 + m] * X[x_b0 + b][t + shift][x_c0 + n],
 *   m < m_valid <= 128, n < n_valid <= BN, frames outside [0, T) read as zeros (the conv's padding);
 * e.g. d(dilated_conv.weight)[:, :, tap] = dz^T x(t + (tap - 1) dil) (wavenet.py:36-38 under autograd).  (G, X) = operand
 * pair `pair` of the launch (up to 4 pairs, so that all weight tensors of one backward pass share a launch): bf16
 * [batch rows][T][channels]; channel offsets multiples of 8; ld_out and out_off multiples of 4; jobs in DEVICE memory. */
typedef struct ds_wgrad_operands {
  const void* g_bf16;
  const void* x_bf16;
  int32_t g_channels, g_batch_rows, x_channels, x_batch_rows;
} ds_wgrad_operands;
typedef struct ds_wgrad_job {
  int32_t g_c0, g_b0, x_c0, x_b0, shift, m_valid, n_valid, ld_out;
  int64_t out_off;
  int32_t pair, reserved;
} ds_wgrad_job;
int ds_wgrad(const ds_wgrad_operands* pairs, int32_t n_pairs, int32_t T, int32_t B, int32_t BN, const ds_wgrad_job* jobs_dev,
             int32_t n_jobs, float* out, void* stream);
/* Debug: device buffer of n_entries x 4 uint64 into which CTA 0 of every fused-layer launch records
 * {role*16+kind, tile*16+idx, t0_ns, t1_ns} (first half: MMA issuer, second half: epilogue). NULL disables. */
int ds_debug_set_trace(void* dev_buf, int32_t n_entries);
/* Debug / bench: launch ONLY the fused kernel of residual layer `layer` (ResidualBlock.forward, wavenet.py:33-48) on the
 * session's current activation buffers -- the per-kernel timing hook bench.py's roofline uses. */
int ds_debug_layer(ds_session* s, int32_t layer, int32_t time_index, void* stream);
/* 0 = two GEMM launches per layer, 5 = one fused kernel per layer (csrc/layer_kernel.cuh: CTA pairs, halo tiles, grouped pipeline
 * stages; 1-4 were earlier generations of that kernel, removed). */
int ds_plan_layer_mode(const ds_plan* plan);
/* 0 bf16, 1 fp32-accurate (ds_wavenet_config.precision, or DSB200_PRECISION=fp32 in the environment). */
int ds_plan_precision(const ds_plan* plan);
/* Debug: raw tcgen05.mma rate micro-benchmark (csrc/mma_bench.cuh). */
int ds_debug_mma_rate(int32_t pair, int32_t grid, int32_t n_mma, int32_t commit_every, long long* out_dev, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DSB200_H_ */
