"""Thin host wrappers of the primitive ops of libdsb200.so (include/dsb200.h: ds_gemm, ds_layernorm, ds_dwconv, ds_cast_bf16,
ds_lincomb).  Tensors are time-major: bf16 activations [B*T, channels], bf16 weights [rows][K], fp32 vectors.  Every call
launches on the current CUDA stream and returns immediately; nothing here computes in PyTorch."""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence, Tuple

import torch

from . import _native as nat

EPI_LINEAR, EPI_GATE, EPI_RES, EPI_OUTF32, EPI_SWIGLU, EPI_GATEBWD = 0, 1, 2, 3, 5, 6
ACT_NONE, ACT_RELU, ACT_GELU, ACT_LRELU, ACT_TANH = 0, 1, 2, 3, 4
DW_NONE, DW_RELU, DW_PRELU, DW_SILU = 0, 1, 2, 3


def stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def tile_n(n: int) -> int:
    """N tile of a GEMM with n output rows (multiple of 64)."""
    if n % 64:
        raise nat.NativeError(f"GEMM width {n} must be a multiple of 64")
    return 256 if n % 256 == 0 else (128 if n % 128 == 0 else 64)


def pad64(n: int) -> int:
    return (n + 63) // 64 * 64


def _p(t: Optional[torch.Tensor]):
    return t.data_ptr() if t is not None else None


def gemm(epi: int, BN: int, B: int, T: int, a: torch.Tensor, w: torch.Tensor, out: torch.Tensor, *, k_per_tap: Optional[int] = None,
         n_taps: int = 1, dil: int = 1, a_c0: int = 0, a1: Optional[torch.Tensor] = None, bias: Optional[torch.Tensor] = None,
         vec: Optional[torch.Tensor] = None, vec_bstride: int = 0, res: Optional[torch.Tensor] = None, act: int = ACT_NONE,
         n_real: int = 0, scale: float = 1.0, out_c0: int = 0, out2: Optional[torch.Tensor] = None, slope2: float = 0.0,
         res_c0: int = 0, stats_out: Optional[torch.Tensor] = None, stats_d: Optional[torch.Tensor] = None,
         ln: Optional[dict] = None, lengths: Optional[torch.Tensor] = None):
    """out[b, t, out_c0 + n] = epilogue(sum_tap sum_k a[b, t + (tap - n_taps // 2) * dil, a_c0 + k] * w[n, tap * K + k]
    + sum_h a1[b, t, h] * w[n, n_taps * K + h]); frames outside [0, T) read as zeros (the conv's zero padding).
    a: [B*T, ld_a] bf16 (K = k_per_tap columns from a_c0), a1: optional second operand [B*T, H] (one tap),
    w: [rows, n_taps * K (+ H)] bf16, out: [B*T, ld_out] bf16 (fp32 for EPI_OUTF32), res: [B*T, ld_res] bf16 (EPI_RES).
    Folded LayerNorm (include/dsb200.h: ds_gemm_desc): stats_out fp32 [B*T, 2 * rows / BN, 4] receives the producer's partial sums
    (stats_d: the next norm's step vector [Bt, rows]); ln = dict(stats, channels, eps, s [rows], e [Bt, rows + 4]) makes the SwiGLU
    epilogue apply the norm.  lengths: int32 [B] device tensor of a ragged batch (outputs beyond an utterance's end are zeros)."""
    a_ch = a.shape[1]
    k = a_ch if k_per_tap is None else k_per_tap
    h = a1.shape[1] if a1 is not None else 0
    rows, w_k = w.shape
    assert w_k == n_taps * k + h and k % 64 == 0 and h % 64 == 0 and rows % BN == 0 and a_c0 + k <= a_ch, (w.shape, n_taps, k, h, BN)
    d = nat.GemmDesc()
    d.epi, d.BN, d.B, d.T = epi, BN, B, T
    d.a0, d.a0_channels, d.a0_c0, d.kb_a0, d.n_taps, d.dil = a.data_ptr(), a_ch, a_c0, k // 64, n_taps, dil
    d.a1, d.a1_channels, d.kb_a1 = _p(a1), h, h // 64
    d.w, d.w_rows, d.w_k, d.b_row0, d.n_tiles = w.data_ptr(), rows, w_k, 0, rows // BN
    d.bias, d.vec_sub, d.vec_add, d.vec_bstride = _p(bias), None, _p(vec), vec_bstride
    d.res_in, d.ld_res = (res.data_ptr() + 2 * res_c0 if res is not None else None), (res.shape[1] if res is not None else 0)
    d.out, d.ld_out, d.out_c0, d.n_real = out.data_ptr(), out.shape[-1], out_c0, n_real
    d.relu, d.scale = act, scale
    d.out2, d.slope2 = _p(out2), slope2
    if stats_out is not None:
        assert stats_out.dtype == torch.float32 and stats_out.numel() == B * T * 2 * (rows // BN) * 4
        d.stats_out, d.stats_d = stats_out.data_ptr(), _p(stats_d)
        d.stats_d_bstride = stats_d.shape[-1] if stats_d is not None and stats_d.shape[0] > 1 else 0
    if ln is not None:
        st, e = ln["stats"], ln["e"]
        assert e.shape[-1] == rows + 4 and ln["s"].numel() == rows
        d.ln_stats, d.ln_parts, d.ln_channels, d.ln_eps = st.data_ptr(), st.numel() // (B * T * 4), ln["channels"], ln["eps"]
        d.ln_s, d.ln_e, d.ln_e_bstride = ln["s"].data_ptr(), e.data_ptr(), (e.shape[-1] if e.shape[0] > 1 else 0)
    if lengths is not None:
        assert lengths.dtype == torch.int32 and lengths.numel() == B and lengths.is_cuda
        d.lengths = lengths.data_ptr()
    nat.check(nat.lib().ds_gemm(C.byref(d), stream()))


def gate_fwd(y: torch.Tensor, y_c0: int, z: torch.Tensor, z_c0: int, channels: int):
    """y[:, y_c0 + c] = sigmoid(z[:, z_c0 + c]) * tanh(z[:, z_c0 + channels + c]); bf16 [rows, ld] buffers."""
    es = 2
    nat.check(nat.lib().ds_gate_fwd(y.data_ptr() + y_c0 * es, y.shape[1], z.data_ptr() + z_c0 * es, z.shape[1], y.shape[0], channels,
                                    stream()))


def gate_bwd(dz: torch.Tensor, dz_c0: int, dy: torch.Tensor, z: torch.Tensor, z_c0: int, channels: int):
    """dz[:, dz_c0 : dz_c0 + 2 channels] = d(sigmoid(g) tanh(f)) / d[g | f] * dy; bf16 [rows, ld] buffers."""
    es = 2
    nat.check(nat.lib().ds_gate_bwd(dz.data_ptr() + dz_c0 * es, dz.shape[1], dy.data_ptr(), dy.shape[1], z.data_ptr() + z_c0 * es,
                                    z.shape[1], dz.shape[0], channels, stream()))


WGRAD_JOB_DTYPE = [("g_c0", "<i4"), ("g_b0", "<i4"), ("x_c0", "<i4"), ("x_b0", "<i4"), ("shift", "<i4"), ("m_valid", "<i4"),
                   ("n_valid", "<i4"), ("ld_out", "<i4"), ("out_off", "<i8"), ("pair", "<i4"), ("reserved", "<i4")]
WGRAD_JOB_BYTES = 48


def wgrad_jobs(jobs, device) -> torch.Tensor:
    """Device copy of a ds_wgrad_job table: jobs = iterable of (g_c0, g_b0, x_c0, x_b0, shift, m_valid, n_valid, ld_out, out_off,
    pair)."""
    import numpy as np
    arr = np.array([tuple(j) + (0,) for j in jobs], dtype=WGRAD_JOB_DTYPE)
    assert arr.itemsize == WGRAD_JOB_BYTES
    return torch.from_numpy(arr.view(np.uint8).copy()).to(device)


def wgrad(out: torch.Tensor, pairs: Sequence[Tuple[torch.Tensor, torch.Tensor]], T: int, B: int, BN: int, jobs: torch.Tensor):
    """out (fp32, flat) tiles = sum over frames of g^T x per job (include/dsb200.h: ds_wgrad).  pairs: up to four (g, x) operand
    pairs, g: [batch rows * T, g_channels] bf16, x: [batch rows * T, x_channels] bf16; jobs from wgrad_jobs()."""
    arr = (nat.WgradOperands * len(pairs))()
    for o, (g, x) in zip(arr, pairs):
        o.g_bf16, o.x_bf16 = g.data_ptr(), x.data_ptr()
        o.g_channels, o.g_batch_rows, o.x_channels, o.x_batch_rows = g.shape[1], g.shape[0] // T, x.shape[1], x.shape[0] // T
    nat.check(nat.lib().ds_wgrad(arr, len(pairs), T, B, BN, jobs.data_ptr(), jobs.numel() // WGRAD_JOB_BYTES, out.data_ptr(), stream()))


def source_conv(y: torch.Tensor, src: torch.Tensor, w_t: torch.Tensor, bias: torch.Tensor, stride: int, pad: int):
    """y[b, n, c] = bias[c] + sum_k w_t[k, c] * src[b, n * stride + k - pad]; y bf16 [B, n_out, ld], src fp32 [B, n_src],
    w_t fp32 [k, ld] (tap-major), bias fp32 [ld]."""
    B, n_out, ld = y.shape
    assert w_t.shape[1] == ld and bias.numel() == ld
    nat.check(nat.lib().ds_source_conv(y.data_ptr(), src.data_ptr(), w_t.data_ptr(), bias.data_ptr(), B, src.shape[1], n_out, ld,
                                       w_t.shape[0], stride, pad, stream()))


def conv_small_weights(w: torch.Tensor) -> torch.Tensor:
    """Conv1d weight [c_out, c_in, k] -> the bf16 shared-memory image [k][c_in / 8][c_out][8] ds_conv_small reads."""
    c_out, c_in, k = w.shape
    return w.detach().float().permute(2, 1, 0).reshape(k, c_in // 8, 8, c_out).permute(0, 1, 3, 2).to(torch.bfloat16).contiguous()


def conv_small(x: torch.Tensor, w_img: torch.Tensor, bias: torch.Tensor, out: torch.Tensor, B: int, T: int, dil: int = 1,
               res: Optional[torch.Tensor] = None, out2: Optional[torch.Tensor] = None, act: int = ACT_NONE, slope: float = 0.0):
    """out = act(conv_k(x) + bias + res) for c_in = c_out in {16, 32} (include/dsb200.h: ds_conv_small); x bf16 [B * T, c_in]."""
    taps, _, c_out, _ = w_img.shape
    nat.check(nat.lib().ds_conv_small(x.data_ptr(), x.shape[1], w_img.data_ptr(), c_out, bias.data_ptr(), _p(res), out.data_ptr(), _p(out2),
                                      B, T, taps, dil, act, slope, stream()))


def mean3_lrelu(y: torch.Tensor, a: torch.Tensor, b: torch.Tensor, c: torch.Tensor, slope: float):
    nat.check(nat.lib().ds_mean3_lrelu(y.data_ptr(), a.data_ptr(), b.data_ptr(), c.data_ptr(), y.numel(), slope, stream()))


def layernorm(y: torch.Tensor, x: torch.Tensor, gamma: torch.Tensor, beta: torch.Tensor, eps: float = 1e-5,
              vec: Optional[torch.Tensor] = None, rows_per_vec: int = 0, add: Optional[torch.Tensor] = None):
    """y = LayerNorm(x + add + vec[row // rows_per_vec]); add bf16 [rows, ch], vec fp32 [n, ch], both optional."""
    rows, ch = x.shape
    nat.check(nat.lib().ds_layernorm(y.data_ptr(), x.data_ptr(), gamma.data_ptr(), beta.data_ptr(), rows, ch, eps, _p(vec), rows_per_vec,
                                     _p(add), stream()))


def dwconv(y: torch.Tensor, x: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor], B: int, T: int, kernel_size: int,
           act: int = DW_NONE, slope: Optional[torch.Tensor] = None):
    nat.check(nat.lib().ds_dwconv(y.data_ptr(), x.data_ptr(), w.data_ptr(), _p(bias), _p(slope), B, T, x.shape[1], kernel_size,
                                  kernel_size // 2, act, stream()))


def cast_bf16(dst: torch.Tensor, src: torch.Tensor):
    """fp32 [rows, cols] -> bf16 [rows, ld >= cols] (zero padded)."""
    rows, cols = src.shape[0], src.shape[1]
    nat.check(nat.lib().ds_cast_bf16(dst.data_ptr(), src.data_ptr(), rows, cols, dst.shape[1], stream()))


def lincomb(dst: torch.Tensor, srcs: Sequence[torch.Tensor], coefs: Sequence[float]):
    ptrs = (C.c_void_p * len(srcs))(*[s.data_ptr() for s in srcs])
    cf = (C.c_float * len(coefs))(*coefs)
    nat.check(nat.lib().ds_lincomb(dst.data_ptr(), ptrs, cf, len(srcs), dst.numel(), stream()))
