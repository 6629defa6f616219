"""Training forward / backward of the WaveNet residual stack on the native kernels.

Reference: ``GaussianDiffusion.p_losses`` (modules/core/ddpm.py:212-219) and ``RectifiedFlow.p_losses``
(modules/core/reflow.py:36-41) call the backbone once under autograd with a per-sample step ``t[B]``; 99 % of that
evaluation's FLOPs are the L residual blocks (modules/backbones/wavenet.py:33-48).  ``WaveNetStack`` is a
``torch.autograd.Function`` over exactly those blocks, time-major ``[B*T, channels]`` bf16 like the sampling path:

forward, per layer l (dilation dl):
    z_l  = [W1_l taps | Wc_l] x [xt_l(t-dl), xt_l(t), xt_l(t+dl), cond] + b1 + bc     ds_gemm EPI_LINEAR, kept for backward
    y_l  = sigmoid(z_gate) * tanh(z_filter)                                             ds_gate_fwd
    xt_{l+1} = (xt_l + W2res_l y_l) / sqrt(2) + rvec_l[b]                               ds_gemm EPI_RES
  with xt_l = x_l + d_l[b] (the step vector is added before the conv's zero padding, wavenet.py:36) and
  rvec_l = (b2res_l - d_l) / sqrt(2) + d_{l+1};  skip_sum = [y_0 | ... | y_{L-1}] x Wskip^T + sum_l b2skip_l (one GEMM, K = L*C).

backward, per layer in reverse, with h_l = dLoss/dx_l / sqrt(2) (all layers kept: the weight gradients read them):
    dy   = W2_l^T [h_{l+1} | dskip];  dz_l = gate'(z_l) * dy                             ds_gemm, gate-backward epilogue
    h_l  = (sum_tap W1_tap^T dz_l(t - (tap-1) dl) + h_{l+1}) / sqrt(2)                   ds_gemm EPI_RES (taps mirrored)
  then dcond = [Wc_0^T | ... ] x [dz_0 | ... | dz_{L-1}] (one GEMM, K = 2LC).

  weight gradients: dW1_l[:, :, tap] = dz_l^T xt_l(t + (tap-1) dl), dWc_l = dz_l^T cond, dW2_l = [h_{l+1} | dskip]^T y_l --
  frame reductions with both operands MN-major (ds_wgrad: the tap shift and the utterance edges are TMA coordinates),
  one launch for all layers, taps and weight tensors (two when C and H need different tile widths).

Every GEMM and the gate passes are this repo's tcgen05 / streaming kernels; torch supplies memory, the weight-layout
permutations and the O(channels) bias / step-vector reductions.  Head (input projection), step-embedding MLP and tail
(skip / output projections, < 1 % of the FLOPs) stay ordinary torch modules around the Function.
"""
from __future__ import annotations

import math
from typing import List

import torch

from . import _native as nat
from . import ops

_bn = ops.tile_n
_RS2 = 0.7071067811865476


def _bf(t: torch.Tensor) -> torch.Tensor:
    return t.to(torch.bfloat16).contiguous()


class WaveNetStack(torch.autograd.Function):
    """skip_sum[B, T, C] = sum_l skip_l(x0, cond, dvecs; weights) -- the L residual blocks of wavenet.py:91-96 before / sqrt(L).

    Arguments (all on one CUDA device):
        x0    [B, T, C]   fp32, relu(input_projection(spec)) time-major
        cond  [B, T, H]   fp32
        dvecs [L, B, C]   fp32, diffusion_projection_l(mlp(emb(t)))                      wavenet.py:35-36
        w1 [L, 2C, C, 3], b1 [L, 2C], wc [L, 2C, H], bc [L, 2C], w2 [L, 2C, C], b2 [L, 2C]  stacked layer parameters
        dilations: list of L ints
    """

    @staticmethod
    @torch.amp.custom_fwd(device_type='cuda', cast_inputs=torch.float32)  # under the trainer's autocast: fp32 in, own bf16 inside
    def forward(ctx, x0, cond, dvecs, w1, b1, wc, bc, w2, b2, dilations: List[int]):
        nat.require_device()
        B, T, Cc = x0.shape
        H, L = cond.shape[2], w1.shape[0]
        rows, dev, bf = B * T, x0.device, torch.bfloat16
        if Cc % 64 or H % 64:
            raise nat.NativeError(f"num_channels={Cc} and hidden_size={H} must be multiples of 64")
        # per-step weight layouts (weights change every optimiser step; 8C^2 + 2HC elements per layer)
        wcat = _bf(torch.cat([w1.permute(0, 1, 3, 2).reshape(L, 2 * Cc, 3 * Cc), wc], dim=2))   # [L][2C][3C + H]
        w2res = _bf(w2[:, :Cc])                                                                    # [L][C][C]
        wskip = _bf(w2[:, Cc:].permute(1, 0, 2).reshape(Cc, L * Cc))                               # [C][L*C]
        bias1 = (b1 + bc).float().contiguous()
        bskip = b2[:, Cc:].sum(0).float().contiguous()
        dv = dvecs.float()
        rvec = ((b2[:, None, :Cc] - dv) * _RS2)[:-1] + dv[1:] if L > 1 else None                  # [L-1][B][C]
        if rvec is not None:
            rvec = rvec.contiguous()
        cond_bf = torch.empty(rows, H, dtype=bf, device=dev)
        ops.cast_bf16(cond_bf, cond.detach().float().contiguous().view(rows, H))
        XT = torch.empty(L, rows, Cc, dtype=bf, device=dev)
        Z = torch.empty(rows, L * 2 * Cc, dtype=bf, device=dev)
        Y = torch.empty(rows, L * Cc, dtype=bf, device=dev)
        XT[0] = (x0.detach() + dv[0][:, None, :]).reshape(rows, Cc)
        for l in range(L):
            ops.gemm(ops.EPI_LINEAR, _bn(2 * Cc), B, T, XT[l], wcat[l], Z, k_per_tap=Cc, n_taps=3, dil=dilations[l], a1=cond_bf,
                     bias=bias1[l], out_c0=l * 2 * Cc)
            ops.gate_fwd(Y, l * Cc, Z, l * 2 * Cc, Cc)
            if l + 1 < L:  # the last layer's residual half is never used (wavenet.py:92-94)
                ops.gemm(ops.EPI_RES, _bn(Cc), B, T, Y, w2res[l], XT[l + 1], k_per_tap=Cc, a_c0=l * Cc, vec=rvec[l], vec_bstride=Cc,
                         res=XT[l], scale=_RS2)
        S = torch.empty(rows, Cc, dtype=bf, device=dev)
        ops.gemm(ops.EPI_LINEAR, _bn(Cc), B, T, Y, wskip, S, bias=bskip)
        ctx.save_for_backward(XT, Z, Y, cond_bf, w1, wc, w2)
        ctx.dilations, ctx.shape = list(dilations), (B, T, Cc, H, L)
        return S.float().view(B, T, Cc)

    @staticmethod
    @torch.amp.custom_bwd(device_type='cuda')
    def backward(ctx, dS):
        XT, Z, Y, cond_bf, w1, wc, w2 = ctx.saved_tensors
        B, T, Cc, H, L = ctx.shape
        dil = ctx.dilations
        rows, dev, bf = B * T, dS.device, torch.bfloat16
        # data-gradient weight layouts
        w2t = _bf(w2.transpose(1, 2))                                                             # [L][C][2C] = [W2res^T | W2skip^T]
        w1t = _bf(w1.flip(-1).permute(0, 2, 3, 1).reshape(L, Cc, 6 * Cc))                         # [L][C][3 taps mirrored x 2C]
        wct = _bf(wc.permute(2, 0, 1).reshape(H, L * 2 * Cc))                                     # [H][L*2C]
        zero_c = torch.zeros(max(2 * Cc, H), dtype=torch.float32, device=dev)
        dS_bf = torch.empty(rows, Cc, dtype=bf, device=dev)
        ops.cast_bf16(dS_bf, dS.float().contiguous().view(rows, Cc))
        HALL = torch.empty(L + 1, rows, Cc, dtype=bf, device=dev)                                 # h_l; h_L = 0 (last residual unused)
        HALL[L].zero_()
        DZ = torch.empty(rows, L * 2 * Cc, dtype=bf, device=dev)
        for l in range(L - 1, -1, -1):
            # dy = W2^T [h_{l+1} | dskip] with the gate's derivative in the epilogue: dz_l straight from the accumulators
            ops.gemm(ops.EPI_GATEBWD, _bn(Cc), B, T, HALL[l + 1], w2t[l], DZ, a1=dS_bf, res=Z, res_c0=l * 2 * Cc, n_real=Cc,
                     out_c0=l * 2 * Cc)
            ops.gemm(ops.EPI_RES, _bn(Cc), B, T, DZ, w1t[l], HALL[l], k_per_tap=2 * Cc, n_taps=3, dil=dil[l], a_c0=l * 2 * Cc,
                     vec=zero_c, vec_bstride=0, res=HALL[l + 1], scale=_RS2)
        dcond = torch.empty(rows, H, dtype=bf, device=dev)
        ops.gemm(ops.EPI_LINEAR, _bn(H), B, T, DZ, wct, dcond, bias=zero_c)
        # ---- weight gradients: frame reductions dz^T x with the tap shift as a TMA coordinate (ds_wgrad), fp32 out
        n1, nc, n2 = L * 3 * 2 * Cc * Cc, L * 2 * Cc * H, L * 2 * Cc * Cc
        dw_all = torch.empty(n1 + nc + n2, dtype=torch.float32, device=dev)
        dw1 = dw_all[:n1].view(L, 3, 2 * Cc, Cc)                                                  # tap-major
        dwc = dw_all[n1:n1 + nc].view(L, 2 * Cc, H)
        dw2 = dw_all[n1 + nc:].view(L, 2 * Cc, Cc)
        pairs = [(DZ, XT.view(L * rows, Cc)), (DZ, cond_bf), (HALL.view((L + 1) * rows, Cc), Y), (dS_bf, Y)]
        for bn, table in _wgrad_tables(B, T, Cc, H, L, tuple(dil), dev):
            ops.wgrad(dw_all, pairs, T, B, bn, table)
        # ---- bias / step-vector gradients: O(channels) reductions
        hsum = torch.sum(HALL.view(L + 1, B, T, Cc), 2, dtype=torch.float32)                      # sum_t h_l per utterance
        db1 = torch.sum(DZ, 0, dtype=torch.float32).view(L, 2 * Cc)
        db2 = torch.empty(L, 2 * Cc, dtype=torch.float32, device=dev)
        db2[:, :Cc] = hsum[1:].sum(1)
        db2[:, Cc:] = torch.sum(dS_bf, 0, dtype=torch.float32)
        # d_l only enters through the conv input: sum_t (sqrt(2) h_l - h_{l+1})
        ddv = math.sqrt(2.0) * hsum[:L] - hsum[1:]
        dx0 = HALL[0].view(B, T, Cc).float() * math.sqrt(2.0)
        return dx0, dcond.float().view(B, T, H), ddv, dw1.permute(0, 2, 3, 1), db1, dwc, db1.clone(), dw2, db2, None


_WGRAD_TABLES = {}


def _wgrad_tables(B, T, Cc, H, L, dil, dev):
    """[(BN, device job table)] of one backward pass (ops.wgrad_jobs), cached per shape: every tile of dW1 (operand pair 0),
    dWc (1), dW2 residual half (2) and skip half (3) with offsets into one flat fp32 buffer [dW1 | dWc | dW2] -- one launch
    when C and H share a tile width.  Layer-major order: the tiles of one layer share their operand columns and run
    concurrently (L2 reuse)."""
    key = (B, T, Cc, H, L, dil, str(dev))
    tabs = _WGRAD_TABLES.get(key)
    if tabs is not None:
        return tabs
    if len(_WGRAD_TABLES) > 64:
        _WGRAD_TABLES.clear()
    bn_c, bn_h = _bn(Cc), _bn(H)
    o_c, o_2 = L * 3 * 2 * Cc * Cc, L * 3 * 2 * Cc * Cc + L * 2 * Cc * H
    jobs = {bn_c: [], bn_h: []}
    for l in range(L):
        for tap in range(3):
            for m0 in range(0, 2 * Cc, 128):
                for n0 in range(0, Cc, bn_c):
                    jobs[bn_c].append((l * 2 * Cc + m0, 0, n0, l * B, (tap - 1) * dil[l], min(128, 2 * Cc - m0), min(bn_c, Cc - n0), Cc,
                                       ((l * 3 + tap) * 2 * Cc + m0) * Cc + n0, 0))
        for m0 in range(0, 2 * Cc, 128):
            for n0 in range(0, H, bn_h):
                jobs[bn_h].append((l * 2 * Cc + m0, 0, n0, 0, 0, min(128, 2 * Cc - m0), min(bn_h, H - n0), H,
                                   o_c + (l * 2 * Cc + m0) * H + n0, 1))
        for m0 in range(0, Cc, 128):
            for n0 in range(0, Cc, bn_c):
                mv, nv = min(128, Cc - m0), min(bn_c, Cc - n0)
                # residual half: rows [0, C) of dW2_l = h_{l+1}^T y_l;  skip half: rows [C, 2C) = dskip^T y_l
                jobs[bn_c].append((m0, (l + 1) * B, l * Cc + n0, 0, 0, mv, nv, Cc, o_2 + (l * 2 * Cc + m0) * Cc + n0, 2))
                jobs[bn_c].append((m0, 0, l * Cc + n0, 0, 0, mv, nv, Cc, o_2 + (l * 2 * Cc + Cc + m0) * Cc + n0, 3))
    tabs = _WGRAD_TABLES[key] = [(bn, ops.wgrad_jobs(j, dev)) for bn, j in jobs.items() if j]
    return tabs


def wavenet_forward_train(module, spec, diffusion_step, cond):
    """``WaveNet.forward`` under autograd on a CUDA device (wavenet.py:75-107): torch head / tail around ``WaveNetStack``."""
    import torch.nn.functional as F
    x = spec.squeeze(1) if module.n_feats == 1 else spec.flatten(start_dim=1, end_dim=2)
    B = x.shape[0]
    x0 = F.relu(module.input_projection(x)).transpose(1, 2)                                       # [B, T, C]
    d = module.mlp(module.diffusion_embedding(diffusion_step))                                    # [B or 1, C]
    if d.shape[0] != B:
        d = d.expand(B, -1)
    layers = module.residual_layers
    st = lambda f: torch.stack([f(l) for l in layers])
    # all L diffusion projections as one batched product (wavenet.py:35): [L, C, C] x [B, C] -> [L, B, C]
    dvecs = torch.baddbmm(st(lambda l: l.diffusion_projection.bias)[:, None, :], d.expand(len(layers), -1, -1),
                          st(lambda l: l.diffusion_projection.weight).transpose(1, 2))
    skip = WaveNetStack.apply(
        x0, cond.transpose(1, 2), dvecs,
        st(lambda l: l.dilated_conv.weight), st(lambda l: l.dilated_conv.bias),
        st(lambda l: l.conditioner_projection.weight[:, :, 0]), st(lambda l: l.conditioner_projection.bias),
        st(lambda l: l.output_projection.weight[:, :, 0]), st(lambda l: l.output_projection.bias),
        [l.dilation for l in layers])
    x = skip.transpose(1, 2) / math.sqrt(len(layers))
    x = module.output_projection(F.relu(module.skip_projection(x)))
    if module.n_feats == 1:
        return x[:, None, :, :]
    return x.reshape(-1, module.n_feats, module.in_dims, x.shape[2])
