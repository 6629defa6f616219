"""Drop-in for the reference's ``modules/core/reflow.py`` (RectifiedFlow and its variance wrappers).

Same constructor signatures, non-persistent ``spec_min``/``spec_max`` buffers (reflow.py:33-34),
``forward`` contract and per-call ``hparams`` reads; the ODE loop (:123-136) is compiled into a Program
and run by the native session.
"""
from __future__ import annotations

import torch
import torch.nn as nn

from ..backbones import build_backbone
from ..hparams import get_hparams
from . import programs
from .curves import ClampedCurve, CurveList, RepeatBins
from .runtime import run_program, to_time_major


class RectifiedFlow(nn.Module):
    def __init__(self, out_dims, num_feats=1, t_start=0., time_scale_factor=1000,
                 backbone_type=None, backbone_args=None,
                 spec_min=None, spec_max=None):
        super().__init__()
        hparams = get_hparams()
        self.velocity_fn: nn.Module = build_backbone(out_dims, num_feats, backbone_type, backbone_args)
        self.out_dims = out_dims
        self.num_feats = num_feats
        self.use_shallow_diffusion = hparams.get('use_shallow_diffusion', False)
        if self.use_shallow_diffusion:
            assert 0. <= t_start <= 1., 'T_start should be in [0, 1].'
        else:
            t_start = 0.
        self.t_start = t_start
        self.time_scale_factor = time_scale_factor
        spec_min = torch.FloatTensor(spec_min)[None, None, :out_dims].transpose(-3, -2)
        spec_max = torch.FloatTensor(spec_max)[None, None, :out_dims].transpose(-3, -2)
        self.register_buffer('spec_min', spec_min, persistent=False)
        self.register_buffer('spec_max', spec_max, persistent=False)
        self._programs = {}

    def p_losses(self, x_end, t, cond):
        x_start = torch.randn_like(x_end)
        x_t = x_start + t[:, None, None, None] * (x_end - x_start)
        v_pred = self.velocity_fn(x_t, t * self.time_scale_factor, cond)
        return v_pred, x_end - x_start

    def forward(self, condition, gt_spec=None, src_spec=None, infer=True):
        cond = condition.transpose(1, 2)
        b, device = condition.shape[0], condition.device
        if not infer:
            spec = self.norm_spec(gt_spec).transpose(-2, -1)
            if self.num_feats == 1:
                spec = spec[:, None, :, :]
            t = self.t_start + (1.0 - self.t_start) * torch.rand((b,), device=device)
            v_pred, v_gt = self.p_losses(spec, t, cond=cond)
            return v_pred, v_gt, t
        if src_spec is not None:
            spec = self.norm_spec(src_spec).transpose(-2, -1)
            if self.num_feats == 1:
                spec = spec[:, None, :, :]
        else:
            spec = None
        x = self.inference(cond, b=b, x_end=spec, device=device)
        return self.denorm_spec(x)

    def _draw_noise(self, b, T, device):
        """randn(b, F, M, T) from the global generator like the reference (ddpm.py:227, reflow.py:106) -- unless a caller
        (sharding.sample_sharded) installed the rows of a world-size independent global batch for this call."""
        pre = getattr(self, '_noise_override', None)
        if pre is not None:
            assert tuple(pre.shape) == (b, self.num_feats, self.out_dims, T), (tuple(pre.shape), (b, self.num_feats, self.out_dims, T))
            return pre.to(device)
        return torch.randn(b, self.num_feats, self.out_dims, T, device=device)

    @torch.no_grad()
    def inference(self, cond, b=1, x_end=None, device=None):
        """cond [B, H, T]; x_end [B, F, M, T] or None -> [B, T, M] / [B, F, T, M] (reflow.py:104-138)."""
        hparams = get_hparams()
        T = cond.shape[2]
        noise = self._draw_noise(b, T, device)
        t_start = hparams.get('T_start_infer', self.t_start)
        if self.use_shallow_diffusion and t_start > 0:
            assert x_end is not None, 'Missing shallow diffusion source.'
            if t_start >= 1.:
                t_start = 1.
                x = x_end
            else:
                x = t_start * x_end + (1 - t_start) * noise
        else:
            t_start = 0.
            x = noise

        algorithm = hparams['sampling_algorithm']
        infer_step = hparams['sampling_steps']
        x_tm = to_time_major(x)
        if t_start < 1:
            key = (algorithm, infer_step, float(t_start), float(self.time_scale_factor))
            prog = self._programs.get(key)
            if prog is None:
                prog = self._programs[key] = programs.reflow_program(algorithm, infer_step, float(t_start),
                                                                     float(self.time_scale_factor))
            if infer_step > 0:
                x_tm = run_program(self.velocity_fn, prog, x_tm, cond, lengths=getattr(self, '_lengths_override', None))
        if self.num_feats == 1:
            return x_tm
        return x_tm.reshape(b, T, self.num_feats, self.out_dims).permute(0, 2, 1, 3)

    def norm_spec(self, x):
        return (x - self.spec_min) / (self.spec_max - self.spec_min) * 2 - 1

    def denorm_spec(self, x):
        return (x + 1) / 2 * (self.spec_max - self.spec_min) + self.spec_min


class RepetitiveRectifiedFlow(RepeatBins, RectifiedFlow):
    """reference reflow.py:147-181 -- core/curves.py:RepeatBins over the rectified-flow sampler."""

    def __init__(self, vmin, vmax, repeat_bins, time_scale_factor=1000, backbone_type=None, backbone_args=None):
        super().__init__(vmin, vmax, repeat_bins, time_scale_factor=time_scale_factor,
                         backbone_type=backbone_type, backbone_args=backbone_args)


class PitchRectifiedFlow(ClampedCurve, RectifiedFlow):
    """reference reflow.py:184-203 -- core/curves.py:ClampedCurve over the rectified-flow sampler."""

    def __init__(self, vmin, vmax, cmin, cmax, repeat_bins, time_scale_factor=1000, backbone_type=None, backbone_args=None):
        super().__init__(vmin, vmax, cmin, cmax, repeat_bins, time_scale_factor=time_scale_factor,
                         backbone_type=backbone_type, backbone_args=backbone_args)


class MultiVarianceRectifiedFlow(CurveList, RectifiedFlow):
    """reference reflow.py:206-261 -- core/curves.py:CurveList over the rectified-flow sampler."""

    def __init__(self, ranges, clamps, repeat_bins, time_scale_factor=1000, backbone_type=None, backbone_args=None):
        super().__init__(ranges, clamps, repeat_bins, time_scale_factor=time_scale_factor,
                         backbone_type=backbone_type, backbone_args=backbone_args)
