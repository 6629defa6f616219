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
from .runtime import run_program, shallow_start, to_time_major


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
        if (src_spec is not None and src_spec.is_cuda and self.num_feats == 1 and src_spec.dim() == 3
                and type(self).norm_spec is RectifiedFlow.norm_spec and self.spec_min.numel() in (1, self.out_dims)):
            # norm_spec -> linear mix with the noise -> time-major as one kernel (runtime.shallow_start)
            return self.denorm_spec(self.inference(cond, b=b, device=device, src_raw=src_spec))
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
    def inference(self, cond, b=1, x_end=None, device=None, src_raw=None):
        """cond [B, H, T]; x_end [B, F, M, T] or None -> [B, T, M] / [B, F, T, M] (reflow.py:104-138).  src_raw: the un-normalised
        shallow-diffusion source [B, T, M] instead of x_end (forward passes it when the start state can be built in one kernel)."""
        hparams = get_hparams()
        T = cond.shape[2]
        noise = self._draw_noise(b, T, device)
        t_start = hparams.get('T_start_infer', self.t_start)
        x_tm = None
        if src_raw is not None and not (self.use_shallow_diffusion and 0. < t_start < 1.):  # no mix to fuse: the plain path
            x_end = self.norm_spec(src_raw).transpose(-2, -1)[:, None, :, :]
            src_raw = None
        if self.use_shallow_diffusion and t_start > 0:
            assert x_end is not None or src_raw is not None, 'Missing shallow diffusion source.'
            if t_start >= 1.:
                t_start = 1.
                x = x_end
            elif src_raw is not None:
                x_tm = shallow_start(src_raw, noise, self.spec_min, self.spec_max, t_start, 1 - t_start)
            else:
                x = t_start * x_end + (1 - t_start) * noise
        else:
            t_start = 0.
            x = noise

        algorithm = hparams['sampling_algorithm']
        infer_step = hparams['sampling_steps']
        if x_tm is None:
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
