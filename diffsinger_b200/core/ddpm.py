"""Drop-in for the reference's ``modules/core/ddpm.py`` sampler classes.

Same constructor signatures, buffers (names / shapes / persistence, ddpm.py:82-115), ``forward``
contract and per-call ``hparams`` reads; ``inference()`` compiles the selected sampler into a
Program (core/programs.py) and runs it through the native session (one CUDA graph per
(sampler config, B, T)).  Training (`infer=False`) goes through the backbone's autograd branch.
"""
from __future__ import annotations

import numpy as np
import torch
from torch import nn

from ..backbones import build_backbone
from ..hparams import get_hparams
from . import programs
from .curves import ClampedCurve, CurveList, RepeatBins
from .runtime import run_program, shallow_start, to_time_major


def linear_beta_schedule(timesteps, max_beta=0.01):
    """reference ddpm.py:28-33 (called without max_beta at :67, so beta_max = 0.01)."""
    return np.linspace(1e-4, max_beta, timesteps)


def cosine_beta_schedule(timesteps, s=0.008):
    """reference ddpm.py:36-46."""
    steps = timesteps + 1
    x = np.linspace(0, steps, steps)
    alphas_cumprod = np.cos(((x / steps) + s) / (1 + s) * np.pi * 0.5) ** 2
    alphas_cumprod = alphas_cumprod / alphas_cumprod[0]
    betas = 1 - (alphas_cumprod[1:] / alphas_cumprod[:-1])
    return np.clip(betas, a_min=0, a_max=0.999)


beta_schedule = {"cosine": cosine_beta_schedule, "linear": linear_beta_schedule}


class GaussianDiffusion(nn.Module):
    ANCESTRAL_BLOCK = 32  # steps per program of the ancestral sampler (one CUDA graph and one ring of noise buffers per block)

    def __init__(self, out_dims, num_feats=1, timesteps=1000, k_step=1000,
                 backbone_type=None, backbone_args=None, betas=None,
                 spec_min=None, spec_max=None):
        super().__init__()
        hparams = get_hparams()
        self.denoise_fn: nn.Module = build_backbone(out_dims, num_feats, backbone_type, backbone_args)
        self.out_dims = out_dims
        self.num_feats = num_feats

        if betas is not None:
            betas = betas.detach().cpu().numpy() if isinstance(betas, torch.Tensor) else betas
        else:
            betas = beta_schedule[hparams['schedule_type']](timesteps)
        alphas = 1. - betas
        alphas_cumprod = np.cumprod(alphas, axis=0)
        alphas_cumprod_prev = np.append(1., alphas_cumprod[:-1])

        self.use_shallow_diffusion = hparams.get('use_shallow_diffusion', False)
        if self.use_shallow_diffusion:
            assert k_step <= timesteps, 'K_step should not be larger than timesteps.'
        self.timesteps = timesteps
        self.k_step = k_step if self.use_shallow_diffusion else timesteps

        def reg(name, arr):
            self.register_buffer(name, torch.tensor(arr, dtype=torch.float32))

        reg('betas', betas)
        reg('alphas_cumprod', alphas_cumprod)
        reg('alphas_cumprod_prev', alphas_cumprod_prev)
        reg('sqrt_alphas_cumprod', np.sqrt(alphas_cumprod))
        reg('sqrt_one_minus_alphas_cumprod', np.sqrt(1. - alphas_cumprod))
        reg('log_one_minus_alphas_cumprod', np.log(1. - alphas_cumprod))
        reg('sqrt_recip_alphas_cumprod', np.sqrt(1. / alphas_cumprod))
        reg('sqrt_recipm1_alphas_cumprod', np.sqrt(1. / alphas_cumprod - 1))
        posterior_variance = betas * (1. - alphas_cumprod_prev) / (1. - alphas_cumprod)
        reg('posterior_variance', posterior_variance)
        reg('posterior_log_variance_clipped', np.log(np.maximum(posterior_variance, 1e-20)))
        reg('posterior_mean_coef1', betas * np.sqrt(alphas_cumprod_prev) / (1. - alphas_cumprod))
        reg('posterior_mean_coef2', (1. - alphas_cumprod_prev) * np.sqrt(alphas) / (1. - alphas_cumprod))

        # spec_min / spec_max: [1, 1, M] or [1, F, 1, M] after the transpose (ddpm.py:104-109)
        spec_min = torch.FloatTensor(spec_min)[None, None, :out_dims].transpose(-3, -2)
        spec_max = torch.FloatTensor(spec_max)[None, None, :out_dims].transpose(-3, -2)
        self.register_buffer('spec_min', spec_min)
        self.register_buffer('spec_max', spec_max)

        self.time_scale_factor = self.timesteps
        self.t_start = 1 - self.k_step / self.timesteps
        factors = torch.LongTensor([i for i in range(1, self.timesteps + 1) if self.timesteps % i == 0])
        self.register_buffer('timestep_factors', factors, persistent=False)
        self._programs = {}

    # ------------------------------------------------------------------ training-side helpers
    def q_sample(self, x_start, t, noise):
        shape = (x_start.shape[0],) + (1,) * (x_start.dim() - 1)
        return (self.sqrt_alphas_cumprod.gather(-1, t).reshape(shape) * x_start
                + self.sqrt_one_minus_alphas_cumprod.gather(-1, t).reshape(shape) * noise)

    def p_losses(self, x_start, t, cond, noise=None):
        if noise is None:
            noise = torch.randn_like(x_start)
        x_noisy = self.q_sample(x_start=x_start, t=t, noise=noise)
        x_recon = self.denoise_fn(x_noisy, t, cond)
        return x_recon, noise

    # ------------------------------------------------------------------ sampling
    def _program(self, algorithm: str, t_max: int, speedup: int):
        key = (algorithm, t_max, speedup)
        prog = self._programs.get(key)
        if prog is None:
            if algorithm == 'ddim':
                prog = programs.ddim_program(self.alphas_cumprod, t_max, speedup)
            elif algorithm == 'pndm':
                prog = programs.pndm_program(self.alphas_cumprod, t_max, speedup)
            elif algorithm == 'dpm-solver':
                prog = programs.dpm_solver_pp_program(self.betas[:t_max], t_max // speedup)
            elif algorithm == 'unipc':
                prog = programs.unipc_program(self.betas[:t_max], t_max // speedup)
            elif algorithm == '__ancestral__':  # block [speedup, t_max) of the ancestral chain (see inference)
                prog = programs.ddpm_program(self, t_max, speedup)
            else:
                raise ValueError(f"Unsupported acceleration algorithm for DDPM: {algorithm}.")
            self._programs[key] = prog
        return prog

    def _draw_noise(self, b, T, device):
        """randn(b, F, M, T) from the global generator like the reference (ddpm.py:227, reflow.py:106) -- unless a caller
        (sharding.sample_sharded) installed the rows of a world-size independent global batch for this call."""
        pre = getattr(self, '_noise_override', None)
        if pre is not None:
            assert tuple(pre.shape) == (b, self.num_feats, self.out_dims, T), (tuple(pre.shape), (b, self.num_feats, self.out_dims, T))
            return pre.to(device)
        return torch.randn(b, self.num_feats, self.out_dims, T, device=device)

    def _host_coef(self, name: str, index: int) -> float:
        """One entry of a schedule buffer as a host float (cached copy: no device sync per call)."""
        buf = getattr(self, name)
        key = (name, buf.data_ptr(), buf._version)
        cache = self.__dict__.setdefault('_coef_cache', {})
        if key not in cache:
            cache.clear()
            cache[key] = buf.detach().cpu()
        return float(cache[key][index])

    def _fusable_start(self, src_spec) -> bool:
        """norm_spec + q_sample + transpose as one kernel (runtime.shallow_start): plain mel spectrograms on the GPU only."""
        return (src_spec is not None and src_spec.is_cuda and self.num_feats == 1 and src_spec.dim() == 3
                and type(self).norm_spec is GaussianDiffusion.norm_spec and self.spec_min.numel() in (1, self.out_dims))

    @torch.no_grad()
    def inference(self, cond, b=1, x_start=None, device=None, src_raw=None):
        """cond [B, H, T]; x_start [B, F, M, T] or None -> [B, T, M] / [B, F, T, M] (ddpm.py:221-351).  src_raw: the un-normalised
        shallow-diffusion source [B, T, M] instead of x_start (forward passes it when the start state can be built in one kernel)."""
        hparams = get_hparams()
        depth = hparams.get('K_step_infer', self.k_step)
        speedup = hparams['diff_speedup']
        if speedup > 0:
            assert depth % speedup == 0, f'Acceleration ratio must be a factor of diffusion depth {depth}.'

        T = cond.shape[2]
        noise = self._draw_noise(b, T, device)
        t_max = min(depth, self.k_step) if self.use_shallow_diffusion else self.k_step

        if src_raw is not None and not 0 < t_max < self.timesteps:  # no mix to fuse: the plain path
            x_start = self.norm_spec(src_raw).transpose(-2, -1)[:, None, :, :]
            src_raw = None
        if t_max >= self.timesteps:
            x = noise
        elif t_max > 0:
            assert x_start is not None or src_raw is not None, 'Missing shallow diffusion source.'
            if src_raw is not None:  # norm_spec -> q_sample -> time-major in one pass
                x = shallow_start(src_raw, noise, self.spec_min, self.spec_max, self._host_coef('sqrt_alphas_cumprod', t_max - 1),
                                  self._host_coef('sqrt_one_minus_alphas_cumprod', t_max - 1))
            else:
                x = self.q_sample(x_start, torch.full((b,), t_max - 1, device=device, dtype=torch.long), noise)
        else:
            assert x_start is not None, 'Missing shallow diffusion source.'
            x = x_start
        as_tm = (lambda v: v) if (src_raw is not None and 0 < t_max < self.timesteps) else to_time_major

        if t_max > 0 and speedup > 1:
            prog = self._program(hparams['diff_accelerator'], t_max, speedup)
            x_tm = run_program(self.denoise_fn, prog, as_tm(x), cond, [], lengths=getattr(self, '_lengths_override', None))
        elif t_max > 0:
            # ancestral chain (ddpm.py:346-349), ANCESTRAL_BLOCK steps per program: the reference holds one noise tensor at a time;
            # pre-drawing all t_max of them (plus their time-major copies and the session buffers) would cost 3 * t_max * B*T*F*M*4
            # bytes.  Fresh noise per step comes from the global generator in noise_like's draw order (ddpm.py:153).
            assert getattr(self, '_lengths_override', None) is None, 'ragged batches are not offered for the ancestral chain'
            x_tm = as_tm(x)
            for hi in range(t_max, 0, -self.ANCESTRAL_BLOCK):
                lo = max(hi - self.ANCESTRAL_BLOCK, 0)
                prog = self._program('__ancestral__', hi, lo)
                extra = [to_time_major(torch.randn(b, self.num_feats, self.out_dims, T, device=device)) for _ in range(hi - lo)]
                x_tm = run_program(self.denoise_fn, prog, x_tm, cond, extra)
                del extra
        else:
            x_tm = as_tm(x)
        # [B, T, F*M] -> [B, T, M] (F == 1) or [B, F, T, M]
        if self.num_feats == 1:
            return x_tm
        return x_tm.reshape(b, T, self.num_feats, self.out_dims).permute(0, 2, 1, 3)

    def forward(self, condition, gt_spec=None, src_spec=None, infer=True):
        """condition [B, T, H] (ddpm.py:353-377)."""
        cond = condition.transpose(1, 2)
        b, device = condition.shape[0], condition.device
        if not infer:
            spec = self.norm_spec(gt_spec).transpose(-2, -1)
            if self.num_feats == 1:
                spec = spec[:, None, :, :]
            t = torch.randint(0, self.k_step, (b,), device=device).long()
            return self.p_losses(spec, t, cond=cond)
        if self._fusable_start(src_spec):
            return self.denorm_spec(self.inference(cond, b=b, device=device, src_raw=src_spec))
        if src_spec is not None:
            spec = self.norm_spec(src_spec).transpose(-2, -1)
            if self.num_feats == 1:
                spec = spec[:, None, :, :]
        else:
            spec = None
        x = self.inference(cond, b=b, x_start=spec, device=device)
        return self.denorm_spec(x)

    def norm_spec(self, x):
        return (x - self.spec_min) / (self.spec_max - self.spec_min) * 2 - 1

    def denorm_spec(self, x):
        return (x + 1) / 2 * (self.spec_max - self.spec_min) + self.spec_min


class RepetitiveDiffusion(RepeatBins, GaussianDiffusion):
    """reference ddpm.py:386-421 -- core/curves.py:RepeatBins over the DDPM sampler."""

    def __init__(self, vmin, vmax, repeat_bins, timesteps=1000, k_step=1000, backbone_type=None, backbone_args=None, betas=None):
        super().__init__(vmin, vmax, repeat_bins, timesteps=timesteps, k_step=k_step,
                         backbone_type=backbone_type, backbone_args=backbone_args, betas=betas)


class PitchDiffusion(ClampedCurve, GaussianDiffusion):
    """reference ddpm.py:424-445 -- core/curves.py:ClampedCurve over the DDPM sampler."""

    def __init__(self, vmin, vmax, cmin, cmax, repeat_bins, timesteps=1000, k_step=1000, backbone_type=None, backbone_args=None,
                 betas=None):
        super().__init__(vmin, vmax, cmin, cmax, repeat_bins, timesteps=timesteps, k_step=k_step,
                         backbone_type=backbone_type, backbone_args=backbone_args, betas=betas)


class MultiVarianceDiffusion(CurveList, GaussianDiffusion):
    """reference ddpm.py:448-505 -- core/curves.py:CurveList over the DDPM sampler."""

    def __init__(self, ranges, clamps, repeat_bins, timesteps=1000, k_step=1000, backbone_type=None, backbone_args=None, betas=None):
        super().__init__(ranges, clamps, repeat_bins, timesteps=timesteps, k_step=k_step,
                         backbone_type=backbone_type, backbone_args=backbone_args, betas=betas)
