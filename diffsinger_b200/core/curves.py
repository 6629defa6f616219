"""Curve adaptors shared by the DDPM and rectified-flow variance samplers.

The variance model predicts scalar curves per frame (pitch deviation, energy, breathiness ...).  The reference feeds them to
the same samplers as a mel spectrogram by repeating every scalar over `repeat_bins` pseudo-bins on the way in and averaging
the bins on the way out (modules/core/ddpm.py:386-505 and modules/core/reflow.py:147-261 state this twice, once per sampler
family).  Here the three behaviours are mixins over ANY sampler that offers `norm_spec` / `denorm_spec` and takes
`out_dims / num_feats / spec_min / spec_max` keywords; core/ddpm.py and core/reflow.py compose them with their sampler, so
each rule exists once.

The composed classes spell out the reference's constructor signatures (names and positional order: the reference's
`filter_kwargs` inspects them, utils/__init__.py:149-163) and forward the sampler-specific arguments as keywords.
"""
from __future__ import annotations

from numbers import Real
from typing import List, Optional, Tuple

import torch

Clamp = Optional[Tuple[Optional[float], Optional[float]]]


def _bounds(vmin, vmax):
    """scalar or per-curve limits -> (num_feats, spec_min, spec_max) in the nesting the samplers' buffers expect"""
    if isinstance(vmin, Real) and isinstance(vmax, Real):
        return 1, [vmin], [vmax]
    if isinstance(vmin, Real) or isinstance(vmax, Real) or len(vmin) != len(vmax):
        raise AssertionError("vmin and vmax must both be scalars or sequences of equal length")
    return len(vmin), [[v] for v in vmin], [[v] for v in vmax]


class RepeatBins:
    """k curves [B, T] / [B, k, T]  <->  k pseudo-spectrograms of `repeat_bins` identical bins."""

    def __init__(self, vmin, vmax, repeat_bins: int, **kw):
        num_feats, lo, hi = _bounds(vmin, vmax)
        self.repeat_bins = repeat_bins  # plain attribute: legal before nn.Module.__init__
        super().__init__(out_dims=repeat_bins, num_feats=num_feats, spec_min=lo, spec_max=hi, **kw)

    def norm_spec(self, x):
        tiled = x.unsqueeze(-1).expand(*x.shape, self.repeat_bins)
        return super().norm_spec(tiled.contiguous())

    def denorm_spec(self, x):
        return super().denorm_spec(x).mean(dim=-1)


class ClampedCurve(RepeatBins):
    """One curve, clipped to [cmin, cmax] before normalisation and after denormalisation (the pitch predictor)."""

    def __init__(self, vmin: float, vmax: float, cmin: float, cmax: float, repeat_bins: int, **kw):
        self.vmin, self.vmax, self.cmin, self.cmax = vmin, vmax, cmin, cmax
        super().__init__(vmin, vmax, repeat_bins, **kw)

    def norm_spec(self, x):
        return super().norm_spec(x.clamp(min=self.cmin, max=self.cmax))

    def denorm_spec(self, x):
        return super().denorm_spec(x).clamp(min=self.cmin, max=self.cmax)


class CurveList(RepeatBins):
    """A LIST of curves with per-curve ranges and optional clamps (the multi-variance predictor)."""

    def __init__(self, ranges: List[Tuple[float, float]], clamps: List[Clamp], repeat_bins: int, **kw):
        if len(ranges) != len(clamps):
            raise AssertionError("one clamp entry (or None) per range")
        self.clamps = clamps
        lo, hi = [r[0] for r in ranges], [r[1] for r in ranges]
        if len(ranges) == 1:
            lo, hi = lo[0], hi[0]
        super().__init__(lo, hi, repeat_bins, **kw)

    def clamp_spec(self, xs):
        return [x if c is None else x.clamp(min=c[0], max=c[1]) for x, c in zip(xs, self.clamps)]

    def norm_spec(self, xs):
        if len(xs) != self.num_feats:
            raise AssertionError(f"expected {self.num_feats} curves, got {len(xs)}")
        xs = self.clamp_spec(xs)
        return super().norm_spec(xs[0] if self.num_feats == 1 else torch.stack(xs, dim=1))

    def denorm_spec(self, xs):
        curves = super().denorm_spec(xs)
        return self.clamp_spec([curves] if self.num_feats == 1 else list(curves.unbind(dim=1)))
