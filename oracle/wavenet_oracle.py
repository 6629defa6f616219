"""TEST INFRASTRUCTURE ONLY -- CPU restatement (torch fp32) of the reference WaveNet denoiser.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl reference``
legs may import this module; the product path (``diffsinger_b200``) never does.

Restates, functionally and from a plain ``state_dict`` (reference key names):
  * ``SinusoidalPosEmb.forward``       modules/commons/common_layers.py:273-280
  * ``ResidualBlock.forward``          modules/backbones/wavenet.py:33-48
  * ``WaveNet.forward``                modules/backbones/wavenet.py:75-107
Parity pinning: ``tests/test_oracle_golden.py`` checks this file against the golden vectors in
``tests/golden/`` that ``oracle/make_golden.py`` produced by running the imported, unmodified
reference in the build container (the reference itself ships no tests or golden vectors --
SURVEY.md section 4 -- so the third-party arithmetic boundary, torch ATen, is "parity unpinned"
beyond those generated fixtures).
"""
from __future__ import annotations

import math
from typing import Dict

import torch
import torch.nn.functional as F


def sinusoidal_pos_emb(t: torch.Tensor, dim: int) -> torch.Tensor:
    """common_layers.py:273-280 -- t [B] (long or float) -> [B, dim] fp32."""
    half = dim // 2
    scale = math.log(10000) / (half - 1)
    freq = torch.exp(torch.arange(half, device=t.device) * -scale)
    arg = t[:, None] * freq[None, :]
    return torch.cat((arg.sin(), arg.cos()), dim=-1)


def step_embedding(sd: Dict[str, torch.Tensor], t: torch.Tensor, C: int) -> torch.Tensor:
    """wavenet.py:89-90 -- sinusoid -> Linear(C,4C) -> Mish -> Linear(4C,C)."""
    e = sinusoidal_pos_emb(t, C)
    e = F.linear(e, sd["mlp.0.weight"], sd["mlp.0.bias"])
    e = F.mish(e)
    return F.linear(e, sd["mlp.2.weight"], sd["mlp.2.bias"])


def residual_block(sd, i: int, x, cond, d_emb, dilation: int):
    """wavenet.py:33-48 -- returns ((x + residual)/sqrt(2), skip)."""
    p = f"residual_layers.{i}."
    C = x.shape[1]
    d = F.linear(d_emb, sd[p + "diffusion_projection.weight"], sd[p + "diffusion_projection.bias"])
    c = F.conv1d(cond, sd[p + "conditioner_projection.weight"], sd[p + "conditioner_projection.bias"])
    y = x + d.unsqueeze(-1)  # added BEFORE the zero padding of the dilated conv (trap T3)
    y = F.conv1d(y, sd[p + "dilated_conv.weight"], sd[p + "dilated_conv.bias"],
                 padding=dilation, dilation=dilation) + c
    gate, filt = y[:, :C], y[:, C:]
    y = torch.sigmoid(gate) * torch.tanh(filt)
    y = F.conv1d(y, sd[p + "output_projection.weight"], sd[p + "output_projection.bias"])
    res, skip = y[:, :C], y[:, C:]
    return (x + res) / math.sqrt(2.0), skip


def count_layers(sd) -> int:
    n = 0
    while f"residual_layers.{n}.dilated_conv.weight" in sd:
        n += 1
    return n


def wavenet_forward(sd: Dict[str, torch.Tensor], spec: torch.Tensor, diffusion_step: torch.Tensor,
                    cond: torch.Tensor, *, n_feats: int, dilation_cycle_length: int) -> torch.Tensor:
    """wavenet.py:75-107 -- spec [B,F,M,T], diffusion_step [B] or [1], cond [B,H,T] -> [B,F,M,T]."""
    L = count_layers(sd)
    C = sd["skip_projection.weight"].shape[0]
    in_dims = spec.shape[2]
    x = spec.squeeze(1) if n_feats == 1 else spec.flatten(1, 2)
    x = F.relu(F.conv1d(x, sd["input_projection.weight"], sd["input_projection.bias"]))
    d_emb = step_embedding(sd, diffusion_step, C)
    skip_sum = None
    for i in range(L):
        x, skip = residual_block(sd, i, x, cond, d_emb, 2 ** (i % dilation_cycle_length))
        skip_sum = skip if skip_sum is None else skip_sum + skip
    x = skip_sum / math.sqrt(L)
    x = F.relu(F.conv1d(x, sd["skip_projection.weight"], sd["skip_projection.bias"]))
    x = F.conv1d(x, sd["output_projection.weight"], sd["output_projection.bias"])
    if n_feats == 1:
        return x[:, None, :, :]
    return x.reshape(-1, n_feats, in_dims, x.shape[2])


class OracleWaveNet:
    """Callable with the reference backbone signature ``(spec, diffusion_step, cond)``."""

    def __init__(self, state_dict, n_feats: int, dilation_cycle_length: int):
        self.sd = {k: v.detach().float().cpu() for k, v in state_dict.items()}
        self.n_feats = n_feats
        self.cycle = dilation_cycle_length
        self.calls = 0

    @torch.no_grad()
    def __call__(self, spec, diffusion_step, cond=None):
        self.calls += 1
        return wavenet_forward(self.sd, spec, diffusion_step, cond,
                               n_feats=self.n_feats, dilation_cycle_length=self.cycle)


def random_state_dict(in_dims: int, n_feats: int, num_layers: int, num_channels: int, hidden: int,
                      seed: int = 1234, out_std: float = 0.05) -> Dict[str, torch.Tensor]:
    """Synthetic weights with the reference's shapes (wavenet.py:52-73).

    Scales follow the reference's default initialisers (kaiming-normal for the three ``Conv1d``
    subclasses, U(+-1/sqrt(fan_in)) for nn.Conv1d / nn.Linear) except ``output_projection.weight``,
    which the reference zero-initialises (wavenet.py:73) and which is re-randomised here so that
    parity checks are not vacuous (SURVEY.md trap T1).
    """
    g = torch.Generator().manual_seed(seed)
    C, H, FM = num_channels, hidden, in_dims * n_feats

    def uni(shape, fan_in):
        b = 1.0 / math.sqrt(fan_in)
        return (torch.rand(shape, generator=g) * 2 - 1) * b

    def kaiming(shape, fan_in):
        return torch.randn(shape, generator=g) * math.sqrt(2.0 / fan_in)

    sd = {
        "input_projection.weight": kaiming((C, FM, 1), FM), "input_projection.bias": uni((C,), FM),
        "mlp.0.weight": uni((4 * C, C), C), "mlp.0.bias": uni((4 * C,), C),
        "mlp.2.weight": uni((C, 4 * C), 4 * C), "mlp.2.bias": uni((C,), 4 * C),
        "skip_projection.weight": kaiming((C, C, 1), C), "skip_projection.bias": uni((C,), C),
        "output_projection.weight": torch.randn((FM, C, 1), generator=g) * out_std,
        "output_projection.bias": uni((FM,), C),
    }
    for i in range(num_layers):
        p = f"residual_layers.{i}."
        sd[p + "dilated_conv.weight"] = uni((2 * C, C, 3), 3 * C)
        sd[p + "dilated_conv.bias"] = uni((2 * C,), 3 * C)
        sd[p + "diffusion_projection.weight"] = uni((C, C), C)
        sd[p + "diffusion_projection.bias"] = uni((C,), C)
        sd[p + "conditioner_projection.weight"] = uni((2 * C, H, 1), H)
        sd[p + "conditioner_projection.bias"] = uni((2 * C,), H)
        sd[p + "output_projection.weight"] = uni((2 * C, C, 1), C)
        sd[p + "output_projection.bias"] = uni((2 * C,), C)
    return sd
