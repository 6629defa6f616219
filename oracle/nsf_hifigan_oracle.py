"""TEST INFRASTRUCTURE ONLY -- CPU restatement (torch fp32) of the reference NSF-HiFiGAN generator from a plain state_dict
(key names of ``Generator.state_dict()`` AFTER ``remove_weight_norm()``: ``conv_pre.*``, ``ups.i.*``,
``resblocks.n.convs1/convs2.m.*``, ``noise_convs.i.*``, ``m_source.l_linear.*``, ``conv_post.*``).

Restates ``SineGen._f02sine`` / ``forward`` (modules/nsf_hifigan/models.py:134-168), ``SourceModuleHnNSF.forward`` (:200-203),
``ResBlock1.forward`` (:60-67) and ``Generator.forward`` (:274-298, ``mini_nsf == False``).  The two random draws of SineGen
(initial harmonic phases, :147-148; additive noise, :166) are inputs, so a run is reproducible from the golden file.
Pinned by tests/test_oracle_golden.py against tests/golden/nsf_hifigan_*.npz (outputs of the imported, unmodified reference).
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F

LRELU_SLOPE = 0.1  # models.py:15


@torch.no_grad()
def harmonic_source(sd, f0, upp: int, sampling_rate: float, rand_ini, noise, sine_amp=0.1, noise_std=0.003):
    """f0 [B, T], rand_ini [1, 1, 9], noise [B, T * upp, 9] -> har_source [B, 1, T * upp]."""
    dim = rand_ini.shape[-1]
    f0 = f0.float().unsqueeze(-1)
    rad = f0 / sampling_rate * torch.arange(1, upp + 1)                                  # :138
    rad2 = torch.fmod(rad[..., -1:].float() + 0.5, 1.0) - 0.5                            # :139
    rad_acc = rad2.cumsum(dim=1).fmod(1.0)                                               # :140
    rad = rad + F.pad(rad_acc[:, :-1, :], (0, 0, 1, 0))                                  # :141
    rad = rad.reshape(f0.shape[0], -1, 1) * torch.arange(1, dim + 1).reshape(1, 1, -1)   # :142-143
    sines = torch.sin(2 * math.pi * (rad + rand_ini)) * sine_amp                         # :146-147, 160
    uv = (f0 > 0).float()                                                                # :161
    uv = F.interpolate(uv.transpose(2, 1), scale_factor=upp, mode="nearest").transpose(2, 1)
    noise_amp = uv * noise_std + (1 - uv) * sine_amp / 3                                 # :163
    sines = sines * uv + noise_amp * noise                                               # :164-165
    merged = torch.tanh(F.linear(sines, sd["m_source.l_linear.weight"].float(), sd["m_source.l_linear.bias"].float()))
    return merged.transpose(1, 2)                                                        # :277


@torch.no_grad()
def generator(sd, cfg, x, har_source):
    """x [B, num_mels, T] (ln mel), har_source [B, 1, T * prod(upsample_rates)] -> [B, 1, T * hop]."""
    sd = {k: v.float() for k, v in sd.items()}
    rates, ksz = list(cfg["upsample_rates"]), list(cfg["upsample_kernel_sizes"])
    rk, rd = list(cfg["resblock_kernel_sizes"]), [list(d) for d in cfg["resblock_dilation_sizes"]]
    x = F.conv1d(x.float(), sd["conv_pre.weight"], sd["conv_pre.bias"], padding=3)       # :278
    for i, (u, k) in enumerate(zip(rates, ksz)):
        x = F.leaky_relu(x, LRELU_SLOPE)
        x = F.conv_transpose1d(x, sd[f"ups.{i}.weight"], sd[f"ups.{i}.bias"], stride=u, padding=(k - u) // 2)
        w = sd[f"noise_convs.{i}.weight"]
        stride = 1
        for r in rates[i + 1:]:
            stride *= r
        pad = stride // 2 if i + 1 < len(rates) else 0
        x = x + F.conv1d(har_source.float(), w, sd[f"noise_convs.{i}.bias"], stride=stride, padding=pad)   # :237-243, 287-288
        xs = None
        for j, (kj, dj) in enumerate(zip(rk, rd)):
            p = f"resblocks.{i * len(rk) + j}."
            y = x
            for m, d in enumerate(dj):                                                   # ResBlock1.forward :60-67
                t = F.leaky_relu(y, LRELU_SLOPE)
                t = F.conv1d(t, sd[p + f"convs1.{m}.weight"], sd[p + f"convs1.{m}.bias"], dilation=d, padding=(kj * d - d) // 2)
                t = F.leaky_relu(t, LRELU_SLOPE)
                t = F.conv1d(t, sd[p + f"convs2.{m}.weight"], sd[p + f"convs2.{m}.bias"], padding=(kj - 1) // 2)
                y = t + y
            xs = y if xs is None else xs + y
        x = xs / len(rk)                                                                 # :294
    x = F.leaky_relu(x)                                                                  # default slope 0.01, :295
    return torch.tanh(F.conv1d(x, sd["conv_post.weight"], sd["conv_post.bias"], padding=3))
