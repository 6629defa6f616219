"""Synthetic WaveNet weights for benchmarks and tools (no checkpoint ships with the reference: `checkpoints/.gitkeep`).

Same shapes and initialiser scales as the reference's `WaveNet.__init__` (modules/backbones/wavenet.py:52-73), with
`output_projection.weight` re-randomised -- the reference zero-initialises it (wavenet.py:73), which would make a
random-init network's output independent of its input (SURVEY.md trap T1)."""
from __future__ import annotations

import math
from typing import Dict

import torch


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



def random_lynx_state_dict(in_dims: int, n_feats: int, num_layers: int, num_channels: int, hidden: int, expansion_factor: int = 2,
                           kernel_size: int = 31, activation: str = "PReLU", seed: int = 1234,
                           out_std: float = 0.05) -> Dict[str, torch.Tensor]:
    """Synthetic LYNXNet weights with the reference's shapes and key names (modules/backbones/lynxnet.py:95-131): default
    initialiser scales, LayerNorm affine / PReLU slopes perturbed away from (1, 0, 0.25), and `output_projection.weight`
    re-randomised (zero-initialised in the reference, lynxnet.py:131)."""
    g = torch.Generator().manual_seed(seed)
    C, H, FM, I = num_channels, hidden, in_dims * n_feats, num_channels * expansion_factor

    def uni(shape, fan_in):
        b = 1.0 / math.sqrt(fan_in)
        return (torch.rand(shape, generator=g) * 2 - 1) * b

    def kaiming(shape, fan_in):
        return torch.randn(shape, generator=g) * math.sqrt(2.0 / fan_in)

    sd = {
        "input_projection.weight": kaiming((C, FM, 1), FM), "input_projection.bias": uni((C,), FM),
        "diffusion_embedding.1.weight": uni((4 * C, C), C), "diffusion_embedding.1.bias": uni((4 * C,), C),
        "diffusion_embedding.3.weight": uni((C, 4 * C), 4 * C), "diffusion_embedding.3.bias": uni((C,), 4 * C),
        "norm.weight": 1.0 + 0.1 * torch.randn(C, generator=g), "norm.bias": 0.05 * torch.randn(C, generator=g),
        "output_projection.weight": torch.randn((FM, C, 1), generator=g) * out_std, "output_projection.bias": uni((FM,), C),
    }
    for i in range(num_layers):
        p, q = f"residual_layers.{i}.", f"residual_layers.{i}.convmodule.net."
        sd[p + "diffusion_projection.weight"] = uni((C, C, 1), C)
        sd[p + "diffusion_projection.bias"] = uni((C,), C)
        sd[p + "conditioner_projection.weight"] = uni((C, H, 1), H)
        sd[p + "conditioner_projection.bias"] = uni((C,), H)
        sd[q + "0.weight"] = 1.0 + 0.1 * torch.randn(C, generator=g)
        sd[q + "0.bias"] = 0.05 * torch.randn(C, generator=g)
        sd[q + "2.weight"] = uni((2 * I, C, 1), C)
        sd[q + "2.bias"] = uni((2 * I,), C)
        sd[q + "4.weight"] = uni((I, 1, kernel_size), kernel_size)
        sd[q + "4.bias"] = uni((I,), kernel_size)
        if activation == "PReLU":
            sd[q + "5.weight"] = 0.05 + 0.4 * torch.rand(I, generator=g)
        sd[q + "6.weight"] = uni((C, I, 1), I)
        sd[q + "6.bias"] = uni((C,), I)
    return sd
