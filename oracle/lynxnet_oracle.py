"""TEST INFRASTRUCTURE ONLY -- CPU restatement (torch fp32) of the reference LYNXNet denoiser from a plain state_dict.

Only ``tests/`` may import this module; the product path never does.  Restates
  * ``LYNXConvModule``          modules/backbones/lynxnet.py:29-70   (LayerNorm -> 1x1 conv -> SwiGLU -> depthwise k conv ->
                                                                       activation -> 1x1 conv)
  * ``LYNXNetResidualLayer``    lynxnet.py:73-93
  * ``LYNXNet.forward``         lynxnet.py:133-172
  * ``SwiGLU``                  modules/commons/common_layers.py  (out, gate = chunk(2); out * silu(gate))
Pinned by tests/test_oracle_golden.py against tests/golden/lynxnet_*.npz, which oracle/make_golden.py produced by running
the imported, unmodified reference in the build container.
"""
from __future__ import annotations

import math
from typing import Dict

import torch
import torch.nn.functional as F

from oracle.wavenet_oracle import sinusoidal_pos_emb


def lynx_layer(sd, i: int, x, cond, d_emb, kernel_size: int, activation: str, front_cond_inject: bool):
    """lynxnet.py:81-93 -- x [B, C, T], cond [B, H, T], d_emb [B or 1, C, 1]."""
    p = f"residual_layers.{i}."
    c = F.conv1d(cond, sd[p + "conditioner_projection.weight"], sd[p + "conditioner_projection.bias"])
    if front_cond_inject:
        x = x + c
        res_x = x
    else:
        res_x = x
        x = x + c
    x = x + F.conv1d(d_emb, sd[p + "diffusion_projection.weight"], sd[p + "diffusion_projection.bias"])
    q = p + "convmodule.net."
    C = x.shape[1]
    y = F.layer_norm(x.transpose(1, 2), (C,), sd[q + "0.weight"], sd[q + "0.bias"]).transpose(1, 2)
    y = F.conv1d(y, sd[q + "2.weight"], sd[q + "2.bias"])
    out, gate = y.chunk(2, dim=1)
    y = out * F.silu(gate)
    inner = y.shape[1]
    pad = kernel_size // 2  # calc_same_padding(odd k) = (k // 2, k // 2)
    y = F.conv1d(y, sd[q + "4.weight"], sd[q + "4.bias"], padding=pad, groups=inner)
    if activation == "PReLU":
        y = F.prelu(y, sd[q + "5.weight"])
    elif activation == "SiLU":
        y = F.silu(y)
    else:
        y = F.relu(y)
    y = F.conv1d(y, sd[q + "6.weight"], sd[q + "6.bias"])
    return y + res_x


class OracleLYNXNet:
    def __init__(self, sd: Dict[str, torch.Tensor], n_feats: int, kernel_size: int = 31, activation: str = "PReLU",
                 strong_cond: bool = False):
        self.sd = {k: v.float() for k, v in sd.items()}
        self.n_feats, self.kernel_size, self.activation, self.strong_cond = n_feats, kernel_size, activation, strong_cond
        self.num_layers = 0
        while f"residual_layers.{self.num_layers}.diffusion_projection.weight" in sd:
            self.num_layers += 1
        self.calls = 0

    @torch.no_grad()
    def __call__(self, spec, diffusion_step, cond):
        """lynxnet.py:133-172 -- spec [B, F, M, T], diffusion_step [B] or [1], cond [B, H, T] -> [B, F, M, T]."""
        sd = self.sd
        self.calls += 1
        x = spec[:, 0] if self.n_feats == 1 else spec.flatten(start_dim=1, end_dim=2)
        x = F.conv1d(x, sd["input_projection.weight"], sd["input_projection.bias"])
        if not self.strong_cond:
            x = F.gelu(x)
        C = x.shape[1]
        e = sinusoidal_pos_emb(diffusion_step.float() if diffusion_step.dtype != torch.long else diffusion_step, C)
        e = F.linear(e, sd["diffusion_embedding.1.weight"], sd["diffusion_embedding.1.bias"])
        e = F.gelu(e)
        e = F.linear(e, sd["diffusion_embedding.3.weight"], sd["diffusion_embedding.3.bias"]).unsqueeze(-1)
        for i in range(self.num_layers):
            x = lynx_layer(sd, i, x, cond, e, self.kernel_size, self.activation, self.strong_cond)
        x = F.layer_norm(x.transpose(1, 2), (C,), sd["norm.weight"], sd["norm.bias"]).transpose(1, 2)
        x = F.conv1d(x, sd["output_projection.weight"], sd["output_projection.bias"])
        if self.n_feats == 1:
            return x[:, None]
        return x.reshape(x.shape[0], self.n_feats, -1, x.shape[2])
