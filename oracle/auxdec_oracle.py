"""TEST INFRASTRUCTURE ONLY -- CPU restatement (torch fp32) of the reference ConvNeXt aux decoder from a plain state_dict
(key names of ``AuxDecoderAdaptor.state_dict()``: ``decoder.inconv.*``, ``decoder.conv.i.*``, ``decoder.outconv.*``).

Restates ``ConvNeXtBlock.forward`` (modules/aux_decoder/convnext.py:41-55), ``ConvNeXtDecoder.forward`` (:79-87) and the
reshape / de-normalisation of ``AuxDecoderAdaptor.forward`` (modules/aux_decoder/__init__.py:50-75).  Pinned by
tests/test_oracle_golden.py against tests/golden/auxdec_*.npz (outputs of the imported, unmodified reference)."""
from __future__ import annotations

import torch
import torch.nn.functional as F


@torch.no_grad()
def aux_decoder(sd, cond, n_feats: int, out_dims: int, spec_min=None, spec_max=None, infer: bool = True):
    """cond [B, T, H] -> [B, T, M] (n_feats == 1) or [B, F, T, M]."""
    sd = {k: v.float() for k, v in sd.items()}
    k = sd["decoder.inconv.weight"].shape[2]
    x = F.conv1d(cond.transpose(1, 2), sd["decoder.inconv.weight"], sd["decoder.inconv.bias"], padding=(k - 1) // 2)
    i = 0
    while f"decoder.conv.{i}.dwconv.weight" in sd:
        p = f"decoder.conv.{i}."
        C = x.shape[1]
        y = F.conv1d(x, sd[p + "dwconv.weight"], sd[p + "dwconv.bias"], padding=3, groups=C).transpose(1, 2)
        y = F.layer_norm(y, (C,), sd[p + "norm.weight"], sd[p + "norm.bias"], 1e-6)
        y = F.linear(F.gelu(F.linear(y, sd[p + "pwconv1.weight"], sd[p + "pwconv1.bias"])), sd[p + "pwconv2.weight"], sd[p + "pwconv2.bias"])
        if p + "gamma" in sd:
            y = sd[p + "gamma"] * y
        x = x + y.transpose(1, 2)
        i += 1
    x = F.conv1d(x, sd["decoder.outconv.weight"], sd["decoder.outconv.bias"], padding=(k - 1) // 2).transpose(1, 2)
    if n_feats > 1:
        x = x.reshape(-1, x.shape[1], n_feats, out_dims).transpose(1, 2)
    if infer and spec_min is not None:
        smin = torch.as_tensor(spec_min, dtype=torch.float32)[None, None, :].transpose(-3, -2)
        smax = torch.as_tensor(spec_max, dtype=torch.float32)[None, None, :].transpose(-3, -2)
        x = x * ((smax - smin) / 2.) + (smax + smin) / 2.
    return x
