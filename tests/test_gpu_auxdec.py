"""GPU parity of the native ConvNeXt aux decoder (diffsinger_b200.AuxDecoderAdaptor: 7-tap implicit GEMMs, depthwise k = 7,
LayerNorm, GELU / layer-scale + residual epilogues through libdsb200.so) against the reference's golden outputs and the CPU
oracle at the default size (6 x 512, configs/acoustic.yaml:85-101).  bf16 path: rel-L2 <= 1e-2 on the decoded mel."""
import os

import numpy as np
import pytest
import torch

import diffsinger_b200 as ds
from conftest import GOLDEN, rel_l2
from oracle.auxdec_oracle import aux_decoder

pytestmark = pytest.mark.gpu
dev = "cuda"
TOL = 1e-2


@pytest.mark.parametrize("tag", ["f1", "f2"])
def test_matches_reference_golden(tag):
    z = np.load(os.path.join(GOLDEN, f"auxdec_{tag}.npz"))
    H, M, Fn, C, L, k = [int(v) for v in z["cfg"]]
    ad = ds.AuxDecoderAdaptor(H, M, Fn, z["spec_min"].tolist(), z["spec_max"].tolist(), "convnext",
                              dict(num_channels=C, num_layers=L, kernel_size=k, dropout_rate=0.1))
    ad.load_state_dict({key[2:]: torch.from_numpy(z[key]) for key in z.files if key.startswith("w.")}, strict=True)
    ad = ad.eval().to(dev)
    with torch.no_grad():
        y = ad(torch.from_numpy(z["cond"]).to(dev), infer=True)
        y_raw = ad(torch.from_numpy(z["cond"]).to(dev), infer=False)
    assert y.shape == z["y"].shape
    assert rel_l2(y_raw.cpu(), torch.from_numpy(z["y_raw"])) < TOL
    assert rel_l2(y.cpu(), torch.from_numpy(z["y"])) < TOL


@pytest.mark.parametrize("B,T", [(2, 300), (1, 5), (3, 129)])
def test_default_size_matches_oracle(B, T):
    torch.manual_seed(0)
    ad = ds.AuxDecoderAdaptor(256, 128, 1, [-12.0], [0.0], "convnext", dict(num_channels=512, num_layers=6, kernel_size=7, dropout_rate=0.1))
    with torch.no_grad():
        for blk in ad.decoder.conv:
            blk.gamma.copy_(0.2 + 0.5 * torch.rand_like(blk.gamma))
    sd = {k: v.clone() for k, v in ad.state_dict().items()}
    ad = ad.eval().to(dev)
    cond = torch.randn(B, T, 256, generator=torch.Generator().manual_seed(T))
    ref = aux_decoder(sd, cond, 1, 128, [-12.0], [0.0], infer=True)
    with torch.no_grad():
        y = ad(cond.to(dev), infer=True)
    err = rel_l2(y.cpu(), ref)
    print(f"aux decoder B={B} T={T} rel-L2 {err:.3e}")
    assert err < TOL
    with pytest.raises(RuntimeError, match="no CPU inference path"):
        with torch.no_grad():
            ds.ConvNeXtDecoder(256, 128)(cond)
