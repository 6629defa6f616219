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


def test_ragged_pipeline_aux_decoder_then_shallow_sampler():
    """The default acoustic pipeline after the encoder (modules/toplevel.py:94-103: aux decoder -> shallow diffusion) over segments of
    different lengths as ragged batches: `sample_segments(diffusion, conds, aux_decoder=...)` must equal the segment-by-segment loop
    (each utterance keeps its own conv edges in the ConvNeXt decoder AND in the LYNXNet denoiser)."""
    from diffsinger_b200.synthetic import random_lynx_state_dict
    H, M, C = 256, 128, 256
    torch.manual_seed(0)
    ad = ds.AuxDecoderAdaptor(H, M, 1, [-12.0], [0.0], "convnext", dict(num_channels=256, num_layers=2, kernel_size=7, dropout_rate=0.1))
    with torch.no_grad():
        for blk in ad.decoder.conv:
            blk.gamma.fill_(0.3)  # layer scale away from its 1e-6 initialisation, or the blocks are invisible
    ad = ad.eval().to(dev)
    ds.hparams.clear()
    ds.hparams.update(hidden_size=H, use_shallow_diffusion=True, sampling_algorithm="euler", sampling_steps=4, T_start=0.4, T_start_infer=0.4,
                      time_scale_factor=1000)
    m = ds.RectifiedFlow(M, 1, t_start=0.4, backbone_type="lynxnet",
                         backbone_args=dict(num_layers=2, num_channels=C, expansion_factor=2, kernel_size=31, strong_cond=True),
                         spec_min=[-12.0], spec_max=[0.0])
    m.velocity_fn.load_state_dict(random_lynx_state_dict(M, 1, 2, C, H, expansion_factor=2, kernel_size=31, activation="PReLU", seed=4))
    m = m.eval().to(dev)
    lens = [300, 77, 515, 257]
    g = torch.Generator().manual_seed(31)
    conds = [torch.randn(1, t, H, generator=g).to(dev) for t in lens]
    torch.manual_seed(5)
    with torch.no_grad():
        alone = [m(c, src_spec=ad(c, infer=True), infer=True).cpu() for c in conds]
    torch.manual_seed(5)
    batched = [y.cpu() for y in ds.sample_segments(m, conds, aux_decoder=ad)]
    assert any(ws.lengths is not None for ws in m.velocity_fn.engine().workspaces.values()), "the ragged path did not run"
    worst = 0.0
    for t, a, b in zip(lens, alone, batched):
        assert a.shape == b.shape == (1, t, M)
        worst = max(worst, rel_l2(b, a), rel_l2(b[:, -24:], a[:, -24:]), rel_l2(b[:, :24], a[:, :24]))
    print(f"[ragged aux decoder + LYNXNet shallow reflow] worst rel-L2 vs one by one: {worst:.3e}")
    assert worst < 1e-2
    # the aux decoder alone: a padded batch with lengths equals the lone run, without them the last frames differ
    t_max = max(lens)
    cond = torch.zeros(2, t_max, H, device=dev)
    cond[0, :lens[0]], cond[1] = conds[0][0], conds[2][0]
    with torch.no_grad():
        lone = ad(conds[0], infer=True).cpu()
        masked = ad(cond, infer=True, lengths=[lens[0], lens[2]]).cpu()
        plain = ad(cond, infer=True).cpu()
    assert rel_l2(masked[0:1, :lens[0]], lone) < 1e-3
    assert rel_l2(plain[0:1, lens[0] - 6:lens[0]], lone[:, -6:]) > 1e-2
