"""NSF-HiFiGAN on the host: the oracle restatement and this package's module (torch-op branch,
checkpoint loading) against golden outputs of the imported, unmodified reference (oracle/make_golden.py: gen_nsf_hifigan)."""
import pytest
import torch

from conftest import centered_rel_l2, load_nsf_golden, rel_l2


@pytest.mark.parametrize("tag", ["a", "b"])
def test_oracle_matches_reference_golden(tag):
    from oracle.nsf_hifigan_oracle import generator, harmonic_source
    d, sd, cfg = load_nsf_golden(tag)
    upp = 1
    for r in cfg["upsample_rates"]:
        upp *= r
    har = harmonic_source(sd, d["f0"], upp, cfg["sampling_rate"], d["rand_ini"], d["noise"])
    assert har.shape == d["har_source"].shape and rel_l2(har, d["har_source"]) < 1e-5
    wav = generator(sd, cfg, d["mel"], d["har_source"])
    assert wav.shape == d["wav"].shape and centered_rel_l2(wav, d["wav"]) < 1e-5


@pytest.mark.parametrize("tag", ["a", "b"])
def test_module_torch_branch_matches_reference_golden(tag):
    from diffsinger_b200.vocoder import Generator
    d, sd, cfg = load_nsf_golden(tag)
    gen = Generator(cfg).eval()
    gen.load_state_dict(sd, strict=True)  # the reference's parameter names (after remove_weight_norm) load as they are
    with torch.no_grad():
        wav = gen(d["mel"], d["f0"], rand_ini=d["rand_ini"], noise=d["noise"])
    assert wav.shape == d["wav"].shape and centered_rel_l2(wav, d["wav"]) < 1e-5


def test_weight_norm_checkpoints_fold_on_load():
    """A reference checkpoint holds (weight_g, weight_v) for every weight-normed conv (models.py:215,224,247)."""
    from diffsinger_b200.vocoder import Generator
    d, sd, cfg = load_nsf_golden("a")
    wn = {}
    for k, v in sd.items():
        is_wn = k.endswith(".weight") and v.dim() == 3 and not k.startswith("noise_convs")
        if is_wn:
            base = k[:-len("weight")]
            scale = torch.rand(v.shape[0], 1, 1) + 0.5
            wn[base + "weight_v"] = v * scale
            wn[base + "weight_g"] = (v * scale).reshape(v.shape[0], -1).norm(dim=1).reshape(-1, 1, 1) / scale
        else:
            wn[k] = v
    gen = Generator(cfg).eval()
    gen.load_state_dict(wn, strict=True)
    for k, v in sd.items():
        assert torch.allclose(gen.state_dict()[k], v, rtol=1e-5, atol=1e-7), k


def test_unsupported_configurations_are_rejected():
    from diffsinger_b200.vocoder import DEFAULT_CONFIG, Generator
    with pytest.raises(NotImplementedError):
        Generator(dict(DEFAULT_CONFIG, resblock="2"))
    with pytest.raises(NotImplementedError):
        Generator(dict(DEFAULT_CONFIG, mini_nsf=True))
