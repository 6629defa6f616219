"""NSF-HiFiGAN native path on the GPU: against golden outputs of the unmodified reference and against
the fp32 torch-op composition of the same module at the published 44.1 kHz configuration.  bf16 activations between ~20
convolutions per stage: centered rel-L2 (error against the waveform's variation) <= 3e-2."""
import pytest
import torch

from conftest import centered_rel_l2, load_nsf_golden

pytestmark = pytest.mark.gpu
WAV_TOL = 3e-2


@pytest.mark.parametrize("tag", ["a", "b"])
def test_native_generator_matches_reference_golden(tag):
    from diffsinger_b200.vocoder import Generator
    d, sd, cfg = load_nsf_golden(tag)
    gen = Generator(cfg).eval()
    gen.load_state_dict(sd, strict=True)
    gen = gen.cuda()
    with torch.no_grad():
        wav = gen(d["mel"].cuda(), d["f0"].cuda(), rand_ini=d["rand_ini"], noise=d["noise"])
    assert wav.shape == d["wav"].shape
    err = centered_rel_l2(wav.cpu(), d["wav"])
    print(f"nsf-hifigan {tag}: centered rel-L2 {err:.3e}, launches {gen.last_launches}")
    assert err < WAV_TOL
    n_stage = len(cfg["upsample_rates"])
    assert gen.last_launches == 2 + n_stage * (2 + 18 + 1) + 1


@pytest.mark.parametrize("B,T", [(1, 40), (2, 23)])
def test_native_generator_published_config(B, T):
    """upsample 8-8-2-2-2 from 512 channels (256 / 128 / 64 / 32 / 16: the last two padded to 64), hop 512."""
    from diffsinger_b200.vocoder import DEFAULT_CONFIG, Generator
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.manual_seed(B * 100 + T)
    gen = Generator(DEFAULT_CONFIG).eval()
    with torch.no_grad():  # the 0.01-std initialisers would make every conv a near no-op
        for n, p in gen.named_parameters():
            if p.dim() == 3 and "noise_convs" not in n and "conv_pre" not in n:
                p.mul_(3.0 if "ups" in n else 5.0)
    gen = gen.cuda()
    mel = torch.randn(B, 128, T, device="cuda") * 2.0 - 4.0
    f0 = 220.0 * torch.exp2(torch.rand(B, T, device="cuda"))
    f0[:, 5:9] = 0.0
    rand_ini = torch.rand(1, 1, 9, device="cuda")
    rand_ini[..., 0] = 0
    noise = torch.randn(B, T * 512, 9, device="cuda")
    with torch.no_grad():
        wav = gen(mel, f0, rand_ini=rand_ini, noise=noise)
        har = gen.m_source(f0, gen.upp, rand_ini, noise).transpose(1, 2)
        ref = gen._forward_torch(mel, har)
    assert wav.shape == ref.shape == (B, 1, T * 512)
    err = centered_rel_l2(wav, ref)
    print(f"nsf-hifigan 44.1k B={B} T={T}: centered rel-L2 {err:.3e}")
    assert err < WAV_TOL


def test_spec2wav_torch_wrapper():
    from diffsinger_b200.vocoder import DEFAULT_CONFIG, Generator, NsfHifiGAN
    cfg = dict(DEFAULT_CONFIG, upsample_initial_channel=128)
    voc = NsfHifiGAN(generator=Generator(cfg))
    voc.to_device("cuda")
    mel = torch.randn(1, 30, 128, device="cuda") - 2.0
    f0 = torch.full((1, 30), 261.6, device="cuda")
    y = voc.spec2wav_torch(mel, f0=f0)
    assert y.shape == (30 * 512,) and torch.isfinite(y).all() and float(y.abs().max()) <= 1.0
