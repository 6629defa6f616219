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


@pytest.mark.parametrize("B,T", [(1, 40), (2, 23), (3, 1)])
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
    f0[:, 5:9] = 0.0  # an unvoiced stretch (no-op for T < 6)
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


def test_graph_replay_and_deepcopy():
    """Second call replays the captured CUDA graph (same result, new inputs honoured); the module stays deep-copyable."""
    import copy
    from diffsinger_b200.vocoder import DEFAULT_CONFIG, Generator
    torch.manual_seed(3)
    gen = Generator(dict(DEFAULT_CONFIG, upsample_initial_channel=128)).eval().cuda()
    mel = torch.randn(1, 128, 25, device="cuda") - 3.0
    f0 = torch.full((1, 25), 300.0, device="cuda")
    ri, nz = torch.zeros(1, 1, 9, device="cuda"), torch.randn(1, 25 * 512, 9, device="cuda")
    with torch.no_grad():
        y1 = gen(mel, f0, rand_ini=ri, noise=nz)
        y2 = gen(mel, f0, rand_ini=ri, noise=nz)
        y3 = gen(mel * 0.5, f0, rand_ini=ri, noise=nz)
        y4 = copy.deepcopy(gen)(mel, f0, rand_ini=ri, noise=nz)
    assert torch.equal(y1, y2) and torch.equal(y1, y4) and not torch.equal(y1, y3)
    assert len(gen._graphs) == 1


def test_spec2wav_torch_wrapper():
    from diffsinger_b200.vocoder import DEFAULT_CONFIG, Generator, NsfHifiGAN
    cfg = dict(DEFAULT_CONFIG, upsample_initial_channel=128)
    voc = NsfHifiGAN(generator=Generator(cfg))
    voc.to_device("cuda")
    mel = torch.randn(1, 30, 128, device="cuda") - 2.0
    f0 = torch.full((1, 30), 261.6, device="cuda")
    y = voc.spec2wav_torch(mel, f0=f0)
    assert y.shape == (30 * 512,) and torch.isfinite(y).all() and float(y.abs().max()) <= 1.0


@pytest.mark.parametrize("c0,rates,ks", [(256, [4, 2, 2], [8, 4, 4]), (128, [4, 2, 2], [8, 4, 4])])
def test_native_generator_narrow_stage_plans(c0, rates, ks, monkeypatch):
    """Stage-width planning: (256 -> 128, 64, 32): last stage narrow through the GEMM up-sampler; (128 -> 64, 32, 16): the last
    two narrow, the final up-sampling through ds_conv_small; conv_post narrow in both."""
    from diffsinger_b200.vocoder import DEFAULT_CONFIG, Generator
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.manual_seed(c0)
    cfg = dict(DEFAULT_CONFIG, upsample_initial_channel=c0, upsample_rates=rates, upsample_kernel_sizes=ks, num_mels=64)
    gen = Generator(cfg).eval()
    assert gen._plan_widths()[1] == ([False, False, True] if c0 == 256 else [False, True, True])
    with torch.no_grad():
        for n, p in gen.named_parameters():
            if p.dim() == 3 and "noise_convs" not in n and "conv_pre" not in n:
                p.mul_(3.0 if "ups" in n else 5.0)
    gen = gen.cuda()
    B, T, upp = 2, 50, 16
    mel = torch.randn(B, 64, T, device="cuda") * 2.0 - 4.0
    f0 = 220.0 * torch.exp2(torch.rand(B, T, device="cuda"))
    rand_ini = torch.zeros(1, 1, 9, device="cuda")
    noise = torch.randn(B, T * upp, 9, device="cuda")
    with torch.no_grad():
        wav = gen(mel, f0, rand_ini=rand_ini, noise=noise)
        ref = gen._forward_torch(mel, gen.m_source(f0, gen.upp, rand_ini, noise).transpose(1, 2))
        monkeypatch.setenv("DSB200_VOC_NARROW", "0")  # the same network with every stage on the 64-padded GEMM path
        gen._packed = None
        wide = gen(mel, f0, rand_ini=rand_ini, noise=noise)
    err, err_wide, mutual = centered_rel_l2(wav, ref), centered_rel_l2(wide, ref), centered_rel_l2(wav, wide)
    print(f"nsf-hifigan c0={c0}: centered rel-L2 narrow {err:.3e}, padded {err_wide:.3e}, narrow vs padded {mutual:.3e}")
    # three stages with 5x-amplified weights: looser than the published configuration, but both layouts must agree
    assert wav.shape == ref.shape and err < 5e-2 and err_wide < 5e-2 and mutual < 3e-2
