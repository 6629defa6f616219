"""The drop-in, installed INSIDE the imported reference: with `diffsinger_b200.install_dropin()`
the reference's own top-level models -- DiffSingerAcoustic / DiffSingerVariance (modules/toplevel.py:32-122, 125-309), built from
the reference's own YAML configs -- construct this package's samplers / backbones / aux decoder, and their `state_dict`s are
interchangeable with the stock models' (same keys in the same order, shapes, dtypes; strict load both ways; sampler buffers
bit-equal) -- the checkpoint ABI of SURVEY.md section 8(b).  CPU only; skipped where /root/reference does not exist (the GPU box)."""
import pytest
import torch

from oracle.ref_harness import load_reference, load_reference_config, reference_available

pytestmark = pytest.mark.skipif(not reference_available(), reason="reference tree not present")

WAVENET = dict(backbone_type="wavenet", backbone_args=dict(num_layers=6, num_channels=128, dilation_cycle_length=4))
CASES = {
    # acoustic.yaml defaults: reflow + LYNXNet + shallow diffusion (ConvNeXt aux decoder)
    "acoustic_default_lynxnet": ("acoustic.yaml", {}),
    "acoustic_reflow_wavenet_shallow": ("acoustic.yaml", dict(WAVENET)),
    "acoustic_ddpm_wavenet": ("acoustic.yaml", dict(WAVENET, diffusion_type="ddpm", use_shallow_diffusion=False, K_step=1000)),
    "acoustic_rel_pos_encoder": ("acoustic.yaml", dict(WAVENET, use_rope=False)),  # encoder configuration this package does not build
    "acoustic_ddpm_wavenet_shallow": ("acoustic.yaml", dict(WAVENET, diffusion_type="ddpm", K_step=400, K_step_infer=400)),
    # variance.yaml defaults: reflow, pitch WaveNet 20x256 cycle 5 + multi-variance WaveNet 10x192
    "variance_reflow": ("variance.yaml", dict(predict_energy=True, predict_breathiness=True)),
    "variance_ddpm": ("variance.yaml", dict(predict_energy=True, predict_breathiness=True, predict_tension=True, diffusion_type="ddpm")),
}


def _build(ref, cfg_name, over):
    import modules.toplevel as mt
    ref.hparams.clear()
    ref.hparams.update(load_reference_config(cfg_name))
    ref.hparams.update(over)
    torch.manual_seed(0)
    if cfg_name.startswith("acoustic"):
        return mt.DiffSingerAcoustic(vocab_size=61, out_dims=ref.hparams["audio_num_mel_bins"])
    return mt.DiffSingerVariance(vocab_size=61)


def _samplers(model):
    return [m for n, m in model.named_children() if n in ("diffusion", "pitch_predictor", "variance_predictor")]


@pytest.mark.parametrize("case", sorted(CASES))
def test_reference_toplevel_models_build_on_the_dropin(case):
    import diffsinger_b200 as dsb
    ref = load_reference()
    cfg_name, over = CASES[case]
    stock = _build(ref, cfg_name, over)
    uninstall = dsb.install_dropin()
    try:
        assert dsb.get_hparams() is ref.hparams  # per-call hparams reads see the reference's dict
        ours = _build(ref, cfg_name, over)
    finally:
        uninstall()
    # the reference's model now holds this package's classes ...
    for smp_stock, smp in zip(_samplers(stock), _samplers(ours)):
        assert type(smp).__module__.startswith("diffsinger_b200.core") and type(smp).__name__ == type(smp_stock).__name__
        bb = smp.denoise_fn if hasattr(smp, "denoise_fn") else smp.velocity_fn
        assert type(bb).__module__.startswith("diffsinger_b200.backbones")
        for attr in ("out_dims", "num_feats", "t_start", "time_scale_factor", "use_shallow_diffusion"):
            assert getattr(smp, attr) == getattr(smp_stock, attr), attr
        if hasattr(smp_stock, "k_step"):
            assert (smp.k_step, smp.timesteps) == (smp_stock.k_step, smp_stock.timesteps)
    assert len(_samplers(ours)) == len(_samplers(stock)) >= 1
    if getattr(stock, "use_shallow_diffusion", False):
        assert type(ours.aux_decoder.decoder).__module__.startswith("diffsinger_b200")
    if cfg_name.startswith("acoustic"):  # the condition producer (configs/acoustic.yaml: use_rope true)
        want = "diffsinger_b200.fastspeech" if ref.hparams.get("use_rope") else "modules.fastspeech.acoustic_encoder"
        assert type(ours.fs2).__module__ == want and type(stock.fs2).__module__.startswith("modules.")
    # ... and after uninstall the stock classes are back
    again = _build(ref, cfg_name, over)
    assert type(_samplers(again)[0]).__module__.startswith("modules.core")
    # checkpoint ABI: identical keys (and order), shapes, dtypes; strict load both ways; sampler buffers bit-equal
    sd_s, sd_o = stock.state_dict(), ours.state_dict()
    assert list(sd_s.keys()) == list(sd_o.keys())
    for k in sd_s:
        assert sd_s[k].shape == sd_o[k].shape and sd_s[k].dtype == sd_o[k].dtype, k
        if any(b in k for b in ("betas", "alphas_cumprod", "posterior_", "spec_min", "spec_max")):
            assert torch.equal(sd_s[k], sd_o[k]), k
    ours.load_state_dict(sd_s, strict=True)
    stock.load_state_dict(ours.state_dict(), strict=True)
    for k, v in ours.state_dict().items():
        assert torch.equal(v, sd_s[k]), k
    # non-persistent buffers the callers read (reflow spec_min/max, ddpm timestep_factors) exist on both sides
    for smp_stock, smp in zip(_samplers(stock), _samplers(ours)):
        assert {n for n, _ in smp.named_buffers()} == {n for n, _ in smp_stock.named_buffers()}
    # no silent CPU path: inference on a CPU module raises instead of running torch ops
    smp = _samplers(ours)[0]
    bb = smp.denoise_fn if hasattr(smp, "denoise_fn") else smp.velocity_fn
    H = ref.hparams["hidden_size"]
    with torch.no_grad(), pytest.raises(RuntimeError, match="no CPU inference path"):
        bb(torch.zeros(1, smp.num_feats, smp.out_dims, 8), torch.zeros(1), torch.zeros(1, H, 8))
