"""TEST INFRASTRUCTURE ONLY -- generates tests/golden/*.npz by running the UNMODIFIED reference.

Run in the build container (needs /root/reference):  python -m oracle.make_golden
The fixtures hold weights (reference state_dict), inputs (cond, the noise the reference drew)
and the reference's outputs, so they can be checked anywhere (CPU oracle tests, GPU parity tests)
without the reference tree.  torch version is stamped into every file.
"""
from __future__ import annotations

import os
from contextlib import contextmanager

import numpy as np
import torch

from oracle.ref_harness import load_reference, set_reference_hparams

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")

SMALL = dict(num_layers=4, num_channels=64, dilation_cycle_length=2)
HIDDEN = 64


@contextmanager
def record_randn(log):
    real = torch.randn

    def fake(*a, **k):
        out = real(*a, **k)
        log.append(out.clone())
        return out

    torch.randn = fake
    try:
        yield
    finally:
        torch.randn = real


def _np(sd):
    return {k: v.detach().cpu().numpy() for k, v in sd.items()}


def _save(name, **arrs):
    os.makedirs(OUT, exist_ok=True)
    arrs["torch_version"] = np.array(torch.__version__)
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **arrs)
    print("wrote", name, {k: getattr(v, "shape", None) for k, v in arrs.items() if not k.startswith("w.")})


def _randomise_out_proj(net, std=0.05):
    torch.nn.init.normal_(net.output_projection.weight, 0.0, std)  # trap T1


_W = {}


def shared_weights(ref, M, F):
    """One weight set per (M, F), saved once as weights_m{M}f{F}.npz and loaded into every model."""
    tag = f"m{M}f{F}"
    existing = os.path.join(OUT, "weights_" + tag + ".npz")
    if tag not in _W and os.path.exists(existing):  # later fixture batches reuse the committed weight sets
        z = np.load(existing)
        _W[tag] = {k: torch.from_numpy(z[k]) for k in z.files if k not in ("cfg", "torch_version")}
    if tag not in _W:
        set_reference_hparams(hidden_size=HIDDEN)
        torch.manual_seed(100 + F)
        net = ref.WaveNet(M, F, **SMALL).eval()
        _randomise_out_proj(net, 0.1)
        _W[tag] = {k: v.clone() for k, v in net.state_dict().items()}
        _save("weights_" + tag, cfg=np.array([M, F, SMALL["num_layers"], SMALL["num_channels"],
                                               SMALL["dilation_cycle_length"], HIDDEN]), **_np(_W[tag]))
    return tag, _W[tag]


def gen_backbone(ref):
    for tag, M, F in (("f1", 16, 1), ("f2", 8, 2)):
        set_reference_hparams(hidden_size=HIDDEN)
        net = ref.WaveNet(M, F, **SMALL).eval()
        wtag, sd = shared_weights(ref, M, F)
        net.load_state_dict(sd)
        torch.manual_seed(5)
        B, T = 2, 37
        spec = torch.randn(B, F, M, T)
        cond = torch.randn(B, T, HIDDEN).transpose(1, 2)  # transposed view like the callers (T8)
        t_long = torch.tensor([950, 3], dtype=torch.long)
        t_float = torch.tensor([412.5])
        with torch.no_grad():
            y_long = net(spec, t_long, cond)
            y_float = net(spec, t_float, cond)
        _save(f"wavenet_{tag}", spec=spec.numpy(), cond=cond.contiguous().numpy(), t_long=t_long.numpy(),
              t_float=t_float.numpy(), y_long=y_long.numpy(), y_float=y_float.numpy(),
              weights=np.array(wtag))


def gen_ddpm(ref):
    M, B, T = 16, 2, 41
    for algo, speedup, b in (("ddim", 50, B), ("pndm", 50, 1), ("dpm-solver", 50, B), ("unipc", 50, B),
                             ("dpm-solver", 125, B), ("ddpm", 1, B), ("ddim_shallow", 20, B)):
        shallow = algo in ("ddpm", "ddim_shallow")
        k_step = 12 if algo == "ddpm" else (400 if shallow else 1000)
        set_reference_hparams(hidden_size=HIDDEN, diff_accelerator=algo.split("_")[0], diff_speedup=speedup,
                              use_shallow_diffusion=shallow, K_step=k_step, K_step_infer=k_step)
        torch.manual_seed(7)
        m = ref.ddpm.GaussianDiffusion(M, 1, 1000, k_step, "wavenet", SMALL, spec_min=[-12.0], spec_max=[0.0]).eval()
        wtag, sd = shared_weights(ref, M, 1)
        m.denoise_fn.load_state_dict(sd)
        cond = torch.randn(b, T, HIDDEN)
        src = torch.rand(b, T, M) * -10.0 if shallow else None
        log = []
        torch.manual_seed(11)
        with torch.no_grad(), record_randn(log):
            y = m(cond, src_spec=src, infer=True)
        arrs = dict(cond=cond.numpy(), noise=log[0].numpy(), y=y.numpy(),
                    meta=np.array([M, 1, 1000, k_step, speedup, int(shallow)]), weights=np.array(wtag))
        if len(log) > 1:
            arrs["step_noise"] = torch.stack(log[1:]).numpy()
        if src is not None:
            arrs["src"] = src.numpy()
        _save("ddpm_" + algo.replace("-", "_") + (f"_s{speedup}" if algo == "dpm-solver" else ""), **arrs)


def gen_reflow(ref):
    M, B, T = 16, 2, 33
    for algo, steps, t_start in (("euler", 20, 0.0), ("rk2", 6, 0.0), ("rk4", 4, 0.0), ("rk5", 3, 0.0), ("euler", 8, 0.4)):
        shallow = t_start > 0
        set_reference_hparams(hidden_size=HIDDEN, sampling_algorithm=algo, sampling_steps=steps,
                              use_shallow_diffusion=shallow, T_start=t_start, T_start_infer=t_start)
        torch.manual_seed(9)
        m = ref.reflow.RectifiedFlow(M, 1, t_start, 1000, "wavenet", SMALL, spec_min=[-12.0], spec_max=[0.0]).eval()
        wtag, sd = shared_weights(ref, M, 1)
        m.velocity_fn.load_state_dict(sd)
        cond = torch.randn(B, T, HIDDEN)
        src = torch.rand(B, T, M) * -10.0 if shallow else None
        log = []
        torch.manual_seed(13)
        with torch.no_grad(), record_randn(log):
            y = m(cond, src_spec=src, infer=True)
        arrs = dict(cond=cond.numpy(), noise=log[0].numpy(), y=y.numpy(),
                    meta=np.array([M, 1, steps, int(shallow)]), t_start=np.array(t_start), weights=np.array(wtag))
        if src is not None:
            arrs["src"] = src.numpy()
        _save(f"reflow_{algo}" + ("_shallow" if shallow else ""), **arrs)


def gen_variance(ref):
    B, T = 1, 29
    # pitch (configs/variance.yaml:62-72 ranges, reduced repeat_bins)
    set_reference_hparams(hidden_size=HIDDEN, diff_accelerator="ddim", diff_speedup=100)
    torch.manual_seed(21)
    m = ref.ddpm.PitchDiffusion(vmin=-8.0, vmax=8.0, cmin=-12.0, cmax=12.0, repeat_bins=16, timesteps=1000,
                                k_step=1000, backbone_type="wavenet", backbone_args=SMALL).eval()
    wtag, sd = shared_weights(ref, 16, 1)
    m.denoise_fn.load_state_dict(sd)
    cond = torch.randn(B, T, HIDDEN)
    log = []
    torch.manual_seed(23)
    with torch.no_grad(), record_randn(log):
        y = m(cond, infer=True)
    _save("var_pitch_ddim", cond=cond.numpy(), noise=log[0].numpy(), y=y.numpy(), weights=np.array(wtag))
    # multi-variance, reflow euler (configs/variance.yaml:74-95 ranges)
    set_reference_hparams(hidden_size=HIDDEN, sampling_algorithm="euler", sampling_steps=10)
    torch.manual_seed(25)
    m = ref.reflow.MultiVarianceRectifiedFlow(ranges=[(-96.0, -12.0), (-96.0, -20.0)], clamps=[(-96.0, 0.0), (-96.0, 0.0)],
                                              repeat_bins=8, backbone_type="wavenet", backbone_args=SMALL).eval()
    wtag, sd = shared_weights(ref, 8, 2)
    m.velocity_fn.load_state_dict(sd)
    log = []
    torch.manual_seed(27)
    with torch.no_grad(), record_randn(log):
        ys = m(cond, infer=True)
    _save("var_multi_euler", cond=cond.numpy(), noise=log[0].numpy(), y=torch.stack(list(ys)).numpy(),
          weights=np.array(wtag))


LYNX_SMALL = dict(num_layers=3, num_channels=64, expansion_factor=2, kernel_size=31)


def gen_lynxnet(ref):
    """LYNXNet (modules/backbones/lynxnet.py), the reference's other registry entry: forward fixtures for the three
    activations, F = 1 / F = 2 and strong_cond, with non-trivial norm / PReLU parameters and a re-randomised output
    projection (it is zero-initialised like WaveNet's, trap T1)."""
    from modules.backbones.lynxnet import LYNXNet
    for tag, M, Fn, act, strong in (("prelu_f1", 16, 1, "PReLU", False), ("silu_f2", 8, 2, "SiLU", False),
                                    ("relu_strong", 16, 1, "ReLU", True)):
        set_reference_hparams(hidden_size=HIDDEN)
        torch.manual_seed(300 + Fn)
        net = LYNXNet(M, Fn, activation=act, strong_cond=strong, **LYNX_SMALL).eval()
        _randomise_out_proj(net, 0.1)
        with torch.no_grad():
            for k, v in net.state_dict().items():  # LayerNorm affine and PReLU slopes away from their (1, 0, 0.25) defaults
                if k.endswith("net.0.weight") or k == "norm.weight":
                    v.copy_(1.0 + 0.2 * torch.randn_like(v))
                elif k.endswith("net.0.bias") or k == "norm.bias":
                    v.copy_(0.1 * torch.randn_like(v))
                elif k.endswith("net.5.weight"):
                    v.copy_(0.05 + 0.4 * torch.rand_like(v))
        torch.manual_seed(5)
        B, T = 2, 45
        spec = torch.randn(B, Fn, M, T)
        cond = torch.randn(B, T, HIDDEN).transpose(1, 2)
        t_long = torch.tensor([950, 3], dtype=torch.long)
        t_float = torch.tensor([412.5])
        with torch.no_grad():
            y_long = net(spec, t_long, cond)
            y_float = net(spec, t_float, cond)
        _save(f"lynxnet_{tag}", spec=spec.numpy(), cond=cond.contiguous().numpy(), t_long=t_long.numpy(), t_float=t_float.numpy(),
              y_long=y_long.numpy(), y_float=y_float.numpy(),
              cfg=np.array([M, Fn, LYNX_SMALL["num_layers"], LYNX_SMALL["num_channels"], LYNX_SMALL["expansion_factor"],
                            LYNX_SMALL["kernel_size"], HIDDEN, int(strong)]), activation=np.array(act),
              **{"w." + k: v.numpy() for k, v in net.state_dict().items()})


def gen_aux_decoder(ref):
    """ConvNeXt aux decoder through AuxDecoderAdaptor (modules/aux_decoder/__init__.py:30-75, convnext.py:58-87) with
    non-trivial layer scales (they are initialised to 1e-6, which would hide the blocks)."""
    from modules.aux_decoder import AuxDecoderAdaptor
    for tag, Fn in (("f1", 1), ("f2", 2)):
        torch.manual_seed(400 + Fn)
        M = 16 // Fn
        smin, smax = ([-12.0] if Fn == 1 else [[-96.0], [-80.0]]), ([0.0] if Fn == 1 else [[-12.0], [-20.0]])
        ad = AuxDecoderAdaptor(HIDDEN, M, Fn, smin, smax, "convnext",
                               dict(num_channels=64, num_layers=2, kernel_size=7, dropout_rate=0.1)).eval()
        with torch.no_grad():
            for blk in ad.decoder.conv:
                blk.gamma.copy_(0.5 + torch.rand_like(blk.gamma))
                blk.norm.weight.copy_(1.0 + 0.2 * torch.randn_like(blk.norm.weight))
                blk.norm.bias.copy_(0.1 * torch.randn_like(blk.norm.bias))
        cond = torch.randn(2, 53, HIDDEN)
        with torch.no_grad():
            y = ad(cond, infer=True)
            y_raw = ad(cond, infer=False)
        _save(f"auxdec_{tag}", cond=cond.numpy(), y=y.numpy(), y_raw=y_raw.numpy(), cfg=np.array([HIDDEN, M, Fn, 64, 2, 7]),
              spec_min=np.array(smin, dtype=np.float32), spec_max=np.array(smax, dtype=np.float32),
              **{"w." + k: v.numpy() for k, v in ad.state_dict().items()})


NSF_SMALL = dict(sampling_rate=44100, num_mels=32, upsample_rates=[4, 2, 2], upsample_kernel_sizes=[8, 4, 4],
                 upsample_initial_channel=64, resblock="1", resblock_kernel_sizes=[3, 7, 11],
                 resblock_dilation_sizes=[[1, 3, 5], [1, 3, 5], [1, 3, 5]], mini_nsf=False, noise_sigma=None)


def gen_nsf_hifigan(ref):
    """NSF-HiFiGAN generator (modules/nsf_hifigan/models.py:206-302) with weight norm removed, and its source module with
    the two random draws of SineGen recorded (torch.rand at :147, torch.randn_like at :166)."""
    from modules.nsf_hifigan.models import Generator
    from modules.nsf_hifigan.env import AttrDict
    for tag, B, T in (("a", 2, 37), ("b", 1, 64)):
        torch.manual_seed(500 + T)
        gen = Generator(AttrDict(NSF_SMALL)).eval()
        with torch.no_grad():  # the 0.01-std initialisers would make every conv a near no-op
            for n, p in gen.named_parameters():
                if n.endswith("weight_v") or (n.endswith("weight") and p.dim() == 3):
                    p.mul_(6.0 if "ups" not in n else 3.0)
        gen.remove_weight_norm()
        mel = torch.randn(B, NSF_SMALL["num_mels"], T) * 2.0 - 4.0
        f0 = 220.0 * torch.exp2(torch.rand(B, T) * 1.5)
        f0[:, T // 3: T // 3 + 5] = 0.0  # an unvoiced stretch
        draws = []
        real_rand, real_randn_like = torch.rand, torch.randn_like

        def rand(*a, **k):
            draws.append(real_rand(*a, **k))
            return draws[-1]

        def randn_like(*a, **k):
            draws.append(real_randn_like(*a, **k))
            return draws[-1]

        torch.rand, torch.randn_like = rand, randn_like
        try:
            with torch.no_grad():
                wav = gen(mel, f0)
        finally:
            torch.rand, torch.randn_like = real_rand, real_randn_like
        rand_ini, noise = draws[0].clone(), draws[1]
        rand_ini[..., 0] = 0  # models.py:148 zeroes the fundamental's phase in place after the draw
        with torch.no_grad():
            torch.manual_seed(1)
            har = None
            real = gen.m_source.l_sin_gen._f02sine
            torch.rand = lambda *a, **k: draws[0].clone()
            torch.randn_like = lambda *a, **k: noise
            try:
                har = gen.m_source(f0, gen.upp).transpose(1, 2)
            finally:
                torch.rand, torch.randn_like = real_rand, real_randn_like
        _save(f"nsf_hifigan_{tag}", mel=mel.numpy(), f0=f0.numpy(), rand_ini=rand_ini.numpy(), noise=noise.numpy(), har_source=har.numpy(),
              wav=wav.numpy(), cfg_json=np.array(__import__("json").dumps(NSF_SMALL)),
              **{"w." + k: v.numpy() for k, v in gen.state_dict().items()})


def gen_extra(ref):
    """Round-2 additions: the wrappers / schedules / shallow solver paths that had no fixture.
      var_multi_ddim      MultiVarianceDiffusion (ddpm.py:448-505), F = 2, DDIM 10 steps
      var_pitch_euler     PitchRectifiedFlow (reflow.py:184-203), Euler 10 steps
      ddpm_*_cosine       schedule_type='cosine' (ddpm.py:36-46) through DDIM, DPM-Solver++ (lambda-clip active,
                          dpm_solver_pytorch.py:114-125) and UniPC (no clip, uni_pc.py:77-86)
      ddpm_*_shallow      DPM-Solver++ / UniPC with use_shallow_diffusion, K_step = 400: NoiseScheduleVP over betas[:t_max]
                          (ddpm.py:246-322, trap T9)"""
    B, T = 1, 29
    cond = None
    # MultiVarianceDiffusion
    set_reference_hparams(hidden_size=HIDDEN, diff_accelerator="ddim", diff_speedup=100)
    torch.manual_seed(31)
    m = ref.ddpm.MultiVarianceDiffusion(ranges=[(-96.0, -12.0), (-96.0, -20.0)], clamps=[(-96.0, 0.0), None], repeat_bins=8,
                                        timesteps=1000, k_step=1000, backbone_type="wavenet", backbone_args=SMALL).eval()
    wtag, sd = shared_weights(ref, 8, 2)
    m.denoise_fn.load_state_dict(sd)
    cond = torch.randn(2, T, HIDDEN)
    log = []
    torch.manual_seed(33)
    with torch.no_grad(), record_randn(log):
        ys = m(cond, infer=True)
    _save("var_multi_ddim", cond=cond.numpy(), noise=log[0].numpy(), y=torch.stack(list(ys)).numpy(), weights=np.array(wtag))
    # PitchRectifiedFlow
    set_reference_hparams(hidden_size=HIDDEN, sampling_algorithm="euler", sampling_steps=10)
    torch.manual_seed(35)
    m = ref.reflow.PitchRectifiedFlow(vmin=-8.0, vmax=8.0, cmin=-12.0, cmax=12.0, repeat_bins=16, backbone_type="wavenet",
                                      backbone_args=SMALL).eval()
    wtag, sd = shared_weights(ref, 16, 1)
    m.velocity_fn.load_state_dict(sd)
    cond = torch.randn(2, T, HIDDEN)
    log = []
    torch.manual_seed(37)
    with torch.no_grad(), record_randn(log):
        y = m(cond, infer=True)
    _save("var_pitch_euler", cond=cond.numpy(), noise=log[0].numpy(), y=y.numpy(), weights=np.array(wtag))
    # cosine schedule and shallow solver paths
    M, B, T = 16, 2, 41
    for name, algo, sched, speedup, k_step in (("ddim_cosine", "ddim", "cosine", 50, 1000),
                                               ("dpm_solver_cosine", "dpm-solver", "cosine", 50, 1000),
                                               ("unipc_cosine", "unipc", "cosine", 50, 1000),
                                               ("dpm_solver_shallow", "dpm-solver", "linear", 20, 400),
                                               ("unipc_shallow", "unipc", "linear", 20, 400)):
        shallow = k_step < 1000
        set_reference_hparams(hidden_size=HIDDEN, diff_accelerator=algo, diff_speedup=speedup, schedule_type=sched,
                              use_shallow_diffusion=shallow, K_step=k_step, K_step_infer=k_step)
        torch.manual_seed(41)
        m = ref.ddpm.GaussianDiffusion(M, 1, 1000, k_step, "wavenet", SMALL, spec_min=[-12.0], spec_max=[0.0]).eval()
        wtag, sd = shared_weights(ref, M, 1)
        m.denoise_fn.load_state_dict(sd)
        cond = torch.randn(B, T, HIDDEN)
        src = torch.rand(B, T, M) * -10.0 if shallow else None
        log = []
        torch.manual_seed(43)
        with torch.no_grad(), record_randn(log):
            y = m(cond, src_spec=src, infer=True)
        assert torch.isfinite(y).all(), name
        arrs = dict(cond=cond.numpy(), noise=log[0].numpy(), y=y.numpy(), meta=np.array([M, 1, 1000, k_step, speedup, int(shallow)]),
                    schedule=np.array(sched), weights=np.array(wtag))
        if src is not None:
            arrs["src"] = src.numpy()
        _save("ddpm_" + name, **arrs)

FS2_HP = dict(hidden_size=64, enc_layers=2, num_heads=2, enc_ffn_kernel_size=3, ffn_act="gelu", dropout=0.1, use_pos_embed=True,
              use_rope=True, rel_pos=True, use_lang_id=False, use_spk_id=True, num_spk=3)


def gen_fs2(ref):
    """The condition producer: FastSpeech2Acoustic (modules/fastspeech/acoustic_encoder.py:14-118) with the rotary-embedding encoder of
    configs/acoustic.yaml at a small width; padded tokens and frames in the batch, every optional per-frame embedding in (a) and
    none of them in (b)."""
    from modules.fastspeech.acoustic_encoder import FastSpeech2Acoustic
    for tag, extra in (("a", dict(use_energy_embed=True, use_breathiness_embed=True, use_key_shift_embed=True, use_speed_embed=True)),
                       ("b", dict(use_energy_embed=False, use_breathiness_embed=False, use_key_shift_embed=False, use_speed_embed=False,
                                  use_spk_id=False))):
        set_reference_hparams(**dict(FS2_HP, **extra))
        torch.manual_seed(600 + len(tag) + (tag == "b"))
        m = FastSpeech2Acoustic(17).eval()
        with torch.no_grad():
            for k, v in m.state_dict().items():  # LayerNorm affine / biases away from their (1, 0) defaults
                if "layer_norm" in k and k.endswith("weight"):
                    v.copy_(1.0 + 0.2 * torch.randn_like(v))
                elif k.endswith("bias"):
                    v.copy_(0.1 * torch.randn_like(v))
        B, Tt, T = 2, 11, 57
        tok = torch.randint(1, 17, (B, Tt))
        tok[1, 8:] = 0
        n_real = [Tt, 8]
        mel2ph = torch.zeros(B, T, dtype=torch.long)
        for b, (n, frames) in enumerate(zip(n_real, (T, 44))):
            mel2ph[b, :frames] = torch.sort(torch.randint(1, n + 1, (frames,))).values
        f0 = torch.rand(B, T) * 300 + 80
        kw = {}
        if extra["use_energy_embed"]:
            kw.update(energy=torch.randn(B, T) * 10 - 40, breathiness=torch.randn(B, T) * 10 - 50, key_shift=torch.randn(B, 1) * 3,
                      speed=torch.rand(B, T) + 0.5)
        if extra.get("use_spk_id", True):
            kw.update(spk_embed_id=torch.tensor([2, 0]))
        with torch.no_grad():
            y = m(tok, mel2ph, f0, **kw)
        _save(f"fs2_{tag}", tokens=tok.numpy(), mel2ph=mel2ph.numpy(), f0=f0.numpy(), y=y.numpy(),
              hp_json=np.array(__import__("json").dumps(dict(FS2_HP, **extra))),
              **{"in." + k: v.numpy() for k, v in kw.items()}, **{"w." + k: v.numpy() for k, v in m.state_dict().items()})


def main():
    ref = load_reference()
    only = os.environ.get("GOLDEN_ONLY", "")
    if not only or "fs2" in only:
        gen_fs2(ref)
    if not only or "extra" in only:
        gen_extra(ref)
    if not only or "nsf" in only:
        gen_nsf_hifigan(ref)
    if not only or "auxdec" in only:
        gen_aux_decoder(ref)
    if not only or "lynxnet" in only:
        gen_lynxnet(ref)
    if only:
        return
    gen_backbone(ref)
    gen_ddpm(ref)
    gen_reflow(ref)
    gen_variance(ref)


if __name__ == "__main__":
    main()
