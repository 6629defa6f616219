"""TEST INFRASTRUCTURE ONLY -- runs the UNMODIFIED reference (imported from /root/reference, or from its copy under
oracle/_ref/ on the GPU box: oracle/build_ref.py) as the CPU baseline of bench.py and of the config #1 tool.

Nothing here is imported by the product.  All functions run on the host CPU in fp32 with the reference's own classes:
`modules.core.GaussianDiffusion` / `RectifiedFlow` (+ the pitch / multi-variance wrappers) over `modules.backbones.wavenet.WaveNet`,
and for BASELINE config #1 the reference's own `.ds` front end: `DiffSingerAcousticInfer.preprocess_input`
(inference/ds_acoustic.py:70-181) -> `FastSpeech2Acoustic` (modules/fastspeech/acoustic_encoder.py:82-118) -> `condition`.
"""
from __future__ import annotations

import json
import os
import time

import torch

from oracle.ref_harness import REFERENCE_ROOT, load_reference, load_reference_config, reference_available, set_reference_hparams


def cpu_info():
    model, n = "unknown", os.cpu_count() or 1
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                model = line.split(":", 1)[1].strip()
                break
    except OSError:
        pass
    return {"model": model, "logical_cores": n}


def build_sampler(kind: str, sd: dict, *, M: int, F: int, L: int, C: int, cycle: int, H: int, K: int, timesteps: int = 1000):
    """kind: 'dpm-solver' | 'ddim' | 'unipc' | 'pndm' (GaussianDiffusion) or 'euler' | 'rk2' | 'rk4' | 'rk5' (RectifiedFlow)."""
    ref = load_reference()
    args = dict(num_layers=L, num_channels=C, dilation_cycle_length=cycle)
    if kind in ("euler", "rk2", "rk4", "rk5"):
        set_reference_hparams(hidden_size=H, sampling_algorithm=kind, sampling_steps=K)
        m = ref.reflow.RectifiedFlow(M, F, 0.0, 1000, "wavenet", args, spec_min=[-12.0], spec_max=[0.0])
        m.velocity_fn.load_state_dict(sd)
    else:
        set_reference_hparams(hidden_size=H, diff_accelerator=kind, diff_speedup=timesteps // K, timesteps=timesteps, K_step=timesteps,
                              K_step_infer=timesteps)
        m = ref.ddpm.GaussianDiffusion(M, F, timesteps, timesteps, "wavenet", args, spec_min=[-12.0], spec_max=[0.0])
        m.denoise_fn.load_state_dict(sd)
    return m.eval()


@torch.no_grad()
def run_sampler(m, cond_bth: torch.Tensor, noise: torch.Tensor | None = None):
    """The reference's public call `diffusion(condition, infer=True)` (ddpm.py:353-377 / reflow.py:43-64); `noise` replaces the
    torch.randn draw inside inference() when given.  Returns (mel, seconds)."""
    real = torch.randn
    if noise is not None:
        torch.randn = lambda *a, **k: noise.clone()
    try:
        t0 = time.perf_counter()
        y = m(cond_bth, infer=True)
        dt = time.perf_counter() - t0
    finally:
        torch.randn = real
    return y, dt


# ------------------------------------------------------------------------------------------------ BASELINE config #1
CONFIG1_OVERRIDES = dict(diffusion_type="ddpm", backbone_type="wavenet",
                         backbone_args=dict(num_layers=20, num_channels=256, dilation_cycle_length=4),
                         use_shallow_diffusion=False, diff_accelerator="ddim", diff_speedup=50, K_step=1000, K_step_infer=1000,
                         timesteps=1000, schedule_type="linear", infer=False, use_spk_id=False)


def config1_model_and_conditions(ds_name: str = "00_", seed: int = 0):
    """Random-init `DiffSingerAcoustic` (GaussianDiffusion + WaveNet 20 x 256, 128 mel; no checkpoint is shipped) built by the
    reference from configs/acoustic.yaml, and the conditions its own encoder produces for every segment of samples/00_*.ds.
    Returns (model, [cond [1, T, 256] per segment], [the encoder's input batch per segment: tokens, mel2ph, f0, ...])."""
    import glob
    ref = load_reference()
    ref.hparams.clear()
    ref.hparams.update(load_reference_config("acoustic.yaml"))
    ref.hparams.update(CONFIG1_OVERRIDES)
    ref.hparams.update(dictionaries={"zh": os.path.join(REFERENCE_ROOT, "dictionaries", "opencpop-extension.txt")},
                       work_dir=os.path.join(REFERENCE_ROOT, "checkpoints", "__none__"))
    import utils.phoneme_utils as pu
    pu._dictionary = None
    from inference.ds_acoustic import DiffSingerAcousticInfer
    from modules.fastspeech.tts_modules import LengthRegulator
    from modules.toplevel import DiffSingerAcoustic
    inf = DiffSingerAcousticInfer(device="cpu", load_model=False, load_vocoder=False)
    inf.variance_checklist, inf.variances_to_embed = [], set()
    inf.phoneme_dictionary = pu.load_phoneme_dictionary()
    torch.manual_seed(seed)
    model = DiffSingerAcoustic(vocab_size=len(inf.phoneme_dictionary), out_dims=ref.hparams["audio_num_mel_bins"]).eval()
    torch.nn.init.normal_(model.diffusion.denoise_fn.output_projection.weight, 0.0, 0.05)  # zero-initialised otherwise (trap T1)
    inf.model, inf.lr = model, LengthRegulator()
    path = sorted(glob.glob(os.path.join(REFERENCE_ROOT, "samples", ds_name + "*.ds")))[0]
    params = json.load(open(path, encoding="utf8"))
    conds, batches = [], []
    grab = model.fs2.register_forward_hook(lambda mod, a, out: conds.append(out.detach().clone()))
    stop = _StopAfterEncoder()
    model.diffusion.register_forward_pre_hook(stop)
    import contextlib
    import io
    try:
        for i, param in enumerate(params):
            with contextlib.redirect_stdout(io.StringIO()):
                batch = inf.preprocess_input(param, idx=i)
            batches.append(batch)
            try:
                with torch.no_grad():
                    inf.forward_model(batch)
            except _Stop:
                pass
    finally:
        grab.remove()
        stop.armed = False
    return model, conds, batches


class _Stop(Exception):
    pass


class _StopAfterEncoder:
    armed = True

    def __call__(self, mod, args):
        if self.armed:
            raise _Stop()
