"""TEST INFRASTRUCTURE ONLY -- imports the *unmodified* reference from /root/reference.

The reference (hrukalive/DiffSinger) is plain Python/PyTorch, so it can be imported in the
build container to (a) validate the oracle restatement in ``oracle/`` and (b) generate the
golden vectors under ``tests/golden/`` (see ``oracle/make_golden.py``).  It can NOT travel to
the GPU box: nothing under ``-m gpu`` tests, ``smoke()`` or ``bench.py`` may call into this
module.  Every public function here raises ``ReferenceUnavailable`` when /root/reference is
absent.

Three third-party packages the reference imports at module import time are absent from this
image (``lightning``, ``librosa``, ``matplotlib``); empty stand-ins are pre-seeded into
``sys.modules`` so that ``modules.backbones`` / ``modules.core`` import without touching the
reference tree.
"""
from __future__ import annotations

import os
import sys
import types

_REF_COPY = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")  # oracle/build_ref.py (git-ignored, travels to the GPU box)
REFERENCE_ROOT = os.environ.get("DS_REFERENCE_ROOT") or ("/root/reference" if os.path.isdir("/root/reference") else _REF_COPY)


class ReferenceUnavailable(RuntimeError):
    pass


def reference_available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "modules", "backbones", "wavenet.py"))


def _stub(name: str, **attrs) -> types.ModuleType:
    mod = sys.modules.get(name)
    if mod is None:
        mod = types.ModuleType(name)
        mod.__path__ = []  # behave like a package
        sys.modules[name] = mod
    for k, v in attrs.items():
        setattr(mod, k, v)
    return mod


def _noop_attr(attr):
    if attr.startswith("__"):
        raise AttributeError(attr)
    return lambda *a, **k: None


def _install_stubs() -> None:
    try:
        import lightning  # noqa: F401
        return
    except ModuleNotFoundError:
        pass

    class _Dummy:
        def __init__(self, *a, **k):
            pass

    def rank_zero_only(fn):
        return fn

    rank_zero_only.rank = 0

    def _print(*a, **k):
        print(*a)

    _stub("lightning")
    _stub("lightning.pytorch", LightningModule=type("LightningModule", (_Dummy,), {}),
          Trainer=type("Trainer", (_Dummy,), {}))
    _stub("lightning.pytorch.callbacks", ModelCheckpoint=type("ModelCheckpoint", (_Dummy,), {}),
          TQDMProgressBar=type("TQDMProgressBar", (_Dummy,), {}))
    _stub("lightning.pytorch.loggers", TensorBoardLogger=type("TensorBoardLogger", (_Dummy,), {}))
    _stub("lightning.pytorch.utilities")
    _stub("lightning.pytorch.utilities.rank_zero", rank_zero_only=rank_zero_only,
          rank_zero_info=_print, rank_zero_debug=_print, rank_zero_warn=_print)
    _stub("lightning.fabric")
    _stub("lightning.fabric.loggers")
    _stub("lightning.fabric.loggers.tensorboard", _TENSORBOARD_AVAILABLE=False)
    for name in ("librosa", "librosa.filters", "matplotlib", "matplotlib.pyplot"):
        if name not in sys.modules:
            try:
                __import__(name)
            except ModuleNotFoundError:
                # any attribute resolves to a no-op callable (`from librosa.filters import mel`, `matplotlib.use(...)`): these
                # packages serve plotting / the DDSP vocoder / MIDI-name parsing, none of which is on the timed path
                _stub(name, use=lambda *a, **k: None).__getattr__ = _noop_attr


_loaded = None


def load_reference():
    """Returns a namespace with the reference's hot-path symbols (imported, unmodified)."""
    global _loaded
    if _loaded is not None:
        return _loaded
    if not reference_available():
        raise ReferenceUnavailable(f"reference tree not found at {REFERENCE_ROOT}")
    _install_stubs()
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    from utils.hparams import hparams  # reference's global config dict
    import modules.backbones as backbones
    from modules.backbones.wavenet import WaveNet, ResidualBlock
    from modules.commons.common_layers import SinusoidalPosEmb
    import modules.core.ddpm as ddpm
    import modules.core.reflow as reflow

    ns = types.SimpleNamespace(
        hparams=hparams, backbones=backbones, WaveNet=WaveNet, ResidualBlock=ResidualBlock,
        SinusoidalPosEmb=SinusoidalPosEmb, ddpm=ddpm, reflow=reflow,
    )
    _loaded = ns
    return ns


def default_hparams(**over) -> dict:
    """The hparams keys the path reads, with the config defaults."""
    hp = dict(hidden_size=256, schedule_type="linear", use_shallow_diffusion=False,
              timesteps=1000, K_step=1000, K_step_infer=1000, diff_speedup=50,
              diff_accelerator="ddim", sampling_algorithm="euler", sampling_steps=20,
              T_start=0.0, T_start_infer=0.0, time_scale_factor=1000, infer=False,
              max_beta=0.02)
    hp.update(over)
    return hp


def set_reference_hparams(**over) -> dict:
    ref = load_reference()
    ref.hparams.clear()
    ref.hparams.update(default_hparams(**over))
    return ref.hparams


def load_reference_config(name: str) -> dict:
    """The reference's YAML config `configs/<name>` with its `base_config` chain merged (utils/hparams.py:64-82), as a plain dict."""
    import yaml

    def rec(path):
        with open(path) as f:
            c = yaml.safe_load(f)
        out = {}
        bases = c.pop("base_config", [])
        for b in ([bases] if isinstance(bases, str) else bases):
            out.update(rec(os.path.join(REFERENCE_ROOT, b)))
        out.update(c)
        return out

    if not reference_available():
        raise ReferenceUnavailable(f"reference tree not found at {REFERENCE_ROOT}")
    return rec(os.path.join(REFERENCE_ROOT, "configs", name))
