import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_collection_modifyitems(config, items):
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(autouse=True)
def _reference_hparams_do_not_leak():
    """A test that imports the reference leaves `utils.hparams.hparams` filled, and diffsinger_b200.get_hparams() adopts a
    non-empty reference dict (hparams.py) -- the NEXT test's `ds.hparams.update(...)` would then be ignored."""
    yield
    mod = sys.modules.get("utils.hparams")
    if mod is not None and isinstance(getattr(mod, "hparams", None), dict):
        mod.hparams.clear()


def load_golden(name):
    """-> (dict of torch tensors / scalars, state_dict of the shared weight set, cfg)."""
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    d = {k: (torch.from_numpy(z[k]) if z[k].dtype.kind in "fi" and z[k].ndim > 0 else z[k]) for k in z.files}
    w = np.load(os.path.join(GOLDEN, "weights_" + str(z["weights"]) + ".npz"))
    sd = {k: torch.from_numpy(w[k]) for k in w.files if k not in ("cfg", "torch_version")}
    M, F, L, C, cycle, H = [int(v) for v in w["cfg"]]
    cfg = dict(in_dims=M, n_feats=F, num_layers=L, num_channels=C, dilation_cycle_length=cycle, hidden=H)
    return d, sd, cfg


def rel_l2(a, b):
    a, b = a.double(), b.double()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


def load_lynx_golden(name):
    """-> (dict of tensors, state_dict, cfg) for a tests/golden/lynxnet_*.npz fixture."""
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    d = {k: torch.from_numpy(z[k]) for k in ("spec", "cond", "t_long", "t_float", "y_long", "y_float")}
    sd = {k[2:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("w.")}
    M, F, L, C, ef, ks, H, strong = [int(v) for v in z["cfg"]]
    cfg = dict(in_dims=M, n_feats=F, num_layers=L, num_channels=C, expansion_factor=ef, kernel_size=ks, hidden=H,
               strong_cond=bool(strong), activation=str(z["activation"]))
    return d, sd, cfg


def load_nsf_golden(tag):
    """-> (dict of tensors, state_dict, cfg dict) for a tests/golden/nsf_hifigan_*.npz fixture."""
    import json
    z = np.load(os.path.join(GOLDEN, f"nsf_hifigan_{tag}.npz"))
    d = {k: torch.from_numpy(z[k]) for k in ("mel", "f0", "rand_ini", "noise", "har_source", "wav")}
    sd = {k[2:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("w.")}
    return d, sd, json.loads(str(z["cfg_json"]))


def centered_rel_l2(a, b):
    """rel-L2 of a - b against the VARIATION of b (outputs that sit on a large constant would flatter a plain rel-L2)."""
    a, b = a.double(), b.double()
    return float((a - b).norm() / (b - b.mean()).norm().clamp_min(1e-30))


def load_fs2_golden(tag):
    """-> (inputs dict, reference output, state_dict, hparams dict) for a tests/golden/fs2_*.npz fixture."""
    import json
    z = np.load(os.path.join(GOLDEN, f"fs2_{tag}.npz"))
    d = {k: torch.from_numpy(z[k]) for k in ("tokens", "mel2ph", "f0")}
    d.update({k[3:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("in.")})
    sd = {k[2:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("w.")}
    return d, torch.from_numpy(z["y"]), sd, json.loads(str(z["hp_json"]))
