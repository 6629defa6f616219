"""Pins the CPU oracle (oracle/*.py) against golden vectors produced by the unmodified reference
(oracle/make_golden.py).  Tolerance: the oracle restates the same fp32 torch ops, so agreement is
expected at fp32 round-off (1e-5 rel-L2 leaves room for op-order differences in the solvers)."""
import pytest
import torch

from conftest import load_golden, rel_l2
from oracle.wavenet_oracle import OracleWaveNet, wavenet_forward
from oracle import samplers_oracle as so
from oracle import solvers_oracle as sv

TOL = 1e-5
SPEC_MIN, SPEC_MAX = torch.tensor([-12.0]), torch.tensor([0.0])


@pytest.mark.parametrize("tag", ["f1", "f2"])
def test_wavenet_forward(tag):
    d, sd, cfg = load_golden("wavenet_" + tag)
    for tk, yk in (("t_long", "y_long"), ("t_float", "y_float")):
        y = wavenet_forward(sd, d["spec"], d[tk], d["cond"], n_feats=cfg["n_feats"],
                            dilation_cycle_length=cfg["dilation_cycle_length"])
        assert y.shape == d[yk].shape
        assert rel_l2(y, d[yk]) < 1e-6


def _model(sd, cfg):
    return OracleWaveNet(sd, cfg["n_feats"], cfg["dilation_cycle_length"])


def _start(d, sch, t_max):
    """x at loop entry (ddpm.py:227-242)."""
    noise = d["noise"]
    if "src" in d:
        src = (d["src"] - SPEC_MIN) / (SPEC_MAX - SPEC_MIN) * 2 - 1
        return so.q_sample(sch, src.transpose(-2, -1)[:, None], t_max - 1, noise)
    return noise


@pytest.mark.parametrize("name", ["ddim", "pndm", "dpm_solver_s50", "dpm_solver_s125", "unipc", "ddpm", "ddim_shallow",
                                  # round 2: schedule_type='cosine' (ddpm.py:36-46; lambda-clip active in DPM-Solver++ only, trap T9)
                                  # and the solvers over a shallow schedule betas[:400] (ddpm.py:246-322)
                                  "ddim_cosine", "dpm_solver_cosine", "unipc_cosine", "dpm_solver_shallow", "unipc_shallow"])
def test_gaussian_diffusion(name):
    d, sd, cfg = load_golden("ddpm_" + name)
    M, F, timesteps, k_step, speedup, shallow = [int(v) for v in d["meta"]]
    sch = so.Schedule(so.cosine_betas(timesteps) if name.endswith("cosine") else so.linear_betas(timesteps))
    model = _model(sd, cfg)
    cond = d["cond"].transpose(1, 2)
    t_max = k_step
    x = _start(d, sch, t_max)
    if name.startswith("ddim"):
        x = so.sample_ddim(model, sch, x, cond, t_max, speedup)
    elif name == "pndm":
        x = so.sample_pndm(model, sch, x, cond, t_max, speedup)
    elif name == "ddpm":
        x = so.sample_ddpm(model, sch, x, cond, t_max, d["step_noise"])
    elif name.startswith("dpm"):
        x = sv.sample_dpm_solver_pp(model, sch.betas[:t_max], x, cond, t_max // speedup)
    else:
        x = sv.sample_unipc(model, sch.betas[:t_max], x, cond, t_max // speedup)
    y = so.finish(x, SPEC_MIN, SPEC_MAX)
    assert y.shape == d["y"].shape
    assert rel_l2(y, d["y"]) < TOL
    expect = {"pndm": t_max // speedup + 1}.get(name, t_max // max(speedup, 1))
    assert model.calls == expect  # NFE as counted on the reference


@pytest.mark.parametrize("name", ["euler", "rk2", "rk4", "rk5", "euler_shallow"])
def test_rectified_flow(name):
    d, sd, cfg = load_golden("reflow_" + name)
    M, F, steps, shallow = [int(v) for v in d["meta"]]
    t_start = float(d["t_start"])
    model = _model(sd, cfg)
    x = d["noise"]
    if shallow:
        src = (d["src"] - SPEC_MIN) / (SPEC_MAX - SPEC_MIN) * 2 - 1
        x = t_start * src.transpose(-2, -1)[:, None] + (1 - t_start) * x
    x = sv.sample_reflow(model, x, d["cond"].transpose(1, 2), steps, name.split("_")[0], t_start)
    assert rel_l2(so.finish(x, SPEC_MIN, SPEC_MAX), d["y"]) < TOL


def test_variance_wrappers():
    # PitchDiffusion: ddim 10 steps, denorm -> mean over repeat bins -> clamp (ddpm.py:386-445)
    d, sd, cfg = load_golden("var_pitch_ddim")
    sch = so.Schedule(so.linear_betas(1000))
    x = so.sample_ddim(_model(sd, cfg), sch, d["noise"], d["cond"].transpose(1, 2), 1000, 100)
    y = so.finish(x, torch.tensor([-8.0]), torch.tensor([8.0])).mean(-1).clamp(-12.0, 12.0)
    assert rel_l2(y, d["y"]) < TOL
    # MultiVarianceRectifiedFlow: euler 10, F=2 (reflow.py:206-261)
    d, sd, cfg = load_golden("var_multi_euler")
    x = sv.sample_reflow(_model(sd, cfg), d["noise"], d["cond"].transpose(1, 2), 10, "euler")
    smin = torch.tensor([[-96.0], [-96.0]])[None, None].transpose(-3, -2)
    smax = torch.tensor([[-12.0], [-20.0]])[None, None].transpose(-3, -2)
    y = so.finish(x, smin, smax).mean(-1)  # [B,F,T]
    ys = torch.stack([v.clamp(-96.0, 0.0) for v in y.unbind(1)])
    assert rel_l2(ys, d["y"]) < TOL
    # MultiVarianceDiffusion: ddim 10, F=2, second curve unclamped (ddpm.py:448-505)
    d, sd, cfg = load_golden("var_multi_ddim")
    x = so.sample_ddim(_model(sd, cfg), sch, d["noise"], d["cond"].transpose(1, 2), 1000, 100)
    y = so.finish(x, smin, smax).mean(-1)
    a, b = y.unbind(1)
    assert rel_l2(torch.stack([a.clamp(-96.0, 0.0), b]), d["y"]) < TOL
    # PitchRectifiedFlow: euler 10 (reflow.py:184-203)
    d, sd, cfg = load_golden("var_pitch_euler")
    x = sv.sample_reflow(_model(sd, cfg), d["noise"], d["cond"].transpose(1, 2), 10, "euler")
    y = so.finish(x, torch.tensor([-8.0]), torch.tensor([8.0])).mean(-1).clamp(-12.0, 12.0)
    assert rel_l2(y, d["y"]) < TOL


@pytest.mark.parametrize("tag", ["prelu_f1", "silu_f2", "relu_strong"])
def test_lynxnet_forward(tag):
    """The LYNXNet restatement against the imported reference's outputs (modules/backbones/lynxnet.py:133-172)."""
    from conftest import load_lynx_golden
    from oracle.lynxnet_oracle import OracleLYNXNet
    d, sd, cfg = load_lynx_golden("lynxnet_" + tag)
    net = OracleLYNXNet(sd, cfg["n_feats"], cfg["kernel_size"], cfg["activation"], cfg["strong_cond"])
    assert net.num_layers == cfg["num_layers"]
    for tk, yk in (("t_long", "y_long"), ("t_float", "y_float")):
        y = net(d["spec"], d[tk], d["cond"])
        assert y.shape == d[yk].shape
        assert rel_l2(y, d[yk]) < 2e-6


@pytest.mark.parametrize("tag", ["f1", "f2"])
def test_aux_decoder(tag):
    """The ConvNeXt aux decoder restatement against the imported reference's outputs (modules/aux_decoder)."""
    import os
    import numpy as np
    from conftest import GOLDEN
    from oracle.auxdec_oracle import aux_decoder
    z = np.load(os.path.join(GOLDEN, f"auxdec_{tag}.npz"))
    H, M, Fn, C, L, k = [int(v) for v in z["cfg"]]
    sd = {key[2:]: torch.from_numpy(z[key]) for key in z.files if key.startswith("w.")}
    cond = torch.from_numpy(z["cond"])
    y = aux_decoder(sd, cond, Fn, M, z["spec_min"], z["spec_max"], infer=True)
    assert y.shape == z["y"].shape and rel_l2(y, torch.from_numpy(z["y"])) < 2e-6
    y = aux_decoder(sd, cond, Fn, M, infer=False)
    assert rel_l2(y, torch.from_numpy(z["y_raw"])) < 2e-6
