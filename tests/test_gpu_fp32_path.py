"""GPU parity of the fp32-accurate path (precision='fp32': 3 x bf16 split products on the tcgen05 GEMMs) against the
reference's golden vectors and the CPU fp32 oracle.  Tolerance (BASELINE.json north_star): per-step denoiser output
relative L2 <= 1e-4.  Also checks the sampler end to end in that mode (final mel, same bar)."""
import pytest
import torch

import diffsinger_b200 as ds
from conftest import load_golden, rel_l2
from oracle.wavenet_oracle import OracleWaveNet, random_state_dict
from oracle import samplers_oracle as so, solvers_oracle as sv

pytestmark = pytest.mark.gpu
dev = "cuda"
STEP_TOL = 1e-4


def _net(sd, cfg):
    ds.hparams.clear()
    ds.hparams.update(hidden_size=cfg["hidden"])
    net = ds.WaveNet(cfg["in_dims"], cfg["n_feats"], num_layers=cfg["num_layers"], num_channels=cfg["num_channels"],
                     dilation_cycle_length=cfg["dilation_cycle_length"])
    net.load_state_dict(sd, strict=True)
    return net.set_precision("fp32").eval().to(dev)


@pytest.mark.parametrize("tag", ["f1", "f2"])
def test_forward_matches_reference_golden_fp32(tag):
    d, sd, cfg = load_golden("wavenet_" + tag)
    net = _net(sd, cfg)
    from diffsinger_b200 import _native as nat
    with torch.no_grad():
        for tk, yk in (("t_long", "y_long"), ("t_float", "y_float")):
            y = net(d["spec"].to(dev), d[tk].to(dev), d["cond"].to(dev))
            assert rel_l2(y.cpu(), d[yk]) < STEP_TOL
    assert nat.lib().ds_plan_precision(net.engine().plan) == 1


@pytest.mark.parametrize("C,L,cycle,M,F,B,T", [
    (512, 20, 4, 128, 1, 2, 300),   # acoustic config #2/#3 backbone
    (256, 20, 4, 128, 1, 1, 517),   # config #1
    (256, 20, 5, 64, 1, 1, 130),    # pitch predictor
    (192, 10, 4, 24, 2, 3, 97),     # multi-variance predictor
])
def test_forward_matches_oracle_real_sizes_fp32(C, L, cycle, M, F, B, T):
    sd = random_state_dict(M, F, L, C, 256, seed=C + L)
    cfg = dict(in_dims=M, n_feats=F, num_layers=L, num_channels=C, dilation_cycle_length=cycle, hidden=256)
    net = _net(sd, cfg)
    g = torch.Generator().manual_seed(5)
    spec = torch.randn(B, F, M, T, generator=g)
    cond = torch.randn(B, T, 256, generator=g).transpose(1, 2)
    for t in (torch.tensor([617.25]), torch.tensor([3.0]), torch.linspace(3.0, 990.0, B)):
        if t.numel() > 1 and B == 1:
            continue
        ref = OracleWaveNet(sd, F, cycle)(spec, t, cond)
        with torch.no_grad():
            y = net(spec.to(dev), t.to(dev), cond.to(dev))
        err = rel_l2(y.cpu(), ref)
        print(f"fp32 path C={C} L={L} t={t.tolist()[:1]} rel-L2 {err:.3e}")
        assert err < STEP_TOL


def test_precision_switch_repacks_and_bf16_differs():
    d, sd, cfg = load_golden("wavenet_f1")
    net = _net(sd, cfg)
    args = (d["spec"].to(dev), d["t_float"].to(dev), d["cond"].to(dev))
    with torch.no_grad():
        y32 = net(*args).cpu()
        net.set_precision("bf16")
        y16 = net(*args).cpu()
        net.set_precision("fp32")
        y32b = net(*args).cpu()
    assert torch.equal(y32, y32b)
    assert rel_l2(y32, d["y_float"]) < STEP_TOL < rel_l2(y16, d["y_float"]) < 2e-2
    with pytest.raises(ValueError):
        net.set_precision("fp16")


@pytest.mark.parametrize("algo", ["ddim", "dpm-solver"])
def test_sampler_fp32_path_final_mel(algo):
    """Whole 20-step run at config #1's backbone in fp32-accurate mode: every step stays within 1e-4, so does the mel."""
    M, F, L, C, H, cycle, B, T = 128, 1, 20, 256, 256, 4, 1, 300
    sd = random_state_dict(M, F, L, C, H, seed=21)
    ds.hparams.clear()
    ds.hparams.update(hidden_size=H, schedule_type="linear", use_shallow_diffusion=False, diff_accelerator=algo,
                      diff_speedup=50, K_step=1000, K_step_infer=1000, timesteps=1000)
    m = ds.GaussianDiffusion(M, F, 1000, 1000, "wavenet", dict(num_layers=L, num_channels=C, dilation_cycle_length=cycle),
                             spec_min=[-12.0], spec_max=[0.0])
    m.denoise_fn.load_state_dict(sd)
    m.denoise_fn.set_precision("fp32")
    m = m.eval().to(dev)
    cond = torch.randn(B, T, H, generator=torch.Generator().manual_seed(1))
    torch.manual_seed(7)
    with torch.no_grad():
        y = m(cond.to(dev), infer=True)
    torch.manual_seed(7)
    noise = torch.randn(B, F, M, T, device=dev).cpu()
    model = OracleWaveNet(sd, F, cycle)
    sch = so.Schedule(so.linear_betas(1000))
    if algo == "ddim":
        x = so.sample_ddim(model, sch, noise, cond.transpose(1, 2), 1000, 50)
    else:
        x = sv.sample_dpm_solver_pp(model, sch.betas, noise, cond.transpose(1, 2), 20)
    ref = so.finish(x, torch.tensor([-12.0]), torch.tensor([0.0]))
    err = rel_l2(y.cpu(), ref)
    print(f"fp32 path {algo} 20 steps final mel rel-L2 {err:.3e}")
    assert err < STEP_TOL


@pytest.mark.parametrize("hoist", ["1", "0"])
@pytest.mark.parametrize("C,L,cycle,B,T", [(512, 5, 4, 2, 300), (256, 6, 5, 3, 517), (512, 3, 3, 1, 77)])
def test_fused_layer_kernel_split_instantiation(monkeypatch, C, L, cycle, B, T, hoist):
    """ds_layer_grp_kernel<true>: the fused residual-layer kernel with 3 x bf16 split products, accurate gate math and hi / lo
    outputs -- forced here for small sessions (by default it is taken from one wave of tile pairs on), same 1e-4 bar; a ragged batch
    must still give every utterance its own conv edge."""
    monkeypatch.setenv("DSB200_FP32_FUSED", "1")
    monkeypatch.setenv("DSB200_FP32_HOIST", hoist)  # conditioner projection once per call (fp32 rows added in the gate epilogue) or in-kernel
    monkeypatch.setenv("DSB200_STACK", "0")
    M, F = 128, 1
    sd = random_state_dict(M, F, L, C, 256, seed=C + L + 1)
    cfg = dict(in_dims=M, n_feats=F, num_layers=L, num_channels=C, dilation_cycle_length=cycle, hidden=256)
    net = _net(sd, cfg)
    g = torch.Generator().manual_seed(6)
    spec = torch.randn(B, F, M, T, generator=g)
    cond = torch.randn(B, T, 256, generator=g).transpose(1, 2)
    for t in (torch.tensor([617.25]), torch.linspace(3.0, 990.0, B)):
        if t.numel() > 1 and B == 1:
            continue
        ref = OracleWaveNet(sd, F, cycle)(spec, t, cond)
        with torch.no_grad():
            y = net(spec.to(dev), t.to(dev), cond.to(dev))
        err = rel_l2(y.cpu(), ref)
        print(f"fp32 fused C={C} L={L} t={t.tolist()[:1]} rel-L2 {err:.3e}")
        assert err < STEP_TOL
    sess = next(reversed(net.engine().sessions.values()))
    assert sess.path == "fused" and sess.last_launches == L + 5  # one kernel per residual layer
    if B > 1:  # ragged: utterance 1 is shorter
        lens = [T, T - 41] + [T] * (B - 2)
        sess.set_lengths(lens)
        x = spec.reshape(B, F * M, T).transpose(1, 2).contiguous().to(dev)
        out = sess.forward(x, 0)
        torch.cuda.synchronize()
        ref1 = OracleWaveNet(sd, F, cycle)(spec[1:2, ..., :lens[1]], torch.linspace(3.0, 990.0, B)[1:2], cond[1:2, :, :lens[1]])
        got = out[1, :lens[1]].t().reshape(1, F, M, lens[1]).cpu()
        assert rel_l2(got, ref1) < STEP_TOL
