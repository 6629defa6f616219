"""GPU parity of the native LYNXNet path (diffsinger_b200.LYNXNet: primitive ops of libdsb200.so composed per layer) against
  (1) golden vectors produced by the unmodified reference (tests/golden/lynxnet_*.npz),
  (2) the CPU oracle at the default acoustic size (6 x 512, expansion 2, k = 31, configs/acoustic.yaml:73-79), and
  (3) the oracle samplers run with the oracle LYNXNet (final mel).
bf16 path tolerances as for WaveNet: per evaluation rel-L2 <= 2e-2, final mel <= 1e-2."""
import pytest
import torch

import diffsinger_b200 as ds
from conftest import load_lynx_golden, rel_l2
from diffsinger_b200.synthetic import random_lynx_state_dict
from oracle.lynxnet_oracle import OracleLYNXNet
from oracle import samplers_oracle as so, solvers_oracle as sv

pytestmark = pytest.mark.gpu
dev = "cuda"
EVAL_TOL = 2e-2
MEL_TOL = 1e-2


def _net(sd, cfg):
    ds.hparams.clear()
    ds.hparams.update(hidden_size=cfg["hidden"])
    net = ds.build_backbone(cfg["in_dims"], cfg["n_feats"], "lynxnet",
                            dict(num_layers=cfg["num_layers"], num_channels=cfg["num_channels"], expansion_factor=cfg["expansion_factor"],
                                 kernel_size=cfg["kernel_size"], activation=cfg["activation"], strong_cond=cfg["strong_cond"]))
    net.load_state_dict(sd, strict=True)
    return net.eval().to(dev)


@pytest.mark.parametrize("tag", ["prelu_f1", "silu_f2", "relu_strong"])
def test_forward_matches_reference_golden(tag):
    d, sd, cfg = load_lynx_golden("lynxnet_" + tag)
    net = _net(sd, cfg)
    with torch.no_grad():
        for tk, yk in (("t_long", "y_long"), ("t_float", "y_float")):
            y = net(d["spec"].to(dev), d[tk].to(dev), d["cond"].to(dev))
            assert y.shape == d[yk].shape
            assert rel_l2(y.cpu(), d[yk]) < EVAL_TOL


@pytest.mark.parametrize("C,L,M,F,k,act,strong,B,T", [
    (512, 6, 128, 1, 31, "PReLU", False, 2, 300),  # default acoustic backbone
    (256, 3, 64, 1, 31, "SiLU", True, 3, 131),
    (192, 2, 24, 2, 7, "ReLU", False, 1, 77),
    (1024, 6, 128, 1, 31, "PReLU", True, 2, 300),  # configs/acoustic.yaml:72-78 as shipped: 6 x 1024, strong_cond (folded LayerNorm)
])
def test_forward_matches_oracle_real_sizes(C, L, M, F, k, act, strong, B, T):
    sd = random_lynx_state_dict(M, F, L, C, 256, expansion_factor=2, kernel_size=k, activation=act, seed=C + L)
    cfg = dict(in_dims=M, n_feats=F, num_layers=L, num_channels=C, expansion_factor=2, kernel_size=k, activation=act,
               strong_cond=strong, hidden=256)
    net = _net(sd, cfg)
    oracle = OracleLYNXNet(sd, F, k, act, strong)
    g = torch.Generator().manual_seed(5)
    spec = torch.randn(B, F, M, T, generator=g)
    cond = torch.randn(B, T, 256, generator=g).transpose(1, 2)
    for t in (torch.tensor([617.25]), torch.linspace(3.0, 990.0, B)):
        if t.numel() > 1 and B == 1:
            continue
        ref = oracle(spec, t, cond)
        with torch.no_grad():
            y = net(spec.to(dev), t.to(dev), cond.to(dev))
        err = rel_l2(y.cpu(), ref)
        print(f"lynxnet C={C} L={L} k={k} {act} rel-L2 {err:.3e}")
        assert err < EVAL_TOL


@pytest.mark.parametrize("algo", ["ddim", "dpm-solver", "euler"])
def test_samplers_with_lynxnet(algo):
    M, F, L, C, H, k, B, T = 128, 1, 6, 512, 256, 31, 2, 200
    sd = random_lynx_state_dict(M, F, L, C, H, expansion_factor=2, kernel_size=k, activation="PReLU", seed=3)
    ds.hparams.clear()
    ds.hparams.update(hidden_size=H, schedule_type="linear", use_shallow_diffusion=False, diff_accelerator=algo, diff_speedup=100,
                      K_step=1000, K_step_infer=1000, timesteps=1000, sampling_algorithm="euler", sampling_steps=10)
    args = dict(num_layers=L, num_channels=C, expansion_factor=2, kernel_size=k)
    if algo == "euler":
        m = ds.RectifiedFlow(M, F, 0.0, 1000, "lynxnet", args, spec_min=[-12.0], spec_max=[0.0])
        bb = m.velocity_fn
    else:
        m = ds.GaussianDiffusion(M, F, 1000, 1000, "lynxnet", args, spec_min=[-12.0], spec_max=[0.0])
        bb = m.denoise_fn
    bb.load_state_dict(sd)
    m = m.eval().to(dev)
    cond = torch.randn(B, T, H, generator=torch.Generator().manual_seed(1))
    torch.manual_seed(7)
    with torch.no_grad():
        y = m(cond.to(dev), infer=True)
        torch.manual_seed(7)
        y2 = m(cond.to(dev), infer=True)  # second call replays the captured graph
    assert torch.equal(y, y2)
    torch.manual_seed(7)
    noise = torch.randn(B, F, M, T, device=dev).cpu()
    oracle = OracleLYNXNet(sd, F, k, "PReLU", False)
    if algo == "ddim":
        x = so.sample_ddim(oracle, so.Schedule(so.linear_betas(1000)), noise, cond.transpose(1, 2), 1000, 100)
    elif algo == "dpm-solver":
        x = sv.sample_dpm_solver_pp(oracle, so.Schedule(so.linear_betas(1000)).betas, noise, cond.transpose(1, 2), 10)
    else:
        x = sv.sample_reflow(oracle, noise, cond.transpose(1, 2), 10, "euler", 0.0, 1000.0)
    ref = so.finish(x, torch.tensor([-12.0]), torch.tensor([0.0]))
    err = rel_l2(y.cpu(), ref)
    print(f"lynxnet {algo} 10 steps final mel rel-L2 {err:.3e}")
    assert err < MEL_TOL
    ws = next(iter(bb.engine().workspaces.values()))
    prog = next(iter(m._programs.values()))
    # strong_cond = False: the conditioner projections (L GEMMs) are hoisted out of the loop, 4 launches per layer inside it
    assert prog.nfe == 10 and ws.last_launches == L + prog.nfe * (4 * L + 4) + (len(prog.ops) - prog.nfe)


def test_lynxnet_batch_chunking_is_bitwise_identical(monkeypatch):
    """LynxEngine.max_chunk_batch: a batch that would not fit runs as consecutive chunks; utterances are independent end to
    end (LayerNorm, depthwise conv and the 1x1 GEMMs are per frame / per utterance, lynxnet.py:82-99), so the rows agree bit
    for bit."""
    M, F, L, C, H, k, B, T = 128, 1, 2, 256, 256, 31, 5, 150
    sd = random_lynx_state_dict(M, F, L, C, H, expansion_factor=2, kernel_size=k, activation="PReLU", seed=4)
    ds.hparams.clear()
    ds.hparams.update(hidden_size=H, schedule_type="linear", use_shallow_diffusion=False, diff_accelerator="dpm-solver",
                      diff_speedup=100, K_step=1000, K_step_infer=1000, timesteps=1000)
    m = ds.GaussianDiffusion(M, F, 1000, 1000, "lynxnet", dict(num_layers=L, num_channels=C, expansion_factor=2, kernel_size=k),
                             spec_min=[-12.0], spec_max=[0.0])
    m.denoise_fn.load_state_dict(sd)
    m = m.eval().to(dev)
    cond = torch.randn(B, T, H, generator=torch.Generator().manual_seed(1)).to(dev)
    outs = []
    for cap in (None, str(2 * T), str(T)):
        if cap is None:
            monkeypatch.delenv("DSB200_MAX_FRAMES", raising=False)
        else:
            monkeypatch.setenv("DSB200_MAX_FRAMES", cap)
        torch.manual_seed(3)
        with torch.no_grad():
            outs.append(m(cond, infer=True).cpu())
    assert torch.equal(outs[0], outs[1]) and torch.equal(outs[0], outs[2])


@pytest.mark.parametrize("C,strong,shallow", [(256, True, True), (512, False, False)])
def test_ragged_batches_of_segments(C, strong, shallow):
    """sample_segments with LYNXNet: segments of different lengths padded into ragged batches; the SwiGLU GEMM stores zeros beyond an
    utterance's end, so the depthwise k = 31 conv sees each utterance's own padding.  Shallow diffusion sources ride along
    (src_spec=[...]).  The batch may run on other GEMM variants than a lone short segment (pair / split-K kernels): equality to bf16
    noise, also on the LAST frames of every utterance, where a wrong edge would show."""
    M, F, L, H, k = 128, 1, 3, 256, 31
    sd = random_lynx_state_dict(M, F, L, C, H, expansion_factor=2, kernel_size=k, activation="PReLU", seed=C)
    ds.hparams.clear()
    ds.hparams.update(hidden_size=H, use_shallow_diffusion=shallow, sampling_algorithm="euler", sampling_steps=6, T_start=0.4 if shallow else 0.0,
                      T_start_infer=0.4 if shallow else 0.0, time_scale_factor=1000)
    m = ds.RectifiedFlow(M, 1, t_start=0.4 if shallow else 0.0, backbone_type="lynxnet",
                         backbone_args=dict(num_layers=L, num_channels=C, expansion_factor=2, kernel_size=k, strong_cond=strong),
                         spec_min=[-12.0], spec_max=[0.0])
    m.velocity_fn.load_state_dict(sd)
    m = m.eval().to(dev)
    lens = [300, 77, 515, 257, 512, 640]
    g = torch.Generator().manual_seed(21)
    conds = [torch.randn(1, t, H, generator=g).to(dev) for t in lens]
    srcs = [(torch.rand(1, t, M, generator=g) * 12 - 12).to(dev) for t in lens] if shallow else None
    kw = dict(src_spec=srcs) if shallow else {}
    torch.manual_seed(8)
    with torch.no_grad():
        alone = [m(c, infer=True, **({"src_spec": srcs[i]} if shallow else {})).cpu() for i, c in enumerate(conds)]
    torch.manual_seed(8)
    batched = [y.cpu() for y in ds.sample_segments(m, conds, **kw)]
    eng = m.velocity_fn.engine()
    assert any(ws.lengths is not None for ws in eng.workspaces.values()), "the ragged path did not run"
    worst = 0.0
    for t, a, b in zip(lens, alone, batched):
        assert a.shape == b.shape and a.shape[1] == t
        worst = max(worst, rel_l2(b, a), rel_l2(b[:, -24:], a[:, -24:]), rel_l2(b[:, :24], a[:, :24]))
    print(f"[LYNXNet ragged, C={C}, strong={strong}, shallow={shallow}] worst rel-L2 vs one by one: {worst:.3e}")
    assert worst < 1e-2
    # the test can see a wrong edge: the same padded batch WITHOUT lengths changes the last frames of a short utterance
    if not shallow:
        t_max, i = max(lens), 1
        cond = torch.zeros(2, t_max, H, device=dev)
        cond[0, :lens[i]] = conds[i][0]
        cond[1] = conds[lens.index(t_max)][0]
        noise = torch.randn(2, F, M, t_max, generator=g).to(dev)
        m._noise_override = noise
        try:
            with torch.no_grad():
                padded = m(cond, infer=True).cpu()
            m._noise_override = noise[:1, ..., :lens[i]].contiguous()
            with torch.no_grad():
                lone = m(conds[i], infer=True).cpu()
        finally:
            m._noise_override = None
        assert rel_l2(padded[0:1, lens[i] - 12:lens[i]], lone[:, -12:]) > 1e-2
        assert rel_l2(padded[0:1, :lens[i] - 40], lone[:, :-40]) < 1e-2
