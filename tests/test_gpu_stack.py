"""The whole-network kernel (csrc/stack_kernel.cuh, session path 'stack'): one launch per backbone evaluation for launch-bound
sessions (C = 256; batch-1 `.ds` segments, the pitch predictor).  Checked against the CPU oracle at the suite's per-evaluation
bf16 bar (2e-2), against the per-layer kernel chain (DSB200_STACK=0) to bf16 round-off, and end to end through the samplers
(final result rel-L2 <= 1e-2).  Cases are ordered from the simplest (one layer, one tile) to the real configurations so that
`-x` stops at the first construct that breaks."""
import pytest
import torch

import diffsinger_b200 as ds
from conftest import rel_l2
from oracle.wavenet_oracle import OracleWaveNet, random_state_dict
from oracle import samplers_oracle as so

pytestmark = pytest.mark.gpu
dev = "cuda"
EVAL_TOL = 2e-2
MEL_TOL = 1e-2


def _net(sd, M, F, L, C, cycle, H=256):
    ds.hparams.clear()
    ds.hparams.update(hidden_size=H)
    net = ds.WaveNet(M, F, num_layers=L, num_channels=C, dilation_cycle_length=cycle)
    net.load_state_dict(sd, strict=True)
    return net.eval().to(dev)


def _path(net):
    return next(reversed(net.engine().sessions.values())).path


@pytest.mark.parametrize("L,cycle,M,F,B,T,C", [(*c, 256) if len(c) == 6 else c for c in [
    (1, 1, 64, 1, 1, 128),     # head + one layer (no residual exchange) + tail, one full tile
    (2, 1, 64, 1, 1, 128),     # one residual exchange
    (4, 4, 128, 1, 1, 97),     # dilations 1..8, ragged single tile (frames >= T must stay zero: conv padding)
    (20, 5, 64, 1, 1, 128),    # pitch predictor depth, dilation 16
    (3, 2, 64, 1, 1, 130),     # two tiles: edge frames through global memory (second tile holds 2 frames)
    (20, 5, 64, 1, 1, 512),    # pitch predictor, four tiles
    (20, 4, 128, 1, 1, 517),   # config #1 backbone, five tiles
    (20, 4, 128, 1, 3, 300),   # batch of three utterances x three tiles
    (6, 3, 24, 2, 2, 200),     # F = 2, F*M = 48 (output rows padded to 64)
    (10, 4, 24, 2, 1, 128, 192),   # multi-variance predictor: C = 192 -> clusters of THREE CTAs
    (10, 4, 24, 2, 2, 300, 192),   # ... several tiles and utterances
    (3, 2, 64, 1, 1, 77, 192),
]])
def test_stack_forward_matches_oracle_and_chain(monkeypatch, L, cycle, M, F, B, T, C):
    sd = random_state_dict(M, F, L, C, 256, seed=L * 7 + T)
    g = torch.Generator().manual_seed(5)
    spec = torch.randn(B, F, M, T, generator=g)
    cond = torch.randn(B, T, 256, generator=g).transpose(1, 2)
    t = torch.tensor([617.25])
    ref = OracleWaveNet(sd, F, cycle)(spec, t, cond)
    net = _net(sd, M, F, L, C, cycle)
    with torch.no_grad():
        y = net(spec.to(dev), t.to(dev), cond.to(dev))
    torch.cuda.synchronize()
    assert _path(net) == "stack"
    sess = next(reversed(net.engine().sessions.values()))
    assert sess.last_launches == 1
    e_oracle = rel_l2(y.cpu(), ref)
    monkeypatch.setenv("DSB200_STACK", "0")
    net2 = _net(sd, M, F, L, C, cycle)
    with torch.no_grad():
        y2 = net2(spec.to(dev), t.to(dev), cond.to(dev))
    assert _path(net2) != "stack"
    e_chain = rel_l2(y.cpu(), y2.cpu())
    print(f"\n[stack C{C} L{L} cycle{cycle} M{M} F{F} B{B} T{T}] vs oracle {e_oracle:.3e}, vs kernel chain {e_chain:.3e} "
          f"(chain vs oracle {rel_l2(y2.cpu(), ref):.3e})")
    assert torch.isfinite(y).all()
    assert e_oracle < EVAL_TOL and e_chain < EVAL_TOL
    if B > 1:  # per-sample diffusion steps (training-style call, ddpm.py:212-219)
        monkeypatch.delenv("DSB200_STACK")
        tb = torch.linspace(3.0, 990.0, B)
        ref = OracleWaveNet(sd, F, cycle)(spec, tb, cond)
        with torch.no_grad():
            y = net(spec.to(dev), tb.to(dev), cond.to(dev))
        assert rel_l2(y.cpu(), ref) < EVAL_TOL


@pytest.mark.parametrize("T", [128, 300])
def test_pitch_predictor_ddim20_on_the_stack_kernel(T):
    """configs[3]: PitchDiffusion (WaveNet 20 x 256, cycle 5, 64 repeat bins), DDIM 20 steps, batch 1: 20 evaluation launches +
    20 sampler updates in one CUDA graph; final curve against the oracle sampler."""
    M, F, L, C, H, cycle = 64, 1, 20, 256, 256, 5
    sd = random_state_dict(M, F, L, C, H, seed=33)
    ds.hparams.clear()
    ds.hparams.update(hidden_size=H, schedule_type="linear", use_shallow_diffusion=False, diff_accelerator="ddim", diff_speedup=50,
                      K_step=1000, K_step_infer=1000, timesteps=1000)
    m = ds.PitchDiffusion(vmin=-8.0, vmax=8.0, cmin=-12.0, cmax=12.0, repeat_bins=M, timesteps=1000, k_step=1000,
                          backbone_type="wavenet", backbone_args=dict(num_layers=L, num_channels=C, dilation_cycle_length=cycle))
    m.denoise_fn.load_state_dict(sd)
    m = m.eval().to(dev)
    g = torch.Generator().manual_seed(3)
    cond = torch.randn(1, T, H, generator=g)
    noise = torch.randn(1, F, M, T, generator=g)
    outs = []
    for _ in range(2):  # capture, then replay
        m._noise_override = noise.to(dev)
        with torch.no_grad():
            outs.append(m(cond.to(dev), infer=True).cpu())
    m._noise_override = None
    assert torch.equal(outs[0], outs[1])
    sess = next(iter(m.denoise_fn.engine().sessions.values()))
    assert sess.path == "stack" and sess.last_launches == 40
    x = so.sample_ddim(OracleWaveNet(sd, F, cycle), so.Schedule(so.linear_betas(1000)), noise, cond.transpose(1, 2), 1000, 50)
    ref = so.finish(x, torch.tensor([-8.0]), torch.tensor([8.0])).mean(-1).clamp(-12.0, 12.0)
    err = rel_l2(outs[0], ref)
    print(f"\n[pitch predictor ddim-20, T={T}, stack kernel] final curve rel-L2 {err:.3e}")
    assert err < MEL_TOL


def test_multi_variance_predictor_euler20_on_three_cta_clusters():
    """configs[3], second sampler: MultiVarianceRectifiedFlow (WaveNet 10 x 192, 2 x 24 bins), Euler 20 steps, batch 1."""
    M, F, L, C, H, cycle, T = 24, 2, 10, 192, 256, 4, 300
    sd = random_state_dict(M, F, L, C, H, seed=34)
    ds.hparams.clear()
    ds.hparams.update(hidden_size=H, sampling_algorithm="euler", sampling_steps=20, T_start=0.0, T_start_infer=0.0,
                      use_shallow_diffusion=False, time_scale_factor=1000)
    m = ds.MultiVarianceRectifiedFlow(ranges=[(-96.0, -12.0), (-96.0, -20.0)], clamps=[(-96.0, 0.0), (-96.0, 0.0)], repeat_bins=M,
                                      backbone_type="wavenet", backbone_args=dict(num_layers=L, num_channels=C, dilation_cycle_length=cycle))
    m.velocity_fn.load_state_dict(sd)
    m = m.eval().to(dev)
    g = torch.Generator().manual_seed(4)
    cond = torch.randn(1, T, H, generator=g)
    noise = torch.randn(1, F, M, T, generator=g)
    m._noise_override = noise.to(dev)
    with torch.no_grad():
        ys = [v.cpu() for v in m(cond.to(dev), infer=True)]
    m._noise_override = None
    sess = next(iter(m.velocity_fn.engine().sessions.values()))
    assert sess.path == "stack" and sess.last_launches == 40
    from oracle import solvers_oracle as sv
    x = sv.sample_reflow(OracleWaveNet(sd, F, cycle), noise, cond.transpose(1, 2), 20, "euler")
    smin = torch.tensor([[-96.0], [-96.0]])[None, None].transpose(-3, -2)
    smax = torch.tensor([[-12.0], [-20.0]])[None, None].transpose(-3, -2)
    ref = torch.stack([v.clamp(-96.0, 0.0) for v in so.finish(x, smin, smax).mean(-1).unbind(1)])
    err = rel_l2(torch.stack(ys), ref)
    print(f"\n[multi-variance predictor euler-20, T={T}, stack kernel, 3-CTA clusters] final curves rel-L2 {err:.3e}")
    assert err < MEL_TOL


@pytest.mark.parametrize("family", ["acoustic_ddim", "pitch_euler"])
def test_ragged_batches_equal_segment_by_segment(family):
    """sample_segments packs batch-1 segments of different lengths into ragged batches (ds_session_set_lengths): every segment
    must come out BIT-identical to running it alone (same noise draw order, its own zero padding at its own end)."""
    H = 256
    lens = [300, 128, 515, 77, 257, 512]
    ds.hparams.clear()
    if family == "acoustic_ddim":
        M, F, L, C, cycle = 128, 1, 4, 256, 4
        ds.hparams.update(hidden_size=H, diff_speedup=100, diff_accelerator="ddim", timesteps=1000, K_step=1000, use_shallow_diffusion=False,
                          schedule_type="linear")
        m = ds.GaussianDiffusion(M, 1, 1000, 1000, "wavenet", dict(num_layers=L, num_channels=C, dilation_cycle_length=cycle),
                                 spec_min=[-12.0], spec_max=[0.0])
        m.denoise_fn.load_state_dict(random_state_dict(M, F, L, C, H, seed=91))
        net = m.denoise_fn
    else:
        M, F, L, C, cycle = 64, 1, 5, 256, 5
        ds.hparams.update(hidden_size=H, sampling_algorithm="euler", sampling_steps=10, T_start=0.0, T_start_infer=0.0,
                          use_shallow_diffusion=False, time_scale_factor=1000)
        m = ds.PitchRectifiedFlow(vmin=-8.0, vmax=8.0, cmin=-12.0, cmax=12.0, repeat_bins=M, backbone_type="wavenet",
                                  backbone_args=dict(num_layers=L, num_channels=C, dilation_cycle_length=cycle))
        m.velocity_fn.load_state_dict(random_state_dict(M, F, L, C, H, seed=92))
        net = m.velocity_fn
    m = m.eval().to(dev)
    g = torch.Generator().manual_seed(17)
    conds = [torch.randn(1, t, H, generator=g).to(dev) for t in lens]
    torch.manual_seed(5)
    with torch.no_grad():
        alone = [m(c, infer=True).cpu() for c in conds]
    torch.manual_seed(5)
    batched = [y.cpu() for y in ds.sample_segments(m, conds)]
    paths = {s.path for s in net.engine().sessions.values()}
    assert paths == {"stack"}
    assert any(s._lengths is not None for s in net.engine().sessions.values()), "the ragged path did not run"
    for t, a, b in zip(lens, alone, batched):
        assert a.shape == b.shape and a.shape[1] == t
        assert torch.equal(a, b), f"segment of {t} frames differs: {rel_l2(b, a):.3e}"
    # a ragged session falls back to full lengths when asked to
    torch.manual_seed(5)
    streams = [y.cpu() for y in ds.sample_segments(m, conds, ragged=False)]
    for a, b in zip(alone, streams):
        assert torch.equal(a, b)


@pytest.mark.parametrize("C,L,mode", [(256, 4, "chain"), (512, 3, "fused"), (512, 3, "unfused")])
def test_ragged_batches_on_the_per_layer_kernels(monkeypatch, C, L, mode):
    """ds_session_set_lengths off the whole-network kernel: the x~ writers of the per-layer kernels (head GEMM, residual GEMM, fused
    layer kernel) store zeros beyond an utterance's end, so each utterance of a padded batch sees its own conv edge.  The batch may
    run on different kernels than a lone short segment (tile shapes, f16 gate), so equality is to bf16 noise -- but it must hold on
    the LAST frames of every utterance, where a wrong edge would show."""
    monkeypatch.setenv("DSB200_STACK", "0")
    if mode != "chain":  # the big-batch kernels for every session size: the fused layer kernel / the 256-wide unfused GEMMs
        monkeypatch.setenv("DSB200_SMALL", "0")
        monkeypatch.setenv("DSB200_MEDIUM", "0")
    if mode == "unfused":
        monkeypatch.setenv("DSB200_LAYER_MODE", "unfused")
    M, F, H, cycle = 128, 1, 256, 3
    lens = [300, 77, 515, 512, 129, 640]
    ds.hparams.clear()
    ds.hparams.update(hidden_size=H, diff_speedup=200, diff_accelerator="ddim", timesteps=1000, K_step=1000, use_shallow_diffusion=False,
                      schedule_type="linear")
    m = ds.GaussianDiffusion(M, 1, 1000, 1000, "wavenet", dict(num_layers=L, num_channels=C, dilation_cycle_length=cycle),
                             spec_min=[-12.0], spec_max=[0.0])
    m.denoise_fn.load_state_dict(random_state_dict(M, F, L, C, H, seed=95))
    m = m.eval().to(dev)
    g = torch.Generator().manual_seed(18)
    conds = [torch.randn(1, t, H, generator=g).to(dev) for t in lens]
    torch.manual_seed(6)
    with torch.no_grad():
        alone = [m(c, infer=True).cpu() for c in conds]
    torch.manual_seed(6)
    batched = [y.cpu() for y in ds.sample_segments(m, conds)]
    sessions = m.denoise_fn.engine().sessions
    assert any(s._lengths is not None for s in sessions.values()), "the ragged path did not run"
    assert {s.path for s in sessions.values()} == {"small" if mode == "chain" else "fused"}
    launches = {s.last_launches for s in sessions.values()}
    assert launches == ({5 * (L + 5) + 5} if mode == "fused" else {5 * (2 * L + 4) + 5}), launches  # DDIM-5; one kernel per layer when fused
    worst = 0.0
    for t, a, b in zip(lens, alone, batched):
        assert a.shape == b.shape and a.shape[1] == t
        worst = max(worst, rel_l2(b, a), rel_l2(b[:, -24:], a[:, -24:]), rel_l2(b[:, :24], a[:, :24]))
    print(f"\n[ragged batch, C={C}, {mode}] worst rel-L2 vs segment by segment (whole / last 24 / first 24 frames): {worst:.3e}")
    assert worst < 5e-3


def test_lengths_are_validated():
    M, F, L, C, H, cycle = 64, 1, 3, 256, 256, 3
    net2 = _net(random_state_dict(M, F, L, C, H, seed=93), M, F, L, C, cycle)
    sess2 = net2.engine().session(2, 200)
    assert sess2.path == "stack"
    with pytest.raises(Exception, match="outside"):
        sess2.set_lengths([100, 201])
    with pytest.raises(Exception, match="n=1"):
        sess2.set_lengths([100])
