"""Edge cases of the product path on the GPU against the CPU oracle: very short and tile-boundary sequence lengths (shorter
than the largest dilation; one frame below / at / above the 128- and 256-frame tiles), on the fused CTA-pair kernel and
on the small-problem GEMM path; ragged batches (each length its own session, padded frames are ordinary frames in the
reference: SURVEY.md trap T4); argument validation surfacing as NativeError instead of a CPU fallback."""
import pytest
import torch

import diffsinger_b200 as ds
from conftest import rel_l2
from diffsinger_b200 import _native as nat
from oracle.wavenet_oracle import OracleWaveNet, random_state_dict

pytestmark = pytest.mark.gpu
dev = "cuda"
EVAL_TOL = 2e-2


def _net(sd, C, L, cycle, M, F, H=256):
    ds.hparams.clear()
    ds.hparams.update(hidden_size=H)
    net = ds.WaveNet(M, F, num_layers=L, num_channels=C, dilation_cycle_length=cycle)
    net.load_state_dict(sd, strict=True)
    return net.eval().to(dev)


@pytest.mark.parametrize("path", ["stack", "small", "fused"])
@pytest.mark.parametrize("T", [1, 5, 17, 127, 128, 129, 255, 256, 257, 383, 600])
def test_sequence_length_edges(T, path, monkeypatch):
    """cycle 5 -> dilations up to 16, so T = 1, 5 sit entirely inside the conv's zero padding.  Every kernel family that can run
    this shape: the whole-network cluster kernel (default here), the narrow-tile GEMM chain, one fused kernel per layer."""
    if path != "stack":
        monkeypatch.setenv("DSB200_STACK", "0")
        monkeypatch.setenv("DSB200_SMALL", "1" if path == "small" else "0")
    C, L, cycle, M, F, B = 256, 6, 5, 64, 1, 3
    sd = random_state_dict(M, F, L, C, 256, seed=31)
    net = _net(sd, C, L, cycle, M, F)
    g = torch.Generator().manual_seed(T)
    spec = torch.randn(B, F, M, T, generator=g)
    cond = torch.randn(B, T, 256, generator=g).transpose(1, 2)
    t = torch.tensor([250.5])
    ref = OracleWaveNet(sd, F, cycle)(spec, t, cond)
    with torch.no_grad():
        y = net(spec.to(dev), t.to(dev), cond.to(dev)).cpu()
    assert next(iter(net.engine().sessions.values())).path == path
    assert y.shape == ref.shape and torch.isfinite(y).all()
    assert rel_l2(y, ref) < EVAL_TOL


def test_session_graph_cache_is_bounded():
    """ds_session_run keys its captured graphs by (ops, buffer addresses); a caller that brings fresh buffers every call must not
    grow the cache without bound (kMaxSessionGraphs = 16, least recently used entry recycled) -- and results stay right."""
    C, L, cycle, M, F, B, T = 256, 2, 2, 64, 1, 1, 96
    sd = random_state_dict(M, F, L, C, 256, seed=3)
    net = _net(sd, C, L, cycle, M, F)
    eng = net.engine()
    sess = eng.session(B, T, max_times=4)
    from diffsinger_b200.engine import Program
    prog = Program(2)
    ti = prog.add_time(321.0)
    prog.eval(1, 0, ti)
    prog.lincomb(0, [(0.5, 0), (-0.25, 1)])
    g = torch.Generator().manual_seed(1)
    cond = torch.randn(B, T, 256, generator=g).to(dev)
    x0 = torch.randn(B, T, F * M, generator=g).to(dev)
    sess.set_cond(cond)
    sess.set_program_times(prog.times)
    want = None
    keep = []
    for i in range(40):  # 40 distinct buffer pairs > 16 cached graphs
        bufs = [x0.clone(), torch.empty_like(x0)]
        keep.append(bufs)
        sess.run(prog, bufs, use_graph=True)
        if want is None:
            ref = [x0.clone(), torch.empty_like(x0)]
            sess.run(prog, ref, use_graph=False)
            want = ref[0].clone()
        assert torch.equal(bufs[0], want)
    sess.run(prog, keep[0], use_graph=True)  # long evicted: captured again
    torch.cuda.synchronize()


def test_ragged_lengths_one_session_each():
    """A ragged set of utterances run one by one (what ds_acoustic.py does with segments): every length gets its own
    session / graph, results match the oracle per utterance, and re-running an earlier length reuses its session."""
    C, L, cycle, M, F = 256, 4, 4, 128, 1
    sd = random_state_dict(M, F, L, C, 256, seed=5)
    net = _net(sd, C, L, cycle, M, F)
    oracle = OracleWaveNet(sd, F, cycle)
    t = torch.tensor([10.0])
    outs = {}
    for T in (286, 504, 286, 1197, 504):
        g = torch.Generator().manual_seed(T)
        spec = torch.randn(1, F, M, T, generator=g)
        cond = torch.randn(1, T, 256, generator=g).transpose(1, 2)
        with torch.no_grad():
            y = net(spec.to(dev), t.to(dev), cond.to(dev)).cpu()
        assert rel_l2(y, oracle(spec, t, cond)) < EVAL_TOL
        if T in outs:
            assert torch.equal(outs[T], y)
        outs[T] = y
    assert len(net.engine().sessions) == 3


def test_invalid_arguments_raise():
    C, L, cycle, M, F = 64, 2, 2, 16, 1
    sd = random_state_dict(M, F, L, C, 64, seed=1)
    net = _net(sd, C, L, cycle, M, F, H=64)
    eng = net.engine()
    with pytest.raises(nat.NativeError):
        eng.session(0, 10)
    with pytest.raises(nat.NativeError):
        eng.session(1, 0)
    sess = eng.session(1, 8)
    x = torch.zeros(1, 8, M, device=dev)
    with pytest.raises(nat.NativeError, match="set_cond"):
        sess.forward(x, 0)
    sess.set_cond(torch.zeros(1, 8, 64, device=dev))
    with pytest.raises(nat.NativeError, match="time index"):
        sess.forward(x, 0)
    # unsupported shapes are rejected at plan creation, not silently emulated
    ds.hparams.update(hidden_size=64)
    bad = ds.WaveNet(16, 1, num_layers=2, num_channels=96, dilation_cycle_length=2).to(dev)
    with pytest.raises(nat.NativeError, match="multiple of 64"):
        bad.engine()
    with pytest.raises(RuntimeError, match="no CPU inference path"):
        ds.WaveNet(16, 1, num_layers=2, num_channels=64, dilation_cycle_length=2).engine()


def test_sample_segments_matches_sequential():
    """Segment-parallel sampling on several streams == segment-by-segment sampling (same noise order, same kernels)."""
    M, F, L, C, H, cycle = 128, 1, 4, 256, 256, 4
    sd = random_state_dict(M, F, L, C, H, seed=13)
    ds.hparams.clear()
    ds.hparams.update(hidden_size=H, schedule_type="linear", use_shallow_diffusion=False, diff_accelerator="ddim",
                      diff_speedup=100, K_step=1000, K_step_infer=1000, timesteps=1000)
    m = ds.GaussianDiffusion(M, F, 1000, 1000, "wavenet", dict(num_layers=L, num_channels=C, dilation_cycle_length=cycle),
                             spec_min=[-12.0], spec_max=[0.0])
    m.denoise_fn.load_state_dict(sd)
    m = m.eval().to(dev)
    g = torch.Generator().manual_seed(2)
    conds = [torch.randn(1, T, H, generator=g).to(dev) for T in (286, 504, 286, 130, 777, 504, 286)]
    torch.manual_seed(11)
    with torch.no_grad():
        seq = [m(c, infer=True).cpu() for c in conds]
    for n in (1, 3):
        torch.manual_seed(11)
        par = ds.sample_segments(m, conds, n_streams=n)
        torch.cuda.synchronize()
        assert all(torch.equal(a, b.cpu()) for a, b in zip(seq, par))
    assert ds.sample_segments(m, []) == []


@pytest.mark.parametrize("C,B,T,path", [(512, 1, 500, "small"), (512, 1, 2800, "small"), (512, 1, 3000, "medium"), (512, 3, 3500, "medium"),
                                        (512, 3, 4000, "fused"), (512, 64, 2000, "fused"), (256, 1, 8000, "small"), (256, 2, 8000, "fused")])
def test_session_path_by_size(C, B, T, path, monkeypatch):
    """Sessions pick their layer kernels from the tile-pair count and the channel width (small / medium
    problems); DSB200_SMALL / DSB200_MEDIUM only override."""
    for v in ("DSB200_SMALL", "DSB200_MEDIUM", "DSB200_LAYER_MODE"):
        monkeypatch.delenv(v, raising=False)
    net = _net(random_state_dict(128, 1, 2, C, 256, seed=1), C, 2, 2, 128, 1)
    assert net.engine().session(B, T).path == path


@pytest.mark.parametrize("family", ["ddpm", "reflow"])
def test_shallow_start_kernel_equals_the_eager_sequence(family):
    """ds_shallow_start (norm_spec -> q_sample / linear mix -> time-major in one kernel) is bit-identical to the eager
    torch sequence it replaces, and the samplers take it for plain mel sources."""
    from diffsinger_b200.core.runtime import shallow_start, to_time_major
    B, T, M = 3, 517, 128
    g = torch.Generator().manual_seed(2)
    src = (torch.rand(B, T, M, generator=g) * 12 - 12).to(dev)
    noise = torch.randn(B, 1, M, T, generator=g).to(dev)
    smin = torch.linspace(-12.0, -10.0, M)[None, None, :].to(dev)
    smax = torch.linspace(0.0, 1.0, M)[None, None, :].to(dev)
    ca, cb = (0.6180339887, 0.7861513778) if family == "ddpm" else (0.4, 1 - 0.4)
    norm = ((src - smin) / (smax - smin) * 2 - 1).transpose(-2, -1)[:, None]
    ref = to_time_major(ca * norm + cb * noise)
    got = shallow_start(src, noise, smin, smax, ca, cb)
    assert torch.equal(got, ref)
    # through the sampler: same final mel with the fused start state and with the eager one
    H, L, C = 256, 3, 256
    ds.hparams.clear()
    ds.hparams.update(hidden_size=H, use_shallow_diffusion=True, K_step=400, K_step_infer=400, diff_speedup=100, diff_accelerator="ddim",
                      timesteps=1000, schedule_type="linear", sampling_algorithm="euler", sampling_steps=5, T_start=0.4, T_start_infer=0.4,
                      time_scale_factor=1000)
    bb = dict(num_layers=L, num_channels=C, dilation_cycle_length=3)
    if family == "ddpm":
        m = ds.GaussianDiffusion(M, 1, 1000, 400, "wavenet", bb, spec_min=[-12.0], spec_max=[0.0])
        net = m.denoise_fn
    else:
        m = ds.RectifiedFlow(M, 1, t_start=0.4, backbone_type="wavenet", backbone_args=bb, spec_min=[-12.0], spec_max=[0.0])
        net = m.velocity_fn
    net.load_state_dict(random_state_dict(M, 1, L, C, H, seed=12))
    m = m.eval().to(dev)
    cond = torch.randn(B, T, H, generator=g).to(dev)
    torch.manual_seed(9)
    fused = m(cond, src_spec=src, infer=True)
    torch.manual_seed(9)
    with torch.no_grad():  # the eager start state, as the reference composes it
        spec = m.norm_spec(src).transpose(-2, -1)[:, None]
        kw = dict(x_start=spec) if family == "ddpm" else dict(x_end=spec)
        eager = m.denorm_spec(m.inference(cond.transpose(1, 2), b=B, device=cond.device, **kw))
    assert torch.equal(fused, eager)
