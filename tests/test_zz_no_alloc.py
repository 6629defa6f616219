"""ds_forward_static (include/dsb200.h): the no-allocation backbone evaluation whose session workspace is a file-scope
`static __device__` array inside libdsb200.so.  It has no counterpart in the reference;
the check is against the existing entry points (ds_session_set_cond / set_times / ds_session_forward over
caller-allocated memory) through the C ABI, on two calls in a row that share the static array.

Collected last on purpose (file name).
Tolerance: both paths launch the same deterministic kernels, so the fp32 outputs are compared at rel-L2 <= 1e-6 (the oracle
comparison of the same output uses the suite's per-evaluation bf16 bar, 2e-2)."""
import ctypes as C

import pytest
import torch

import diffsinger_b200 as ds
from diffsinger_b200 import _native as nat
from conftest import rel_l2
from oracle.wavenet_oracle import OracleWaveNet, random_state_dict

pytestmark = pytest.mark.gpu
dev = "cuda"
SAME_TOL = 1e-6
EVAL_TOL = 2e-2


def _net(C_, L_, cycle, M, F, H=256, seed=21):
    sd = random_state_dict(M, F, L_, C_, H, seed=seed)
    ds.hparams.clear()
    ds.hparams.update(hidden_size=H)
    net = ds.WaveNet(M, F, num_layers=L_, num_channels=C_, dilation_cycle_length=cycle)
    net.load_state_dict(sd, strict=True)
    return net.eval().to(dev), sd


def _static_forward(eng, cond_bth, t, x_btc, per_sample=False, out=None):
    out = torch.empty_like(x_btc) if out is None else out
    B, T = x_btc.shape[0], x_btc.shape[1]
    st = torch.cuda.current_stream().cuda_stream
    rc = nat.lib().ds_forward_static(eng.plan, cond_bth.data_ptr(), t.data_ptr(), int(per_sample), x_btc.data_ptr(), out.data_ptr(),
                                     B, T, st)
    return rc, out


@pytest.mark.parametrize("force_fused", [False, True])
def test_static_scratch_forward_matches_session_forward_twice(monkeypatch, force_fused):
    if force_fused:
        monkeypatch.setenv("DSB200_SMALL", "0")  # the one-kernel-per-layer path instead of the narrow-tile chain
    C_, L_, cycle, M, F, B, T = 256, 6, 4, 128, 1, 2, 391
    net, sd = _net(C_, L_, cycle, M, F)
    eng = net.engine()
    assert nat.lib().ds_static_scratch_bytes() == 32 << 20
    assert nat.lib().ds_session_workspace_bytes(eng.plan, B, T, 1, 0) < nat.lib().ds_static_scratch_bytes()
    g = torch.Generator().manual_seed(17)
    calls = []
    for i in range(2):  # two calls in a row share the static array (it keeps the first call's contents)
        x = torch.randn(B, T, F * M, generator=g).to(dev)
        cond = torch.randn(B, T, 256, generator=g).to(dev)
        t = torch.tensor([617.25 - 400.0 * i], device=dev)
        calls.append((x, cond, t))
    outs = []
    ys = [torch.empty_like(c[0]) for c in calls]  # the caller passes inputs AND outputs: allocated before the calls
    for i, (x, cond, t) in enumerate(calls):
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info()[0]
        alloc0 = torch.cuda.memory_allocated()
        rc, y = _static_forward(eng, cond, t, x, out=ys[i])
        assert rc == 0, nat.lib().ds_last_error()
        torch.cuda.synchronize()
        if i > 0:  # the first call may page kernels in; the second allocates nothing anywhere on the device
            assert torch.cuda.memory_allocated() == alloc0
            assert free0 - torch.cuda.mem_get_info()[0] < (1 << 20)
        outs.append(y)
    # existing entry points over caller (torch) memory
    sess = eng.session(B, T, max_times=1, per_sample=False)
    for (x, cond, t), y in zip(calls, outs):
        sess.set_cond(cond)
        sess.set_times(t)
        ref = sess.forward(x, 0)
        torch.cuda.synchronize()
        assert torch.isfinite(y).all()
        assert rel_l2(y.cpu(), ref.cpu()) <= SAME_TOL
    # and the oracle, at the suite's per-evaluation bf16 bar
    x, cond, t = calls[1]
    spec = x.cpu().transpose(1, 2).reshape(B, F, M, T)
    want = OracleWaveNet(sd, F, cycle)(spec, t.cpu(), cond.cpu().transpose(1, 2))
    got = outs[1].cpu().transpose(1, 2).reshape(B, F, M, T)
    assert rel_l2(got, want) < EVAL_TOL


def test_static_scratch_per_sample_times_and_size_limit():
    C_, L_, cycle, M, F, B, T = 192, 4, 4, 24, 2, 3, 97
    net, _ = _net(C_, L_, cycle, M, F, seed=5)
    eng = net.engine()
    g = torch.Generator().manual_seed(3)
    x = torch.randn(B, T, F * M, generator=g).to(dev)
    cond = torch.randn(B, T, 256, generator=g).to(dev)
    tb = torch.linspace(3.0, 990.0, B).to(dev)
    for _ in range(2):
        rc, y = _static_forward(eng, cond, tb, x, per_sample=True)
        assert rc == 0, nat.lib().ds_last_error()
    sess = eng.session(B, T, max_times=1, per_sample=True)
    sess.set_cond(cond)
    sess.set_times(tb.reshape(1, -1))
    ref = sess.forward(x, 0)
    torch.cuda.synchronize()
    assert rel_l2(y.cpu(), ref.cpu()) <= SAME_TOL
    # beyond the array: rejected with a clear error, nothing launched
    Bb, Tb = 64, 2000
    assert nat.lib().ds_session_workspace_bytes(eng.plan, Bb, Tb, 1, 0) > nat.lib().ds_static_scratch_bytes()
    dummy = torch.empty(8, device=dev)
    rc = nat.lib().ds_forward_static(eng.plan, dummy.data_ptr(), dummy.data_ptr(), 0, dummy.data_ptr(), dummy.data_ptr(), Bb, Tb,
                                     torch.cuda.current_stream().cuda_stream)
    assert rc == -1
    msg = nat.lib().ds_last_error().decode()
    assert "static scratch" in msg and "DS_STATIC_SCRATCH_BYTES" in msg
    torch.cuda.synchronize()
