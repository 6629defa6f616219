"""GPU unit tests of the tcgen05 GEMM kernel through the C ABI debug hook (ds_debug_gemm), one per
epilogue, against plain torch fp32 math on the same bf16-rounded operands.
Tolerance: operands are identical bf16 values and accumulation is fp32 in both, so the only
differences are summation order, the bf16 rounding of the stored output (2^-9 relative) and, for the
gate, tanh.approx (2^-11): rel-L2 <= 4e-3 for bf16 outputs, 1e-5 for the fp32 output."""
import ctypes as C

import pytest
import torch

from conftest import rel_l2
from diffsinger_b200 import _native as nat

pytestmark = pytest.mark.gpu
dev = "cuda"


def _desc(**kw):
    d = nat.GemmDesc()
    for k, v in kw.items():
        setattr(d, k, v.data_ptr() if isinstance(v, torch.Tensor) else v)
    return d


def _run(d):
    nat.check(nat.lib().ds_debug_gemm(C.byref(d), torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()


def _bf(*shape, scale=1.0):
    return (torch.randn(*shape, device=dev) * scale).to(torch.bfloat16)


# the last case has >= 4 x 148 tiles: it runs on the CTA-pair kernel (gemm_pair_kernel.cuh) with an odd number of K blocks and
# a ragged last 256-frame tile
@pytest.mark.parametrize("BN,B,T,K,N", [(64, 2, 200, 128, 64), (128, 1, 37, 64, 256), (256, 3, 7000, 512, 512),
                                        (256, 6, 3300, 192, 1024),
                                        # few tiles x long K: the split-K cluster kernel (gemm_splitk_kernel.cuh), ragged last tile
                                        (64, 1, 300, 4096, 512), (128, 2, 200, 4160, 256)])
def test_linear(BN, B, T, K, N):
    torch.manual_seed(0)
    a, w = _bf(B, T, K), _bf(N, K, scale=K ** -0.5)
    bias, vadd = torch.randn(N, device=dev), torch.randn(B, N, device=dev)
    out = torch.full((B, T, N), float("nan"), device=dev, dtype=torch.bfloat16)
    d = _desc(epi=0, BN=BN, B=B, T=T, a0=a, a0_channels=K, a0_c0=0, kb_a0=K // 64, n_taps=1, dil=0, kb_a1=0,
              w=w, w_rows=N, w_k=K, b_row0=0, n_tiles=N // BN, bias=bias, vec_add=vadd, vec_bstride=N,
              out=out, ld_out=N, out_c0=0, relu=1, scale=0.5)
    _run(d)
    ref = torch.relu(a.float() @ w.float().T + bias) * 0.5 + vadd[:, None, :]
    assert torch.isfinite(out.float()).all()
    assert rel_l2(out.float(), ref) < 4e-3


@pytest.mark.parametrize("BN,dil,T", [(128, 1, 300), (256, 8, 129), (128, 16, 40), (256, 4, 38000)])  # the last one: CTA-pair kernel
def test_gate_dilated(BN, dil, T):
    """3 shifted taps (zero padded per utterance) + K-concatenated conditioner, sigmoid*tanh epilogue."""
    torch.manual_seed(1)
    B, Cc, H = 2, 128, 64
    half = BN // 2
    n_tiles = 2  # 2*half gated channels per test
    Cz = n_tiles * half
    x, cond = _bf(B, T, Cc), _bf(B, T, H)
    K = 3 * Cc + H
    w = _bf(n_tiles * BN, K, scale=K ** -0.5)
    bias = torch.randn(n_tiles * BN, device=dev) * 0.1
    z = torch.full((B, T, Cz + 64), float("nan"), device=dev, dtype=torch.bfloat16)
    d = _desc(epi=1, BN=BN, B=B, T=T, a0=x, a0_channels=Cc, a0_c0=0, kb_a0=Cc // 64, n_taps=3, dil=dil,
              a1=cond, a1_channels=H, kb_a1=H // 64, w=w, w_rows=n_tiles * BN, w_k=K, b_row0=0, n_tiles=n_tiles,
              bias=bias, out=z, ld_out=Cz + 64, out_c0=64, relu=0, scale=1.0)
    _run(d)
    xf = x.float()
    taps = []
    for k in range(3):
        s = (k - 1) * dil
        sh = torch.zeros_like(xf)
        if s < 0:
            sh[:, -s:] = xf[:, :T + s] if T + s > 0 else 0
        elif s > 0:
            sh[:, :T - s] = xf[:, s:] if T - s > 0 else 0
        else:
            sh = xf
        taps.append(sh)
    feat = torch.cat(taps + [cond.float()], dim=-1)  # [B,T,K]
    y = feat @ w.float().T + bias
    y = y.reshape(B, T, n_tiles, BN)
    ref = (torch.sigmoid(y[..., :half]) * torch.tanh(y[..., half:])).reshape(B, T, Cz)
    got = z[..., 64:].float()
    assert torch.isfinite(got).all()
    assert rel_l2(got, ref) < 4e-3
    assert torch.isnan(z[..., :64].float()).all()  # columns outside out_c0.. untouched


def test_residual_and_f32_out():
    torch.manual_seed(2)
    B, T, Cc = 2, 333, 128
    zin, w = _bf(B, T, 3 * Cc), _bf(2 * Cc, Cc, scale=Cc ** -0.5)
    xt = _bf(B, T, Cc)
    bias = torch.randn(2 * Cc, device=dev)
    dsub, dadd = torch.randn(B, Cc, device=dev), torch.randn(B, Cc, device=dev)
    xt0 = xt.clone()
    # residual epilogue, in place on xt, A = channels [Cc, 2Cc) of zin, weight rows [Cc, 2Cc)
    d = _desc(epi=2, BN=64, B=B, T=T, a0=zin, a0_channels=3 * Cc, a0_c0=Cc, kb_a0=Cc // 64, n_taps=1, dil=0, kb_a1=0,
              w=w, w_rows=2 * Cc, w_k=Cc, b_row0=Cc, n_tiles=Cc // 64, bias=bias[Cc:].contiguous(), vec_sub=dsub,
              vec_add=dadd, vec_bstride=Cc, res_in=xt, ld_res=Cc, out=xt, ld_out=Cc, out_c0=0, relu=0, scale=0.5 ** 0.5)
    _run(d)
    acc = zin[..., Cc:2 * Cc].float() @ w[Cc:].float().T + bias[Cc:]
    ref = (xt0.float() + zin[..., Cc:2 * Cc].float() @ w[Cc:].float().T) * (0.5 ** 0.5) + dadd[:, None]  # bias/vec_sub are folded into vec_add by the caller
    assert rel_l2(xt.float(), ref) < 4e-3
    # fp32 output with n_real < BN
    n_real = 48
    wout = torch.zeros(64, Cc, device=dev, dtype=torch.bfloat16)
    wout[:n_real] = _bf(n_real, Cc, scale=Cc ** -0.5)
    bout = torch.zeros(64, device=dev)
    bout[:n_real] = torch.randn(n_real, device=dev)
    out = torch.full((B, T, n_real), float("nan"), device=dev)
    d = _desc(epi=3, BN=64, B=B, T=T, a0=xt0, a0_channels=Cc, a0_c0=0, kb_a0=Cc // 64, n_taps=1, dil=0, kb_a1=0,
              w=wout, w_rows=64, w_k=Cc, b_row0=0, n_tiles=1, bias=bout, out=out, ld_out=n_real, out_c0=0,
              n_real=n_real, relu=0, scale=1.0)
    _run(d)
    ref = xt0.float() @ wout[:n_real].float().T + bout[:n_real]
    assert torch.isfinite(out).all()
    assert rel_l2(out, ref) < 1e-5
