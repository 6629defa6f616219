"""GPU unit tests of the primitive ops of the C ABI (ds_gemm epilogues GELU / SwiGLU / residual-with-scale, ds_layernorm,
ds_dwconv, ds_cast_bf16) against plain torch fp32 math on the same bf16-rounded operands.  Tolerances: bf16 output
rounding (2^-9 relative) dominates -> rel-L2 <= 4e-3."""
import ctypes as C

import pytest
import torch
import torch.nn.functional as F

from conftest import rel_l2
from diffsinger_b200 import _native as nat

pytestmark = pytest.mark.gpu
dev = "cuda"


def _st():
    return torch.cuda.current_stream().cuda_stream


def _bf(*shape, scale=1.0):
    return (torch.randn(*shape, device=dev) * scale).to(torch.bfloat16)


def _desc(**kw):
    d = nat.GemmDesc()
    for k, v in kw.items():
        setattr(d, k, v.data_ptr() if isinstance(v, torch.Tensor) else v)
    return d


# <= 512 channels: bulk-copy kernel (3-slot ring per warp: single rows, odd row counts, > 6 wraps of the ring, a partial second
# chunk); wider rows: register kernel
@pytest.mark.parametrize("rows,Cc", [(1000, 512), (77, 192), (5, 2048), (300, 8), (1, 512), (4737, 256), (30001, 512), (1001, 264),
                                     (2, 1024)])
def test_layernorm(rows, Cc):
    torch.manual_seed(0)
    x = _bf(rows, Cc) * 3 + 1
    g, b = torch.randn(Cc, device=dev), torch.randn(Cc, device=dev)
    guard = 4096  # sentinel elements on both sides of the output
    ybig = torch.full((rows * Cc + 2 * guard,), 1234.0, device=dev, dtype=torch.bfloat16)
    y = ybig[guard:guard + rows * Cc].view(rows, Cc)
    nat.check(nat.lib().ds_layernorm(y.data_ptr(), x.data_ptr(), g.data_ptr(), b.data_ptr(), rows, Cc, 1e-5, None, 0, None, _st()))
    ref = F.layer_norm(x.float(), (Cc,), g, b, 1e-5)
    assert rel_l2(y.float(), ref) < 4e-3
    assert (ybig[:guard] == 1234.0).all() and (ybig[-guard:] == 1234.0).all()
    y = y.clone()  # the variants below write a plain buffer
    # with a per-utterance additive vector (rows_per_vec frames share one)
    rpv = max(1, rows // 3)
    nvec = (rows + rpv - 1) // rpv
    vec = torch.randn(nvec, Cc, device=dev)
    nat.check(nat.lib().ds_layernorm(y.data_ptr(), x.data_ptr(), g.data_ptr(), b.data_ptr(), rows, Cc, 1e-5, vec.data_ptr(), rpv, None, _st()))
    ref = F.layer_norm(x.float() + vec.repeat_interleave(rpv, 0)[:rows], (Cc,), g, b, 1e-5)
    assert rel_l2(y.float(), ref) < 4e-3
    # ... and a second bf16 operand
    add = _bf(rows, Cc)
    nat.check(nat.lib().ds_layernorm(y.data_ptr(), x.data_ptr(), g.data_ptr(), b.data_ptr(), rows, Cc, 1e-5, vec.data_ptr(), rpv,
                                     add.data_ptr(), _st()))
    ref = F.layer_norm(x.float() + add.float() + vec.repeat_interleave(rpv, 0)[:rows], (Cc,), g, b, 1e-5)
    assert rel_l2(y.float(), ref) < 4e-3


# k = 31 / 7 with channels % 32 == 0 run on the mma.sync Toeplitz kernel (128-frame tiles, chains of up to 16 tiles per CTA), the rest
# on the fp32-FMA kernel: single frames, tile boundaries, a 17-tile utterance (two chained CTAs), asymmetric padding, 32 and 48 channels
@pytest.mark.parametrize("k,pad,act,B,T,CH", [(31, 15, 2, 2, 700, 1024), (31, 15, 3, 1, 40, 256), (7, 3, 0, 3, 129, 512),
                                            (3, 1, 1, 1, 5, 64), (31, 15, 0, 2, 1, 32), (31, 15, 1, 1, 127, 64), (31, 15, 2, 3, 128, 32),
                                            (7, 3, 3, 2, 257, 96), (31, 30, 0, 1, 300, 64), (31, 0, 2, 2, 300, 64), (7, 6, 1, 1, 2100, 32),
                                            (31, 15, 2, 70, 33, 32), (31, 15, 2, 1, 200, 48), (7, 3, 0, 1, 64, 16)])
def test_dwconv(k, pad, act, B, T, CH):
    torch.manual_seed(1)
    x = _bf(B, T, CH)
    w = torch.randn(CH, k, device=dev) * k ** -0.5
    bias = torch.randn(CH, device=dev)
    slope = torch.rand(CH, device=dev) * 0.5
    guard = 4096  # sentinel elements on both sides of the output: the kernel writes nothing outside [B, T, CH]
    ybig = torch.full((B * T * CH + 2 * guard,), 1234.0, device=dev, dtype=torch.bfloat16)
    y = ybig[guard:guard + B * T * CH].view(B, T, CH)
    nat.check(nat.lib().ds_dwconv(y.data_ptr(), x.data_ptr(), w.data_ptr(), bias.data_ptr(), slope.data_ptr(), B, T, CH, k, pad, act, _st()))
    assert (ybig[:guard] == 1234.0).all() and (ybig[-guard:] == 1234.0).all()
    ref = F.conv1d(F.pad(x.float().transpose(1, 2), (pad, k - 1 - pad)), w[:, None, :], bias, groups=CH)  # pad = zeros on the left
    ref = {0: lambda v: v, 1: F.relu, 2: lambda v: torch.where(v > 0, v, slope[None, :, None] * v), 3: F.silu}[act](ref).transpose(1, 2)
    assert y.shape == ref.shape
    assert rel_l2(y.float(), ref) < 4e-3


def test_gemm_gelu_swiglu_and_residual_scale():
    torch.manual_seed(2)
    B, T, K, N = 2, 300, 128, 256
    a, w = _bf(B, T, K), _bf(N, K, scale=K ** -0.5)
    bias = torch.randn(N, device=dev)
    lib = nat.lib()
    # GELU(acc + bias)
    out = torch.empty(B, T, N, device=dev, dtype=torch.bfloat16)
    d = _desc(epi=0, BN=128, B=B, T=T, a0=a, a0_channels=K, a0_c0=0, kb_a0=K // 64, n_taps=1, dil=0, kb_a1=0, w=w, w_rows=N, w_k=K,
              b_row0=0, n_tiles=N // 128, bias=bias, out=out, ld_out=N, out_c0=0, relu=2, scale=1.0)
    nat.check(lib.ds_gemm(C.byref(d), _st()))
    acc = a.float() @ w.float().T + bias
    assert rel_l2(out.float(), F.gelu(acc)) < 4e-3
    # SwiGLU: rows packed per BN=128 pass as 64 "out" rows + the 64 "gate" rows of the same channels
    half = N // 2
    perm = torch.cat([torch.cat([torch.arange(p * 64, p * 64 + 64), half + torch.arange(p * 64, p * 64 + 64)]) for p in range(N // 128)]).to(dev)
    wp, bp = w[perm].contiguous(), bias[perm].contiguous()
    z = torch.empty(B, T, half, device=dev, dtype=torch.bfloat16)
    d = _desc(epi=5, BN=128, B=B, T=T, a0=a, a0_channels=K, a0_c0=0, kb_a0=K // 64, n_taps=1, dil=0, kb_a1=0, w=wp, w_rows=N, w_k=K,
              b_row0=0, n_tiles=N // 128, bias=bp, out=z, ld_out=half, out_c0=0, relu=0, scale=1.0)
    nat.check(lib.ds_gemm(C.byref(d), _st()))
    assert rel_l2(z.float(), acc[..., :half] * F.silu(acc[..., half:])) < 4e-3
    # residual with scale 1: out = acc + res + per-batch vector
    res, vec = _bf(B, T, N), torch.randn(B, N, device=dev)
    o = torch.empty_like(res)
    d = _desc(epi=2, BN=128, B=B, T=T, a0=a, a0_channels=K, a0_c0=0, kb_a0=K // 64, n_taps=1, dil=0, kb_a1=0, w=w, w_rows=N, w_k=K,
              b_row0=0, n_tiles=N // 128, vec_add=vec, vec_bstride=N, res_in=res, ld_res=N, out=o, ld_out=N, out_c0=0, relu=0, scale=1.0)
    nat.check(lib.ds_gemm(C.byref(d), _st()))
    assert rel_l2(o.float(), a.float() @ w.float().T + res.float() + vec[:, None]) < 4e-3


def test_gemm_pair_kernel_swiglu_and_residual():
    """The same epilogues at a size the CTA-pair kernel takes (BN = 256, >= 4 x 148 tiles; gemm_pair_kernel.cuh)."""
    torch.manual_seed(3)
    B, T, K, N = 5, 2000, 128, 2048
    a, w = _bf(B, T, K), _bf(N, K, scale=K ** -0.5)
    bias = torch.randn(N, device=dev)
    lib = nat.lib()
    acc = a.float() @ w.float().T + bias
    half = N // 2
    perm = torch.cat([torch.cat([torch.arange(p * 128, p * 128 + 128), half + torch.arange(p * 128, p * 128 + 128)])
                      for p in range(N // 256)]).to(dev)
    wp, bp = w[perm].contiguous(), bias[perm].contiguous()
    z = torch.empty(B, T, half, device=dev, dtype=torch.bfloat16)
    d = _desc(epi=5, BN=256, B=B, T=T, a0=a, a0_channels=K, a0_c0=0, kb_a0=K // 64, n_taps=1, dil=0, kb_a1=0, w=wp, w_rows=N, w_k=K,
              b_row0=0, n_tiles=N // 256, bias=bp, out=z, ld_out=half, out_c0=0, relu=0, scale=1.0)
    nat.check(lib.ds_gemm(C.byref(d), _st()))
    assert rel_l2(z.float(), acc[..., :half] * F.silu(acc[..., half:])) < 4e-3
    res, vec = _bf(B, T, N), torch.randn(B, N, device=dev)
    o = torch.empty_like(res)
    d = _desc(epi=2, BN=256, B=B, T=T, a0=a, a0_channels=K, a0_c0=0, kb_a0=K // 64, n_taps=1, dil=0, kb_a1=0, w=w, w_rows=N, w_k=K,
              b_row0=0, n_tiles=N // 256, vec_add=vec, vec_bstride=N, res_in=res, ld_res=N, out=o, ld_out=N, out_c0=0, relu=0, scale=0.5)
    nat.check(lib.ds_gemm(C.byref(d), _st()))
    assert rel_l2(o.float(), (a.float() @ w.float().T + res.float()) * 0.5 + vec[:, None]) < 4e-3


def test_cast_bf16_pads():
    x = torch.randn(37, 48, device=dev)
    y = torch.full((37, 64), 7.0, device=dev, dtype=torch.bfloat16)
    nat.check(nat.lib().ds_cast_bf16(y.data_ptr(), x.data_ptr(), 37, 48, 64, _st()))
    assert torch.equal(y[:, :48], x.to(torch.bfloat16)) and (y[:, 48:] == 0).all()
    with pytest.raises(nat.NativeError):
        nat.check(nat.lib().ds_dwconv(y.data_ptr(), y.data_ptr(), x.data_ptr(), None, None, 1, 8, 64, 5, 2, 0, _st()))


@pytest.mark.parametrize("BN", [64, 128, 256])
@pytest.mark.parametrize("B,T", [(3, 150), (1, 64), (2, 333)])
def test_wgrad_shifted_frame_reductions(BN, B, T):
    """ds_wgrad: out = sum_{b,t} G[b,t,:]^T X[b,t+shift,:] with zero padding outside [0, T) -- the weight gradients of a dilated
    conv tap / 1x1 conv (wavenet.py:36-44 under autograd).  fp32 accumulation of bf16 products: rel-L2 <= 1e-5 vs fp64."""
    from diffsinger_b200 import ops
    torch.manual_seed(B * 1000 + T + BN)
    Lg, Cg, Cx = 2, 320, 448              # G: two "layers" side by side in one buffer; X with its own batch-row offset
    G = _bf(B * T, Lg * Cg)
    X = _bf(2 * B * T, Cx)                # batch rows [B, 2B) are used
    jobs, refs = [], []
    ld_out, off = Cx, 0
    out = torch.full((Lg * 3 * Cg * ld_out,), float("nan"), device=dev)
    Gv, Xv = G.view(B, T, Lg * Cg).double(), X.view(2 * B, T, Cx).double()[B:]
    for l in range(Lg):
        for shift in (-4, 0, 16):
            lo, hi = max(0, -shift), min(T, T - shift)
            ref = torch.einsum("btm,btn->mn", Gv[:, lo:hi, l * Cg:(l + 1) * Cg], Xv[:, lo + shift:hi + shift]) if hi > lo else \
                torch.zeros(Cg, Cx, dtype=torch.float64, device=dev)
            refs.append((off, ref))
            for m0 in range(0, Cg, 128):
                for n0 in range(0, Cx, BN):
                    jobs.append((l * Cg + m0, 0, n0, B, shift, min(128, Cg - m0), min(BN, Cx - n0), ld_out, off + m0 * ld_out + n0, 1))
            off += Cg * ld_out
    decoy = _bf(B * T, 64)  # operand pair 0 is not the one the jobs name
    ops.wgrad(out, [(decoy, decoy), (G, X)], T, B, BN, ops.wgrad_jobs(jobs, dev))
    for o, ref in refs:
        got = out[o:o + Cg * ld_out].view(Cg, ld_out)
        assert torch.isfinite(got).all()
        assert rel_l2(got, ref) < 1e-5


@pytest.mark.parametrize("C", [16, 32])
@pytest.mark.parametrize("taps,dil", [(3, 1), (7, 3), (11, 5), (11, 1), (3, 5)])
def test_conv_small(C, taps, dil):
    """ds_conv_small: narrow-channel k-tap dilated conv on tcgen05 (no-swizzle plane layout, one halo tile per 128 samples)
    against F.conv1d on the same bf16-rounded operands; LeakyReLU, residual + activated copy, tanh-to-fp32 epilogues."""
    from diffsinger_b200 import ops
    torch.manual_seed(C * 100 + taps * 10 + dil)
    B, T = 3, 517
    x = _bf(B * T, C)
    w = torch.randn(C, C, taps, device=dev) * (1.0 / (C * taps) ** 0.5)
    bias = torch.randn(C, device=dev) * 0.1
    res = _bf(B * T, C)
    w_img = ops.conv_small_weights(w)
    ref = F.conv1d(x.float().view(B, T, C).transpose(1, 2), w.to(torch.bfloat16).float(), bias, dilation=dil,
                   padding=(taps // 2) * dil).transpose(1, 2).reshape(B * T, C)
    # conv1 of a ResBlock: LeakyReLU epilogue
    y = torch.empty_like(x)
    ops.conv_small(x, w_img, bias, y, B, T, dil=dil, act=ops.ACT_LRELU, slope=0.1)
    assert rel_l2(y.float(), F.leaky_relu(ref, 0.1)) < 4e-3
    # conv2: residual + activated copy
    y2 = torch.empty_like(x)
    ops.conv_small(x, w_img, bias, y, B, T, dil=dil, res=res, out2=y2, slope=0.1)
    assert rel_l2(y.float(), ref + res.float()) < 4e-3
    assert rel_l2(y2.float(), F.leaky_relu(ref + res.float(), 0.1)) < 4e-3
    # in place on the residual
    r2 = res.clone()
    ops.conv_small(x, w_img, bias, r2, B, T, dil=dil, res=r2)
    assert rel_l2(r2.float(), ref + res.float()) < 4e-3
    # conv_post: tanh of output channel 0 -> fp32 [B * T]
    wav = torch.empty(B * T, device=dev)
    ops.conv_small(x, w_img, bias, wav, B, T, dil=dil, act=ops.ACT_TANH)
    assert rel_l2(wav, torch.tanh(ref[:, 0])) < 1e-3


def test_gate_fwd_bwd_and_fused_gate_backward_epilogue():
    """ds_gate_fwd / ds_gate_bwd (wavenet.py:41-42 and its derivative) on column slices of wider buffers, and the same
    derivative fused behind a GEMM (ds_gemm epilogue 6), against torch autograd on the bf16-rounded pre-activations."""
    from diffsinger_b200 import ops
    torch.manual_seed(5)
    B, T, Cc, L = 2, 150, 128, 3
    rows = B * T
    Z = _bf(rows, L * 2 * Cc) * 2
    Y = torch.zeros(rows, L * Cc, dtype=torch.bfloat16, device=dev)
    DZ = torch.zeros_like(Z)
    l = 1
    z = Z[:, l * 2 * Cc:(l + 1) * 2 * Cc].float().requires_grad_(True)
    y_ref = torch.sigmoid(z[:, :Cc]) * torch.tanh(z[:, Cc:])
    dy = _bf(rows, Cc)
    y_ref.backward(dy.float())
    ops.gate_fwd(Y, l * Cc, Z, l * 2 * Cc, Cc)
    assert rel_l2(Y[:, l * Cc:(l + 1) * Cc].float(), y_ref.detach()) < 4e-3
    assert float(Y[:, :l * Cc].abs().max()) == 0 and float(Y[:, (l + 1) * Cc:].abs().max()) == 0
    ops.gate_bwd(DZ, l * 2 * Cc, dy, Z, l * 2 * Cc, Cc)
    assert rel_l2(DZ[:, l * 2 * Cc:(l + 1) * 2 * Cc].float(), z.grad) < 4e-3
    # fused: dy = A W^T with W = identity, so the epilogue sees exactly dy
    DZ2 = torch.zeros_like(Z)
    eye = torch.eye(Cc, device=dev, dtype=torch.bfloat16)
    ops.gemm(ops.EPI_GATEBWD, 128, B, T, dy, eye, DZ2, res=Z, res_c0=l * 2 * Cc, n_real=Cc, out_c0=l * 2 * Cc)
    assert rel_l2(DZ2[:, l * 2 * Cc:(l + 1) * 2 * Cc].float(), z.grad) < 6e-3
    assert float(DZ2[:, :l * 2 * Cc].abs().max()) == 0
