"""Training path: the native forward / backward of the WaveNet residual stack against fp32 torch
autograd of the same module (``train_path='torch'``: plain differentiable torch ops, wavenet.py:33-48,75-107)."""
import copy

import pytest
import torch

from conftest import rel_l2

pytestmark = pytest.mark.gpu

# bf16 activations / gradients between kernels, fp32 accumulation inside them.  The residual stack itself is smooth
# (sigmoid * tanh); through the whole module the two ReLUs (wavenet.py:88,98) sit on values that carry bf16 rounding, so a
# fraction ~3e-3 of their masks flips against the fp32 run and every gradient behind them inherits sqrt(3e-3) ~ 5e-2 of
# relative noise -- the same property bf16 autocast has in the reference's trainer.
OUT_TOL, STACK_GRAD_TOL, MODULE_GRAD_TOL = 1e-2, 1e-2, 9e-2


def _model(C, L, M, F, cycle, hidden=256):
    import diffsinger_b200 as ds
    from diffsinger_b200.synthetic import random_state_dict
    ds.hparams.update(hidden_size=hidden)
    m = ds.WaveNet(M, F, num_layers=L, num_channels=C, dilation_cycle_length=cycle)
    m.load_state_dict(random_state_dict(M, F, L, C, hidden, seed=3), strict=True)
    return m.cuda()


COS_FLOOR = 0.995  # per tensor: a wrong / missing term (e.g. a bias gradient off by a factor) fails this even inside a loose rel-L2


def _check(grads, tol):
    errs = {}
    for name, g_ref in grads[1].items():
        g = grads[0][name]
        assert g is not None and g.shape == g_ref.shape, name
        errs[name] = rel_l2(g, g_ref)
        if float(g_ref.double().norm()) > 0:
            cos = float((g.double() * g_ref.double()).sum() / (g.double().norm() * g_ref.double().norm()).clamp_min(1e-30))
            assert cos > COS_FLOOR, (name, cos)
    worst = max(errs, key=errs.get)
    print(f"worst gradient rel-L2 {errs[worst]:.3e} ({worst})")
    bad = {n: f"{e:.3e}" for n, e in errs.items() if not e < tol}
    assert not bad, bad


SHAPES = [
    (512, 20, 128, 1, 4, 4, 700),  # the acoustic backbone at its real depth / width (configs[1]), ragged last tile
    (256, 4, 128, 1, 4, 2, 300),
    (192, 3, 24, 2, 2, 3, 131),   # ragged tile, multi-variance shape, dilations 1, 2, 1
    (512, 2, 128, 1, 4, 1, 257),
    (256, 5, 64, 1, 5, 2, 140),   # pitch predictor shape: dilation 16 next to short utterances
    (256, 4, 128, 1, 4, 1, 7),    # shorter than the largest dilation and than one 64-frame reduction block
    (192, 2, 24, 2, 2, 5, 1),     # single-frame utterances: every off-centre tap is padding
]


@pytest.mark.parametrize("C,L,M,F,cycle,B,T", SHAPES)
def test_stack_matches_fp32_autograd(C, L, M, F, cycle, B, T):
    """WaveNetStack (native forward, data gradients, ds_wgrad weight gradients) against the same L residual blocks as fp32
    torch ops (wavenet.py:33-48, 91-96), TF32 off."""
    import math
    from diffsinger_b200.training import WaveNetStack
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.manual_seed(11)
    m = _model(C, L, M, F, cycle)
    layers = m.residual_layers
    x0 = torch.randn(B, T, C, device='cuda').relu()
    cond = torch.randn(B, T, 256, device='cuda')
    emb = torch.randn(B, C, device='cuda')
    w = torch.randn(B, T, C, device='cuda')
    outs, grads = [], []
    for native in (True, False):
        m.zero_grad()
        xs, cs, es = (v.clone().requires_grad_(True) for v in (x0, cond, emb))
        if native:
            st = lambda f: torch.stack([f(l) for l in layers])
            skip = WaveNetStack.apply(
                xs, cs, torch.stack([l.diffusion_projection(es) for l in layers]),
                st(lambda l: l.dilated_conv.weight), st(lambda l: l.dilated_conv.bias),
                st(lambda l: l.conditioner_projection.weight[:, :, 0]), st(lambda l: l.conditioner_projection.bias),
                st(lambda l: l.output_projection.weight[:, :, 0]), st(lambda l: l.output_projection.bias),
                [l.dilation for l in layers])
        else:
            x, c, skip = xs.transpose(1, 2), cs.transpose(1, 2), 0
            for l in layers:
                x, sk = l(x, c, es)
                skip = skip + sk
            skip = skip.transpose(1, 2)
        (skip * w).sum().backward()
        outs.append(skip.detach())
        g = {"x0": xs.grad, "cond": cs.grad, "emb": es.grad}
        g.update({n: p.grad.clone() for n, p in layers.named_parameters() if p.grad is not None})
        grads.append(g)
    # the last layer's residual half is unused (wavenet.py:92-94): its parameters get a zero gradient on the native path
    # and none from autograd
    for n in list(grads[0]):
        if n not in grads[1]:
            assert float(grads[0][n].abs().max()) == 0.0, n
    err = rel_l2(outs[0], outs[1])
    print(f"stack forward rel-L2 {err:.3e}")
    assert err < OUT_TOL
    _check(grads, STACK_GRAD_TOL)


@pytest.mark.parametrize("C,L,M,F,cycle,B,T", SHAPES[1:4])
def test_module_forward_backward_match_torch_autograd(C, L, M, F, cycle, B, T):
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.manual_seed(11)
    native = _model(C, L, M, F, cycle)
    ref = copy.deepcopy(native)
    ref.train_path = 'torch'
    spec = torch.randn(B, F, M, T, device='cuda')
    cond = torch.randn(B, 256, T, device='cuda')
    t = torch.randint(0, 1000, (B,), device='cuda')
    w = torch.randn(B, F, M, T, device='cuda')
    outs, grads = [], []
    for m in (native, ref):
        s, c = spec.clone().requires_grad_(True), cond.clone().requires_grad_(True)
        y = m(s, t, c)
        (y * w).sum().backward()
        outs.append(y.detach())
        g = {"spec": s.grad, "cond": c.grad}
        g.update({n: p.grad for n, p in m.named_parameters()})
        grads.append(g)
    assert outs[0].shape == outs[1].shape == (B, F, M, T)
    err = rel_l2(outs[0], outs[1])
    print(f"train forward rel-L2 {err:.3e}")
    assert err < OUT_TOL
    _check(grads, MODULE_GRAD_TOL)


def test_native_stack_under_autocast():
    """The reference's trainer may run the step under bf16 autocast (Lightning precision plugin): the Function takes fp32
    copies of autocast's bf16 tensors, and gradients reach the fp32 parameters."""
    torch.manual_seed(2)
    m = _model(256, 3, 128, 1, 4)
    spec = torch.randn(2, 1, 128, 160, device='cuda')
    cond = torch.randn(2, 256, 160, device='cuda')
    t = torch.randint(0, 1000, (2,), device='cuda')
    y0 = m(spec, t, cond)
    with torch.autocast('cuda', dtype=torch.bfloat16):
        y1 = m(spec, t, cond)
    y1.float().square().mean().backward()
    assert rel_l2(y1.detach().float(), y0.detach().float()) < 3e-2
    for n, p in m.named_parameters():
        assert p.grad is not None and p.grad.dtype == torch.float32 and torch.isfinite(p.grad).all(), n


def test_p_losses_step_trains():
    """One optimiser step of GaussianDiffusion.p_losses (ddpm.py:212-219) through the native stack lowers the loss."""
    import diffsinger_b200 as ds
    from diffsinger_b200.synthetic import random_state_dict
    ds.hparams.update(hidden_size=256, schedule_type='linear', use_shallow_diffusion=False, diff_accelerator='ddim',
                      diff_speedup=50, K_step=1000, timesteps=1000)
    m = ds.GaussianDiffusion(128, 1, 1000, 1000, 'wavenet', dict(num_layers=4, num_channels=256, dilation_cycle_length=4),
                             spec_min=[-12.], spec_max=[0.]).cuda()
    m.denoise_fn.load_state_dict(random_state_dict(128, 1, 4, 256, 256, seed=5), strict=True)
    opt = torch.optim.SGD(m.denoise_fn.parameters(), lr=1e-3)
    torch.manual_seed(0)
    cond = torch.randn(2, 200, 256, device='cuda')
    gt = torch.rand(2, 200, 128, device='cuda') * -12.0
    losses = []
    for _ in range(3):
        torch.manual_seed(1)  # same t / noise every iteration
        x_recon, noise = m(cond, gt_spec=gt, infer=False)
        loss = torch.nn.functional.mse_loss(x_recon, noise)
        opt.zero_grad()
        loss.backward()
        opt.step()
        losses.append(loss.item())
    print("p_losses:", losses)
    assert losses[-1] < losses[0]


def test_reflow_and_variance_p_losses_match_torch_path():
    """RectifiedFlow.p_losses (reflow.py:36-41; float per-sample t) and MultiVarianceDiffusion training (ddpm.py:212-219 through the
    repeat-bins wrapper, F = 2) through the native stack: same loss as the torch-op path on the same draws, finite gradients."""
    import diffsinger_b200 as ds
    from diffsinger_b200.synthetic import random_state_dict
    ds.hparams.clear()
    ds.hparams.update(hidden_size=256, schedule_type='linear', use_shallow_diffusion=False, diff_accelerator='ddim', diff_speedup=50,
                      K_step=1000, timesteps=1000, sampling_algorithm='euler', sampling_steps=20)
    cond = torch.randn(2, 180, 256, device='cuda', generator=torch.Generator('cuda').manual_seed(0))
    gt = torch.rand(2, 180, 128, device='cuda', generator=torch.Generator('cuda').manual_seed(1)) * -12.0
    rf = ds.RectifiedFlow(128, 1, 0.0, 1000, 'wavenet', dict(num_layers=3, num_channels=256, dilation_cycle_length=4),
                          spec_min=[-12.], spec_max=[0.]).cuda()
    rf.velocity_fn.load_state_dict(random_state_dict(128, 1, 3, 256, 256, seed=6), strict=True)
    losses = {}
    for path in ('native', 'torch'):
        rf.velocity_fn.train_path = path
        rf.zero_grad()
        torch.manual_seed(4)
        v_pred, v_gt, t = rf(cond, gt_spec=gt, infer=False)
        assert t.dtype == torch.float32 and t.shape == (2,)
        loss = torch.nn.functional.mse_loss(v_pred, v_gt)
        loss.backward()
        losses[path] = loss.item()
        assert all(torch.isfinite(p.grad).all() for p in rf.velocity_fn.parameters())
    assert abs(losses['native'] - losses['torch']) < 2e-3 * abs(losses['torch']), losses

    mv = ds.MultiVarianceDiffusion(ranges=[(-96.0, -12.0), (-96.0, -20.0)], clamps=[(-96.0, 0.0), (-96.0, 0.0)], repeat_bins=24,
                                   timesteps=1000, k_step=1000, backbone_type='wavenet',
                                   backbone_args=dict(num_layers=2, num_channels=192, dilation_cycle_length=2)).cuda()
    mv.denoise_fn.load_state_dict(random_state_dict(24, 2, 2, 192, 256, seed=7), strict=True)
    curves = [torch.rand(2, 180, device='cuda') * -60.0 - 20.0 for _ in range(2)]
    losses = {}
    for path in ('native', 'torch'):
        mv.denoise_fn.train_path = path
        mv.zero_grad()
        torch.manual_seed(5)
        x_recon, noise = mv(cond, curves, infer=False)
        loss = torch.nn.functional.mse_loss(x_recon, noise)
        loss.backward()
        losses[path] = loss.item()
        assert all(torch.isfinite(p.grad).all() for p in mv.denoise_fn.parameters())
    assert abs(losses['native'] - losses['torch']) < 2e-3 * abs(losses['torch']), losses
