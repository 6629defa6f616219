"""Debug: per-tensor gradient errors of the native training path vs torch autograd (fp32, TF32 off, and fp64)."""
import copy, sys
import torch
sys.path.insert(0, "tests")
import diffsinger_b200 as ds
from diffsinger_b200.synthetic import random_state_dict

torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
C, L, M, F, cycle, B, T = 256, 4, 128, 1, 4, 2, 300
ds.hparams.update(hidden_size=256)
m = ds.WaveNet(M, F, num_layers=L, num_channels=C, dilation_cycle_length=cycle)
m.load_state_dict(random_state_dict(M, F, L, C, 256, seed=3), strict=True)
m = m.cuda()
ref = copy.deepcopy(m); ref.train_path = 'torch'
ref64 = copy.deepcopy(m).double(); ref64.train_path = 'torch'
torch.manual_seed(11)
spec = torch.randn(B, F, M, T, device='cuda'); cond = torch.randn(B, 256, T, device='cuda')
t = torch.randint(0, 1000, (B,), device='cuda'); w = torch.randn(B, F, M, T, device='cuda')
res = []
for mod, dt in ((m, torch.float32), (ref, torch.float32), (ref64, torch.float64)):
    s, c = spec.to(dt).detach().clone().requires_grad_(True), cond.to(dt).detach().clone().requires_grad_(True)
    y = mod(s, t.to(dt) if dt == torch.float64 else t, c)
    (y * w.to(dt)).sum().backward()
    g = {"out": y.detach(), "spec": s.grad, "cond": c.grad}
    g.update({n: p.grad for n, p in mod.named_parameters()})
    res.append(g)
rl = lambda a, b: float((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-30))
for n in res[0]:
    print(f"{n:60s} native-vs-64 {rl(res[0][n], res[2][n]):.3e}   fp32-vs-64 {rl(res[1][n], res[2][n]):.3e}  |g| {float(res[2][n].norm()):.3e}")
