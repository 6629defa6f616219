"""Per-job timeline of CTA 0 of the fused layer kernel (debug trace hook) at configs[1] shapes."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import diffsinger_b200 as ds  # noqa: E402
from diffsinger_b200 import _native as nat  # noqa: E402
from diffsinger_b200.synthetic import random_state_dict  # noqa: E402

B, T, C, L = int(os.environ.get("PB", 64)), 2000, 512, 2
ds.hparams.update(hidden_size=256)
net = ds.WaveNet(128, 1, num_layers=L, num_channels=C, dilation_cycle_length=4)
net.load_state_dict(random_state_dict(128, 1, L, C, 256, seed=1))
net = net.eval().cuda()
spec = torch.randn(B, 1, 128, T, device="cuda")
cond = torch.randn(B, T, 256, device="cuda").transpose(1, 2)
t = torch.tensor([500.0], device="cuda")
cap = 256
trace = torch.zeros(cap * 4, dtype=torch.int64, device="cuda")
with torch.no_grad():
    net(spec, t, cond)
    nat.lib().ds_debug_set_trace(trace.data_ptr(), cap)
    net(spec, t, cond)
torch.cuda.synchronize()
tr = trace.cpu().view(cap, 4)
rows = [r.tolist() for r in tr if r[2] != 0]
t00 = min(r[2] for r in rows)
names = {0: "epi  P", 1: "epi  C", 16: "mma  P", 17: "mma  C"}
rows.sort(key=lambda r: r[2])
for r in rows[:int(os.environ.get("NROWS", 26))]:
    print(f"{names[r[0]]}{r[1] % 16} tile{r[1] // 16:2d}  t0={(r[2] - t00) / 1e3:8.2f}us  t1={(r[3] - t00) / 1e3:8.2f}us  dt={(r[3] - r[2]) / 1e3:6.2f}")
