"""Cold (new shapes: sessions + graphs are created) vs warm sampler calls for ragged batches of segments."""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import sweep  # noqa: E402
import diffsinger_b200 as ds  # noqa: E402
from diffsinger_b200.synthetic import random_state_dict  # noqa: E402

sweep.hp(diff_accelerator="ddim")
m = ds.GaussianDiffusion(128, 1, 1000, 1000, "wavenet", dict(num_layers=20, num_channels=256, dilation_cycle_length=4), spec_min=[-12.0], spec_max=[0.0])
m.denoise_fn.load_state_dict(random_state_dict(128, 1, 20, 256, 256, seed=1))
m = m.eval().cuda()
g = torch.Generator().manual_seed(0)
with torch.no_grad():
    m(torch.randn(1, 300, 256, device="cuda"), infer=True)  # library, weights, first session
torch.cuda.synchronize()
for trial in range(3):
    lens = [int(v) for v in torch.randint(280, 1900, (30,), generator=g)]
    conds = [torch.randn(1, t, 256, device="cuda") for t in lens]
    ts = []
    for rep in range(3):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        with torch.no_grad():
            ds.sample_segments(m, conds)
        torch.cuda.synchronize()
        ts.append((time.perf_counter() - t0) * 1e3)
    print(f"trial {trial}: {sum(lens)} frames in 30 segments: cold {ts[0]:.1f} ms, then {ts[1]:.1f} / {ts[2]:.1f} ms", flush=True)
