"""One sampling run at configs[1] shapes with few steps -- target for the ncu launch list."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import diffsinger_b200 as ds  # noqa: E402
from diffsinger_b200.synthetic import random_state_dict  # noqa: E402

B, T, C, L = int(os.environ.get("PB", 64)), int(os.environ.get("PT", 2000)), 512, 20
ds.hparams.update(hidden_size=256, schedule_type="linear", use_shallow_diffusion=False, diff_accelerator="dpm-solver",
                  diff_speedup=250, K_step=1000, K_step_infer=1000, timesteps=1000)
m = ds.GaussianDiffusion(128, 1, 1000, 1000, "wavenet", dict(num_layers=L, num_channels=C, dilation_cycle_length=4),
                         spec_min=[-12.0], spec_max=[0.0])
m.denoise_fn.load_state_dict(random_state_dict(128, 1, L, C, 256, seed=1234))
m = m.eval().cuda()
cond = torch.randn(B, T, 256, device="cuda")
for _ in range(2):
    with torch.no_grad():
        y = m(cond, infer=True)
torch.cuda.synchronize()
print("ok", float(y.mean()))
