"""Where a pitch-predictor sampler call spends its time: the whole public call vs the program run inside it."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import sweep  # noqa: E402
import diffsinger_b200 as ds  # noqa: E402
from diffsinger_b200.core import runtime  # noqa: E402
from diffsinger_b200.synthetic import random_state_dict  # noqa: E402

sweep.hp(diff_accelerator="ddim")
m = ds.PitchDiffusion(vmin=-8.0, vmax=8.0, cmin=-12.0, cmax=12.0, repeat_bins=64, backbone_type="wavenet",
                      backbone_args=dict(num_layers=20, num_channels=256, dilation_cycle_length=5))
m.denoise_fn.load_state_dict(random_state_dict(64, 1, 20, 256, 256, seed=1))
m = m.eval().cuda()
for T in (128, 512):
    cond = torch.randn(1, T, 256, device="cuda")
    with torch.no_grad():
        full = sweep.timed(lambda: m(cond, infer=True), 20)
    spans = []
    orig = runtime.run_program
    import diffsinger_b200.core.ddpm as ddpm_mod

    def timed_run(*a, **k):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        out = orig(*a, **k)
        e1.record()
        spans.append((e0, e1))
        return out

    ddpm_mod.run_program = timed_run
    with torch.no_grad():
        for _ in range(20):
            m(cond, infer=True)
    torch.cuda.synchronize()
    ddpm_mod.run_program = orig
    inner = sum(a.elapsed_time(b) for a, b in spans) / len(spans)
    sess = next(reversed(m.denoise_fn.engine().sessions.values()))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    prog = next(iter(m._programs.values()))
    bufs = runtime.sess_buffers(sess, prog.n_bufs)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(20):
        sess.run(prog, bufs)
    e1.record()
    torch.cuda.synchronize()
    print(f"T={T}: public call {full:.3f} ms; run_program inside it {inner:.3f} ms; graph replay alone {e0.elapsed_time(e1) / 20:.3f} ms", flush=True)
