#!/usr/bin/env python
"""Time of one backbone evaluation through ds_stack_kernel: an Euler-20 rectified-flow program (20 evaluations + 20 updates in
the session's CUDA graph, conditioner hoisted once per call), 10 calls back to back."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import diffsinger_b200 as ds  # noqa: E402
from diffsinger_b200.synthetic import random_state_dict  # noqa: E402


def one(C, L, cycle, M, T, B=1, F=1):
    ds.hparams.clear()
    ds.hparams.update(hidden_size=256)
    net = ds.WaveNet(M, F, num_layers=L, num_channels=C, dilation_cycle_length=cycle)
    net.load_state_dict(random_state_dict(M, F, L, C, 256, seed=1))
    net = net.eval().cuda()
    spec, cond, t = torch.randn(B, F, M, T, device="cuda"), torch.randn(B, 256, T, device="cuda"), torch.full((B,), 500.0, device="cuda")
    from diffsinger_b200.core import programs
    from diffsinger_b200.core.runtime import run_program, to_time_major
    prog = programs.reflow_program("euler", 20, 0.0, 1000.0)
    x = to_time_major(spec)
    for _ in range(3):
        run_program(net, prog, x.clone(), cond)  # first call captures the session graph
    xs = [x.clone() for _ in range(10)]
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for xi in xs:
        run_program(net, prog, xi, cond)
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1000 / (10 * 20)
    path = next(reversed(net.engine().sessions.values())).path
    print(f"C={C} L={L} M={M}x{F} B={B} T={T}: {us:.1f} us / evaluation incl. the sampler update ({us / L:.2f} us / layer), path={path}, Euler-20 = {us * 20 / 1000:.2f} ms", flush=True)


if __name__ == "__main__":
    for T in (128, 512, 1024, 2048):
        one(256, 20, 5, 64, T)     # pitch predictor
    one(256, 20, 4, 128, 640)      # acoustic 20 x 256
    for T in (128, 512):
        one(192, 10, 4, 24, T, F=2)  # multi-variance predictor
