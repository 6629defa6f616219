#!/usr/bin/env python
"""BASELINE.json configs[0] shapes on the GPU: the 30 segments of samples/00 (frame counts from SURVEY.md section 8c), batch 1,
acoustic WaveNet 20 x 256, DDIM 20 steps, one public-API call per segment (a new (B, T) session per distinct length)."""
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import diffsinger_b200 as ds  # noqa: E402
from diffsinger_b200.synthetic import random_state_dict  # noqa: E402

SEGMENTS = [504, 522, 539, 720, 1197, 550, 550, 515, 550, 539, 539, 540, 1917, 549, 550, 515, 687, 549, 549, 549, 460, 286, 578,
            545, 515, 566, 572, 542, 516, 579]


def main():
    C = int(os.environ.get("C", 256))
    ds.hparams.clear()
    ds.hparams.update(hidden_size=256, schedule_type="linear", use_shallow_diffusion=False, diff_accelerator="ddim", diff_speedup=50,
                      K_step=1000, K_step_infer=1000, timesteps=1000)
    m = ds.GaussianDiffusion(128, 1, 1000, 1000, "wavenet", dict(num_layers=20, num_channels=C, dilation_cycle_length=4),
                             spec_min=[-12.0], spec_max=[0.0])
    m.denoise_fn.load_state_dict(random_state_dict(128, 1, 20, C, 256, seed=1234))
    m = m.eval().cuda()
    conds = [torch.randn(1, T, 256, device="cuda") for T in SEGMENTS]
    with torch.no_grad():
        m(conds[0][:, :100], infer=True)  # library load, weight packing
    torch.cuda.synchronize()
    for rnd in ("first pass (new sessions)", "second pass (cached sessions / graphs)"):
        t0 = time.perf_counter()
        per = []
        with torch.no_grad():
            for c in conds:
                t1 = time.perf_counter()
                y = m(c, infer=True)
                torch.cuda.synchronize()
                per.append((time.perf_counter() - t1) * 1e3)
        dt = time.perf_counter() - t0
        print(json.dumps(dict(case=f"config1 C={C} {rnd}", segments=len(SEGMENTS), frames=sum(SEGMENTS), total_ms=round(dt * 1e3, 1),
                              frames_per_s=round(sum(SEGMENTS) / dt, 1), ms_min=round(min(per), 2), ms_max=round(max(per), 2))), flush=True)
    streams_case(m, conds, C)


def streams_case(m, conds, C):
    for n in (1, 2, 4, 8):
        for rnd in range(2):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            ys = ds.sample_segments(m, conds, n_streams=n)
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
        print(json.dumps(dict(case=f"config1 C={C} sample_segments n_streams={n} (second pass)", segments=len(SEGMENTS),
                              frames=sum(SEGMENTS), total_ms=round(dt * 1e3, 1), frames_per_s=round(sum(SEGMENTS) / dt, 1))), flush=True)


if __name__ == "__main__":
    main()
