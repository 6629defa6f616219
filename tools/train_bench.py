"""Training step (forward + backward of the WaveNet backbone) on one B200: the native residual stack
(diffsinger_b200/training.py) beside the same module as torch ops in fp32 and under bf16 autocast (what the reference's
trainer runs, ddpm.py:212-219 through Lightning's precision plugin).  Prints one JSON line per arm.

    python tools/train_bench.py [B T [C L]]
"""
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import diffsinger_b200 as ds  # noqa: E402
from diffsinger_b200.synthetic import random_state_dict


def main():
    a = [int(v) for v in sys.argv[1:]]
    B, T = (a + [32, 1500])[:2] if len(a) < 2 else a[:2]
    C, L = a[2:4] if len(a) >= 4 else (512, 20)
    M, H = 128, 256
    ds.hparams.update(hidden_size=H)
    m = ds.WaveNet(M, 1, num_layers=L, num_channels=C, dilation_cycle_length=4)
    m.load_state_dict(random_state_dict(M, 1, L, C, H, seed=3), strict=True)
    m = m.cuda()
    torch.manual_seed(0)
    spec = torch.randn(B, 1, M, T, device='cuda')
    cond = torch.randn(B, H, T, device='cuda')
    t = torch.randint(0, 1000, (B,), device='cuda')
    w = torch.randn(B, 1, M, T, device='cuda')
    macs = L * (8 * C * C + 2 * H * C) + C * C + 2 * M * C
    flop_step = 3 * 2 * macs * B * T  # forward + data gradient + weight gradient

    def step(path, autocast):
        m.train_path = path
        m.zero_grad(set_to_none=True)
        with torch.autocast('cuda', dtype=torch.bfloat16, enabled=autocast):
            y = m(spec, t, cond)
        (y.float() * w).sum().backward()

    arms = os.environ.get("ARMS", "native,torch_bf16_autocast,torch_fp32").split(",")
    n_warm, n = int(os.environ.get("WARM", 2)), int(os.environ.get("STEPS", 5))
    for name, path, autocast in (("native", "native", False), ("torch_bf16_autocast", "torch", True), ("torch_fp32", "torch", False)):
        if name not in arms:
            continue
        try:
            for _ in range(n_warm):
                step(path, autocast)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(n):
                step(path, autocast)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / n
            print(json.dumps(dict(arm=name, B=B, T=T, C=C, L=L, ms_per_step=round(ms, 2), frames_per_s=round(B * T / ms * 1e3, 1),
                                  tflops=round(flop_step / ms / 1e9, 1), peak_mem_gb=round(torch.cuda.max_memory_allocated() / 2**30, 2))),
                  flush=True)
        except torch.OutOfMemoryError as e:
            print(json.dumps(dict(arm=name, error="out of memory")), flush=True)
        torch.cuda.reset_peak_memory_stats()


if __name__ == "__main__":
    main()
