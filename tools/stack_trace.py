#!/usr/bin/env python
"""Per-layer timeline of ds_stack_kernel (csrc/stack_kernel.cuh): CTA 0 records %globaltimer at ten points per layer
(MMA issuer: first x~ block in / all x~ blocks in / last conv weights in / first z block in / all z blocks in / last 1x1 weights in;
epilogue: gate accumulator ready / z published / residual accumulator ready / x~ published)."""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import diffsinger_b200 as ds  # noqa: E402
from diffsinger_b200 import _native as nat  # noqa: E402
from diffsinger_b200.synthetic import random_state_dict  # noqa: E402

NAMES = ["x0 in", "x all in", "W1 last in", "z0 in", "z all in", "W2 last in", "acc1 ready", "z published", "acc2 ready", "x published"]


def main():
    T = int(os.environ.get("T", 128))
    L, cycle, M = 20, 5, 64
    ds.hparams.clear()
    ds.hparams.update(hidden_size=256)
    net = ds.WaveNet(M, 1, num_layers=L, num_channels=256, dilation_cycle_length=cycle)
    net.load_state_dict(random_state_dict(M, 1, L, 256, 256, seed=1))
    net = net.eval().cuda()
    spec, cond, t = torch.randn(1, 1, M, T, device="cuda"), torch.randn(1, 256, T, device="cuda"), torch.tensor([500.0], device="cuda")
    with torch.no_grad():
        for _ in range(3):
            net(spec, t, cond)
    buf = torch.zeros(L * 10 + 64, dtype=torch.int64, device="cuda")
    nat.lib().ds_debug_set_trace(C.c_void_p(buf.data_ptr()), (L * 10 + 64) // 4)
    with torch.no_grad():
        net(spec, t, cond)
    torch.cuda.synchronize()
    nat.lib().ds_debug_set_trace(None, 0)
    fine = buf.cpu()[L * 10:L * 10 + 36].reshape(12, 3)
    v = buf.cpu()[:L * 10].reshape(L, 10)
    t0 = int(v[0, 0])
    print(f"T={T}  (ns relative to layer 0 'x0 in'; columns: " + " | ".join(NAMES) + ")")
    for l in range(L):
        print(f"layer {l:2d}: " + " ".join(f"{int(v[l, k]) - t0:7d}" if int(v[l, k]) else "      -" for k in range(10)))
    per = [(int(v[l + 1, 0]) - int(v[l, 0])) for l in range(L - 2)]
    print("ns per layer (x0 in -> next x0 in):", per)
    # the production (uninstrumented) kernel, timed from outside
    nat.lib().ds_debug_set_trace(None, 0)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sess = next(iter(net.engine().sessions.values()))
    x_tm = spec.reshape(1, M, T).transpose(1, 2).contiguous()
    out = torch.empty_like(x_tm)
    for _ in range(3):
        sess.forward(x_tm, 0, out)
    e0.record()
    for _ in range(20):
        sess.forward(x_tm, 0, out)
    e1.record()
    torch.cuda.synchronize()
    print(f"production kernel: {e0.elapsed_time(e1) / 20 * 1e3:.1f} us per evaluation (20 back-to-back launches), path {sess.path}")


if __name__ == "__main__":
    main()
