"""Times the HBM-bound primitive ops behind the LYNXNet row alone at the config #2 shape (B64 x T2000):
ds_dwconv (k = 31, 1024 channels, PReLU) and ds_layernorm (512 channels), CUDA events on the launching stream, L2 flushed by
the size of the operands (262 / 131 MB per pass).  Usage: python tools/prim_bench.py [dwconv|layernorm|all] [iters]
Prints one JSON line per op with the achieved GB/s against the algorithmic bytes."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from diffsinger_b200 import ops  # noqa: E402


def timed(fn, iters):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e-3


def main():
    what = sys.argv[1] if len(sys.argv) > 1 else "all"
    iters = int(sys.argv[2]) if len(sys.argv) > 2 else 20
    B, T = 64, 2000
    dev = "cuda:0"
    g = torch.Generator(device=dev).manual_seed(0)
    if what in ("dwconv", "all"):
        CH, K = 1024, 31
        # two input buffers alternate so that consecutive launches never find their input in L2
        xs = [torch.randn(B * T, CH, device=dev, generator=g).to(torch.bfloat16) for _ in range(2)]
        y = torch.empty_like(xs[0])
        w = torch.randn(CH, K, device=dev, generator=g) * 0.2
        bias = torch.randn(CH, device=dev, generator=g)
        slope = torch.full((CH,), 0.25, device=dev)
        state = {"i": 0}

        def run():
            state["i"] ^= 1
            ops.dwconv(y, xs[state["i"]], w, bias, B, T, K, ops.DW_PRELU, slope)

        s = timed(run, iters)
        algo = B * T * CH * 4.0
        print(json.dumps({"op": "ds_dwconv", "k": K, "channels": CH, "frames": B * T, "us": s * 1e6, "algo_bytes": algo,
                          "GBps": algo / s * 1e-9, "fp32_GFMA_per_s": B * T * CH * K / s * 1e-9}))
    if what in ("layernorm", "all"):
        C = 512
        xs = [torch.randn(B * T, C, device=dev, generator=g).to(torch.bfloat16) for _ in range(2)]
        y = torch.empty_like(xs[0])
        gamma = torch.randn(C, device=dev, generator=g)
        beta = torch.randn(C, device=dev, generator=g)
        state = {"i": 0}

        def run():
            state["i"] ^= 1
            ops.layernorm(y, xs[state["i"]], gamma, beta)

        s = timed(run, iters)
        algo = B * T * C * 4.0
        print(json.dumps({"op": "ds_layernorm", "channels": C, "frames": B * T, "us": s * 1e6, "algo_bytes": algo,
                          "GBps": algo / s * 1e-9}))
        adds = [torch.randn(B * T, C, device=dev, generator=g).to(torch.bfloat16) for _ in range(2)]  # LYNXNet: + conditioner projection

        def run_add():
            state["i"] ^= 1
            ops.layernorm(y, xs[state["i"]], gamma, beta, add=adds[state["i"]])

        s = timed(run_add, iters)
        algo = B * T * C * 6.0
        print(json.dumps({"op": "ds_layernorm+add", "channels": C, "frames": B * T, "us": s * 1e6, "algo_bytes": algo,
                          "GBps": algo / s * 1e-9}))


if __name__ == "__main__":
    main()
