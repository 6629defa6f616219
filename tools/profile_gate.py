"""Runs only the dominant kernel (gate GEMM at configs[1] shapes) a few times -- the ncu --set full target."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402

if __name__ == "__main__":
    ms, flops = bench.time_gate_kernel(64, 2000, 512, 256, iters=4)
    print(f"gate kernel {ms:.4f} ms, {flops / ms / 1e9:.1f} TFLOP/s")
