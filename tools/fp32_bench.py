#!/usr/bin/env python
"""fp32-accurate mode (3 x bf16 split): the fused layer kernel's split instantiation vs the unfused GEMMs (DSB200_FP32_FUSED=0/1)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import sweep  # noqa: E402

for B, T in ((8, 1000), (64, 2000)):
    for fused, hoist in (("0", "0"), ("1", "0"), ("1", "1")):
        os.environ["DSB200_FP32_FUSED"], os.environ["DSB200_FP32_HOIST"] = fused, hoist
        print(f"DSB200_FP32_FUSED={fused} DSB200_FP32_HOIST={hoist}", end=" ", flush=True)
        sweep.acoustic("ddpm", "dpm-solver", B, T, iters=2, precision="fp32")
