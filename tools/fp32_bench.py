#!/usr/bin/env python
"""fp32-accurate mode (3 x bf16 split): the fused layer kernel's split instantiation vs the unfused GEMMs (DSB200_FP32_FUSED=0/1)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import sweep  # noqa: E402

for B, T in ((8, 1000), (64, 2000)):
    for fused in ("0", "1"):
        os.environ["DSB200_FP32_FUSED"] = fused
        print(f"DSB200_FP32_FUSED={fused}", end=" ", flush=True)
        sweep.acoustic("ddpm", "dpm-solver", B, T, iters=2, precision="fp32")
