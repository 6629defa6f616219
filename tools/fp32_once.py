"""One fp32-accurate sampler call at configs[1] shapes on the fused layer kernel's split instantiation (for ncu)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import sweep  # noqa: E402

sweep.acoustic("ddpm", "dpm-solver", 64, 2000, iters=1, precision="fp32")
