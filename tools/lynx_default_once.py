"""One sampler call of the shipped default LYNXNet config (launch list for ncu)."""
import sys
import os
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import sweep
sweep.lynx("dpm-solver", 32, 2000, iters=1, C=1024, strong=True)
