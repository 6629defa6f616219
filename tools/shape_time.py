"""Sampler-call time by batch shape at 20 x 512 (which kernel path a ragged group lands on, wave quantisation of the fused kernel)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import torch, sweep, diffsinger_b200 as ds
from diffsinger_b200.synthetic import random_state_dict
sweep.hp(diff_accelerator="dpm-solver")
m = ds.GaussianDiffusion(128, 1, 1000, 1000, "wavenet", dict(num_layers=20, num_channels=512, dilation_cycle_length=4), spec_min=[-12.0], spec_max=[0.0])
m.denoise_fn.load_state_dict(random_state_dict(128, 1, 20, 512, 256, seed=1))
m = m.eval().cuda()
for B, T in ((2, 1917), (27, 720), (1, 286), (4, 1917), (26, 579), (26, 640), (30, 720), (12, 550), (10, 539), (6, 720), (16, 600), (8, 600), (32, 600), (48, 600)):
    cond = torch.randn(B, T, 256, device="cuda")
    with torch.no_grad():
        ms = sweep.timed(lambda: m(cond, infer=True), 2)
    s = next(reversed(m.denoise_fn.engine().sessions.values()))
    print(f"B={B} T={T} tiles={B*-(-T//128)} pairs={B*-(-T//256)} path={s.path}: {ms:.1f} ms  ({ms/(B*-(-T//128)):.3f} ms/tile)", flush=True)
L = [504, 522, 539, 720, 1197, 550, 550, 515, 550, 539, 539, 540, 1917, 549, 550, 515, 687, 549, 549, 549, 460, 286, 578, 545, 515, 566, 572, 542, 516, 579]
conds = [torch.randn(1, t, 256, device="cuda") for t in L]
from diffsinger_b200 import segments as sg
for ov in (62, 150, 20):
    sg.CALL_OVERHEAD_TILES = ov
    with torch.no_grad():
        ms = sweep.timed(lambda: ds.sample_segments(m, conds), 2)
    groups = sg._ragged_groups(L, 512, 0.25, 0, ov)
    print(f"B=ragged overhead={ov}: {ms:.1f} ms, groups {[(len(g), L[g[0]]) for g in groups]}", flush=True)
    for s in m.denoise_fn.engine().sessions.values():
        if s._lengths is not None:
            print(f"B=  session B={s.B} T={s.T} path={s.path}", flush=True)
