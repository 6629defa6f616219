#!/usr/bin/env python
"""Throughput sweep over BASELINE.json configs #2-#5 on ONE GPU (multi-GPU = utterance shards of these, bench.py):
   python tools/sweep.py [acoustic|reflow|variance|grid] ...   -> one JSON line per case (frames/s, ms, launches).
Synthetic inputs; timing = CUDA events around the public sampler call, cond resident in HBM."""
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import diffsinger_b200 as ds  # noqa: E402
from diffsinger_b200.synthetic import random_state_dict, random_lynx_state_dict  # noqa: E402

dev = "cuda"


def hp(**kw):
    ds.hparams.clear()
    ds.hparams.update(hidden_size=256, schedule_type="linear", use_shallow_diffusion=False, diff_speedup=50, K_step=1000,
                      K_step_infer=1000, timesteps=1000, sampling_steps=20, time_scale_factor=1000, T_start=0.0,
                      T_start_infer=0.0)
    ds.hparams.update(kw)


def timed(fn, iters):
    fn()
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def sessions_launches(backbone):
    return sum(s.last_launches for s in backbone.engine().sessions.values())


def acoustic(kind, algo, B, T, C=512, L=20, iters=3, precision="bf16"):
    M = 128
    if kind == "ddpm":
        hp(diff_accelerator=algo)
        m = ds.GaussianDiffusion(M, 1, 1000, 1000, "wavenet", dict(num_layers=L, num_channels=C, dilation_cycle_length=4),
                                 spec_min=[-12.0], spec_max=[0.0])
        bb = m.denoise_fn
    else:
        hp(sampling_algorithm=algo)
        m = ds.RectifiedFlow(M, 1, 0.0, 1000, "wavenet", dict(num_layers=L, num_channels=C, dilation_cycle_length=4),
                             spec_min=[-12.0], spec_max=[0.0])
        bb = m.velocity_fn
    bb.load_state_dict(random_state_dict(M, 1, L, C, 256, seed=1234))
    bb.set_precision(precision)
    m = m.eval().to(dev)
    cond = torch.randn(B, T, 256, device=dev)
    with torch.no_grad():
        ms = timed(lambda: m(cond, infer=True), iters)
    out = dict(case=f"{kind}/{algo}", precision=precision, C=C, L=L, B=B, T=T, ms=round(ms, 3), frames_per_s=round(B * T / ms * 1e3, 1),
               chunks=len(bb.engine().sessions), launches=sessions_launches(bb))
    print(json.dumps(out), flush=True)
    del m, cond
    torch.cuda.empty_cache()
    return out


def lynx(algo, B, T, iters=3, C=512, strong=False):
    """The reference's default acoustic backbone family (LYNXNet, expansion 2, k = 31).  configs/acoustic.yaml:72-78 ships
    6 x 1024 with strong_cond = true; 6 x 512 without it is the round-1 measurement point."""
    M, L = 128, 6
    hp(diff_accelerator=algo)
    m = ds.GaussianDiffusion(M, 1, 1000, 1000, "lynxnet", dict(num_layers=L, num_channels=C, expansion_factor=2, kernel_size=31,
                                                                strong_cond=strong),
                             spec_min=[-12.0], spec_max=[0.0])
    m.denoise_fn.load_state_dict(random_lynx_state_dict(M, 1, L, C, 256, seed=1234))
    m = m.eval().to(dev)
    cond = torch.randn(B, T, 256, device=dev)
    with torch.no_grad():
        ms = timed(lambda: m(cond, infer=True), iters)
    ws = next(iter(m.denoise_fn.engine().workspaces.values()))
    flops = 2.0 * B * T * 20 * (L * (256 * C + C * 4 * C + 2 * C * C) + 2 * M * C)  # 1x1 convs only (depthwise / norms are HBM-bound)
    print(json.dumps(dict(case=f"lynxnet/{algo}", C=C, L=L, strong_cond=strong, B=B, T=T, ms=round(ms, 3), frames_per_s=round(B * T / ms * 1e3, 1),
                          gemm_tflops=round(flops / ms / 1e9, 1), launches=ws.last_launches)), flush=True)
    del m, cond
    torch.cuda.empty_cache()


def lynx_segments(C=1024, strong=True):
    """The 30 segment lengths of config #1 on the shipped default backbone: one by one (the reference's loop) vs ragged batches."""
    M, L = 128, 6
    hp(diff_accelerator="dpm-solver")
    m = ds.GaussianDiffusion(M, 1, 1000, 1000, "lynxnet", dict(num_layers=L, num_channels=C, expansion_factor=2, kernel_size=31,
                                                                strong_cond=strong),
                             spec_min=[-12.0], spec_max=[0.0])
    m.denoise_fn.load_state_dict(random_lynx_state_dict(M, 1, L, C, 256, seed=1234))
    m = m.eval().to(dev)
    lens = [504, 522, 539, 720, 1197, 550, 550, 515, 550, 539, 539, 540, 1917, 549, 550, 515, 687, 549, 549, 549, 460, 286, 578, 545, 515,
            566, 572, 542, 516, 579]
    conds = [torch.randn(1, t, 256, device=dev) for t in lens]
    with torch.no_grad():
        torch.manual_seed(7)
        one = [m(c, infer=True) for c in conds]
        torch.manual_seed(7)
        rag = ds.sample_segments(m, conds)
        err = max(float((a - b).norm() / b.norm()) for a, b in zip(rag, one))
        ms_one = timed(lambda: [m(c, infer=True) for c in conds], 1)
        ms_rag = timed(lambda: ds.sample_segments(m, conds), 1)
    print(json.dumps(dict(case="lynxnet/segments30", C=C, L=L, strong_cond=strong, frames=sum(lens), one_by_one_ms=round(ms_one, 1),
                          ragged_ms=round(ms_rag, 1), ragged_frames_per_s=round(sum(lens) / ms_rag * 1e3, 1),
                          ragged_vs_one_by_one_rel_l2_max=round(err, 7))), flush=True)


def variance(T, iters=20):
    """config #4: PitchDiffusion (C256 L20 cycle5, 64 repeat bins) + MultiVarianceDiffusion (C192 L10 cycle4, 2 x 24 bins), B=1."""
    for algo in ("ddim", "dpm-solver", "unipc"):
        hp(diff_accelerator=algo)
        pitch = ds.PitchDiffusion(vmin=-8.0, vmax=8.0, cmin=-12.0, cmax=12.0, repeat_bins=64, timesteps=1000, k_step=1000,
                                  backbone_type="wavenet", backbone_args=dict(num_layers=20, num_channels=256, dilation_cycle_length=5))
        pitch.denoise_fn.load_state_dict(random_state_dict(64, 1, 20, 256, 256, seed=1))
        mv = ds.MultiVarianceDiffusion(ranges=[(-96.0, -12.0), (-96.0, -20.0)], clamps=[(-96.0, 0.0), (-96.0, 0.0)], repeat_bins=24,
                                       timesteps=1000, k_step=1000, backbone_type="wavenet",
                                       backbone_args=dict(num_layers=10, num_channels=192, dilation_cycle_length=4))
        mv.denoise_fn.load_state_dict(random_state_dict(24, 2, 10, 192, 256, seed=2))
        pitch, mv = pitch.eval().to(dev), mv.eval().to(dev)
        cond = torch.randn(1, T, 256, device=dev)
        with torch.no_grad():
            for name, m in (("pitch", pitch), ("multivar", mv)):
                ms = timed(lambda: m(cond, infer=True), iters)
                t0 = time.perf_counter()
                for _ in range(iters):
                    m(cond, infer=True)
                torch.cuda.synchronize()
                wall = (time.perf_counter() - t0) / iters * 1e3
                print(json.dumps(dict(case=f"variance/{name}/{algo}", B=1, T=T, ms=round(ms, 3), wall_ms=round(wall, 3),
                                      ms_per_sampler_step=round(ms / 20, 4), frames_per_s=round(T / ms * 1e3, 1),
                                      launches_in_graph=sessions_launches(m.denoise_fn))), flush=True)


def main():
    what = sys.argv[1:] or ["acoustic", "lynxnet", "reflow", "variance", "grid"]
    if what[0] == "one":  # one case, e.g. for a launch list:  sweep.py one ddpm dpm-solver B T [C L]
        a = what[1:]
        acoustic(a[0], a[1], int(a[2]), int(a[3]), *(int(v) for v in a[4:6]))
        return
    if "acoustic" in what:
        for algo in ("ddim", "pndm", "dpm-solver", "unipc"):
            if algo == "pndm":
                acoustic("ddpm", algo, 1, 2000)  # the reference's PNDM is batch-1 only
            acoustic("ddpm", algo, 64, 2000)
        acoustic("ddpm", "dpm-solver", 8, 1000, iters=2, precision="fp32")
    if "lynxnet" in what:
        for B, T in ((64, 2000), (8, 1000), (1, 500)):
            lynx("dpm-solver", B, T)
        for B, T in ((32, 2000), (1, 500)):  # the shipped default: 6 x 1024, strong_cond
            lynx("dpm-solver", B, T, C=1024, strong=True)
        lynx_segments()
    if "reflow" in what:
        for algo in ("euler", "rk2", "rk4"):
            acoustic("reflow", algo, 64, 2000, iters=2)
    if "variance" in what:
        for T in (128, 256, 512):
            variance(T)
    if "grid" in what:  # config #5 (one GPU's share)
        for T in (500, 1000, 2000, 4000, 8000):
            for B in (1, 4, 16, 64, 256, 1024):
                if B * T > 2_100_000:
                    continue  # bounded so the whole sweep stays within minutes; larger cases chunk identically
                acoustic("ddpm", "dpm-solver", B, T, iters=1 if B * T > 200_000 else 3)


if __name__ == "__main__":
    main()
