#!/usr/bin/env python
"""Headline benchmark: denoised mel frames/s of the diffusion-denoiser sampling loop.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl native|reference]

A "step" is one complete sampling run (DPM-Solver++ 2M, 20 backbone evaluations) over one batch of
synthetic utterances: BASELINE.json configs[1] = acoustic WaveNet 20 layers x 512 ch, 128 mel bins,
batch 64 x 2000 frames per GPU, bf16 tensor-core path.  Prints ONE JSON line (rank 0).
The reference arm (--impl reference) times the CPU oracle port of the same path (oracle/*.py, torch
fp32 on the box's host cores) on a bounded sample of the same workload.
"""
import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

CFG = dict(C=512, L=20, cycle=4, M=128, F=1, H=256, B=64, T=2000, K=20, speedup=50, timesteps=1000,
           algo="dpm-solver")


def algorithmic_flops_per_frame_eval(C, L, H, FM):
    """2 * [L*(8C^2 + 2HC) + C^2 + 2*F*M*C] (un-hoisted conditioner projection)."""
    return 2.0 * (L * (8 * C * C + 2 * H * C) + C * C + 2 * FM * C)


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        p = json.load(open(path))
        return dict(bf16=p["bf16_tflops"], bf16_sustained=p["bf16_tflops_sustained"], hbm=p["hbm_gbs"], src="measured")
    return dict(bf16=1590.0, bf16_sustained=1400.0, hbm=6650.0, src="fallback")


class ClockSampler(threading.Thread):
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index=0):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag = index, [], False

    def run(self):
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                self.rows.append([v.strip() for v in out.strip().split(",")])
            except Exception:
                pass
            time.sleep(0.2)

    def summary(self):
        self.stop_flag = True
        sm, mx, reasons = [], 0, set()
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx = max(mx, float(r[1]))
            except (ValueError, IndexError):
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons),
                "samples": len(sm)}


def cpu_oracle_run(B, T, threads):
    """Reference arm / cpu_baseline: the oracle port of GaussianDiffusion(dpm-solver) + WaveNet on the host."""
    from oracle.wavenet_oracle import OracleWaveNet, random_state_dict
    from oracle import samplers_oracle as so, solvers_oracle as sv
    torch.set_num_threads(threads)
    c = CFG
    sd = random_state_dict(c["M"], c["F"], c["L"], c["C"], c["H"], seed=1234)
    model = OracleWaveNet(sd, c["F"], c["cycle"])
    sch = so.Schedule(so.linear_betas(c["timesteps"]))
    g = torch.Generator().manual_seed(5678)
    cond = torch.randn(B, T, c["H"], generator=g).transpose(1, 2)
    noise = torch.randn(B, c["F"], c["M"], T, generator=g)
    t0 = time.perf_counter()
    x = sv.sample_dpm_solver_pp(model, sch.betas, noise, cond, c["K"])
    so.finish(x, torch.tensor([-12.0]), torch.tensor([0.0]))
    dt = time.perf_counter() - t0
    return B * T / dt, dt


def run_reference(args, rank, world):
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    B, T = 1, 500  # bounded sample of the B64 x T2000 workload (CPU throughput is flat in B)
    vals = []
    for i in range(args.warmup + args.steps):
        v, dt = cpu_oracle_run(B, T, threads)
        if i >= args.warmup:
            vals.append((v, dt))
    value = sum(v for v, _ in vals) / len(vals)
    ms = 1e3 * sum(dt for _, dt in vals) / len(vals)
    sample = f"B={B} x T={T} frames, {CFG['algo']} {CFG['K']} steps, fp32 torch-CPU oracle port"
    line = {
        "impl": "reference", "metric": "denoised mel frames/sec", "value": value, "unit": "frames/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_name(), "sample": sample},
        "cpu_baseline": {"value": value, "unit": "frames/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def workload_name():
    c = CFG
    if c.get("sampler") == "reflow-euler":  # BASELINE configs[2]: same backbone, RectifiedFlow, Euler, utterances sharded over the GPUs
        return (f"acoustic RectifiedFlow+WaveNet {c['L']}x{c['C']}ch, {c['M']} mel, batch {c['B']}x{c['T']} frames per GPU, "
                f"Euler {c['K']} steps (configs[2])")
    return (f"acoustic GaussianDiffusion+WaveNet {c['L']}x{c['C']}ch, {c['M']} mel, batch {c['B']}x{c['T']} frames per GPU, "
            f"DPM-Solver++ {c['K']} steps (configs[1])")


LAYER_KERNELS = {0: "ds_gemm_kernel<256, EPI_GATE> + ds_gemm_kernel<256, EPI_RES>", 1: "ds_layer_kernel<256, true>",
                 2: "ds_layer_kernel<256, false>", 3: "ds_layer_pair_kernel", 4: "ds_layer_halo_kernel", 5: "ds_layer_grp_kernel"}


def time_layer_kernel(session, L, iters=3):
    """CUDA-event timing (on the launching stream) of the dominant kernel: ONE launch = one fused residual layer
    (dilated conv + conditioner projection + gate + residual 1x1) over the whole B x T batch.  Launches walk the
    layers 0..L-2 like the real loop does, on the activations the last evaluation left in the session."""
    from diffsinger_b200 import _native as nat
    lib = nat.lib()
    st = torch.cuda.current_stream()
    n = 0
    for l in range(L - 1):
        nat.check(lib.ds_debug_layer(session.handle, l, 0, st.cuda_stream))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record(st)
    for _ in range(iters):
        for l in range(L - 1):
            nat.check(lib.ds_debug_layer(session.handle, l, 0, st.cuda_stream))
            n += 1
    e1.record(st)
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def measured_traffic(kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of `kernel` from the committed ncu --set full capture."""
    path = os.path.join(ROOT, "profiles", "layer_kernel_traffic.json")
    try:
        return json.load(open(path)).get(kernel)
    except (OSError, ValueError):
        return None


def run_native(args, rank, world, local_rank):
    import torch.distributed as dist
    import diffsinger_b200 as ds
    from diffsinger_b200.synthetic import random_state_dict
    c = CFG
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    ds.hparams.clear()
    ds.hparams.update(hidden_size=c["H"], schedule_type="linear", use_shallow_diffusion=False, diff_accelerator=c["algo"],
                      diff_speedup=c["speedup"], K_step=c["timesteps"], K_step_infer=c["timesteps"], timesteps=c["timesteps"])
    bb_args = dict(num_layers=c["L"], num_channels=c["C"], dilation_cycle_length=c["cycle"])
    if c.get("sampler") == "reflow-euler":
        ds.hparams.update(sampling_algorithm="euler", sampling_steps=c["K"], time_scale_factor=1000, T_start=0.0, T_start_infer=0.0)
        model = ds.RectifiedFlow(c["M"], c["F"], 0.0, 1000, "wavenet", bb_args, spec_min=[-12.0], spec_max=[0.0])
        model.denoise_fn = model.velocity_fn  # one name for the backbone below
    else:
        model = ds.GaussianDiffusion(c["M"], c["F"], c["timesteps"], c["timesteps"], "wavenet", bb_args, spec_min=[-12.0], spec_max=[0.0])
    model.denoise_fn.load_state_dict(random_state_dict(c["M"], c["F"], c["L"], c["C"], c["H"], seed=1234))
    model = model.eval().to(dev)
    B, T = c["B"], c["T"]
    g = torch.Generator().manual_seed(5678 + rank)
    cond_host = torch.randn(B, T, c["H"], generator=g).pin_memory()
    out_host = torch.empty(B, T, c["M"]).pin_memory()
    cond_dev = cond_host.to(dev)
    gathered = torch.empty(world * B, T, c["M"], device=dev) if world > 1 else None

    def step_resident():
        with torch.no_grad():
            y = model(cond_dev, infer=True)
        if world > 1:
            dist.all_gather_into_tensor(gathered, y.contiguous())
        return y

    def step_e2e():
        cd = cond_host.to(dev, non_blocking=True)
        with torch.no_grad():
            y = model(cd, infer=True)
        if world > 1:
            dist.all_gather_into_tensor(gathered, y.contiguous())
        out_host.copy_(y, non_blocking=True)
        torch.cuda.current_stream().synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sess = lambda: next(iter(model.denoise_fn.engine().sessions.values()))
    for _ in range(max(args.warmup, 3)):
        step_resident()
    barrier()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches = 0
    barrier()
    e0.record()
    for _ in range(args.steps):
        step_resident()
        launches += sess().last_launches
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1) / args.steps
    # end to end through the public API with host buffers: H2D of cond, sampling, D2H of the mel, all on the current stream,
    # timed on the device (events bracket K steps; step_e2e synchronises the stream after every step like a caller would)
    step_e2e()
    barrier()
    e0.record()
    for _ in range(args.steps):
        step_e2e()
    e1.record()
    barrier()
    ms_e2e = e0.elapsed_time(e1) / args.steps
    clocks = sampler.summary() if sampler else None
    t = torch.tensor([ms, ms_e2e], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, ms_e2e = float(t[0]), float(t[1])
    if rank != 0:
        return
    pk = peaks()
    frames = world * B * T
    value = frames / (ms * 1e-3)
    fpe = algorithmic_flops_per_frame_eval(c["C"], c["L"], c["H"], c["F"] * c["M"])
    step_tflops = B * T * c["K"] * fpe / (ms * 1e-3) / 1e12  # per GPU
    from diffsinger_b200 import _native as nat
    mode = int(nat.lib().ds_plan_layer_mode(model.denoise_fn.engine().plan))
    layer_ms = time_layer_kernel(sess(), c["L"])
    # algorithmic FLOPs of one launch: 2 * B*T * [2C*(3C+H) + C*C]  (the skip half of output_projection runs in the
    # deferred skip GEMM, so it is NOT counted here)
    layer_flops = 2.0 * B * T * (2 * c["C"] * (3 * c["C"] + c["H"]) + c["C"] * c["C"])
    achieved = layer_flops / (layer_ms * 1e-3) / 1e12
    threads = os.cpu_count() or 1
    cpu_baseline = None  # rank 0 at N = 1 only: a bounded sample of the same workload on the box's host cores
    if world == 1:
        cpu_v, cpu_dt = cpu_oracle_run(1, 500, threads)
        cpu_baseline = {"value": cpu_v, "unit": "frames/s", "cores": threads, "kind": "port",
                        "sample": "B=1 x T=500 frames, dpm-solver 20 steps, fp32 torch-CPU oracle port"}
    line = {
        "metric": "denoised mel frames/sec", "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16", "data": "synthetic",
        "config": {"workload": workload_name(), "l2": "working set per step (gated-activation store 2.6 GB/GPU) exceeds the 126 MB L2",
                   "parallelism": f"utterance shards, {world} x (B={B})", "nfe_per_step": c["K"]},
        "e2e": {"value": frames / (ms_e2e * 1e-3), "unit": "frames/s", "h2d_bytes_per_step": cond_host.numel() * 4,
                "d2h_bytes_per_step": out_host.numel() * 4},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": {"bound": "tensor", "achieved": achieved, "peak": pk["bf16"], "unit": "TFLOP/s", "frac": achieved / pk["bf16"],
                     "traffic": measured_traffic(LAYER_KERNELS[mode]), "kernel": LAYER_KERNELS[mode], "kernel_ms": layer_ms,
                     "flops_per_launch": layer_flops, "peak_source": pk["src"],
                     "step_tflops_per_gpu": step_tflops, "step_frac_of_sustained": step_tflops / pk["bf16_sustained"]},
        "cpu_baseline": cpu_baseline,
    }
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--sampler", default="dpm-solver", choices=["dpm-solver", "reflow-euler"],
                    help="dpm-solver = BASELINE configs[1] (the contract's workload); reflow-euler = configs[2] (native arm only)")
    ap.add_argument("--batch", type=int, default=None)
    ap.add_argument("--frames", type=int, default=None)
    args = ap.parse_args()
    CFG["sampler"] = args.sampler
    if args.batch:
        CFG["B"] = args.batch
    if args.frames:
        CFG["T"] = args.frames
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        # NCCL's own log lines (e.g. the version banner under NCCL_DEBUG=VERSION) go to stderr: stdout carries ONE JSON line
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":  # that level printf()s its banner to stdout
            os.environ["NCCL_DEBUG"] = "WARN"
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    try:
        run_native(args, rank, world, local_rank)
    finally:
        if world > 1:
            import torch.distributed as dist
            dist.destroy_process_group()


if __name__ == "__main__":
    main()
