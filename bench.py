#!/usr/bin/env python
"""Headline benchmark: denoised mel frames/s of the diffusion-denoiser sampling loop.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl native|reference]

A "step" is one complete sampling run (DPM-Solver++ 2M, 20 backbone evaluations) over one batch of synthetic utterances:
BASELINE.json configs[1] = acoustic WaveNet 20 layers x 512 ch, 128 mel bins, batch 64 x 2000 frames per GPU, bf16
tensor-core path.  Prints ONE JSON line on stdout (rank 0); everything else any library prints goes to stderr.

  value      device-timed throughput with `cond` resident in HBM (public API: diffsinger_b200.sample_sharded -> the sampler's
             `forward(cond, infer=True)`; N > 1: utterance shards + one NCCL all-gather of the mels per step on a side stream)
  e2e        the same call with `cond` in pinned host memory and the mel read back to the host every step (copies double-
             buffered on a copy stream, inside the timed region)
  roofline   the dominant kernel (one fused residual layer per launch) timed over >= 2000 launches right after the timed loop
  extra      the other BASELINE configs: configs[2] (RectifiedFlow Euler-20, every N), configs[3] (variance predictors, latency),
             configs[4] corners, and config #1 on the real samples/00 .ds segments (needs oracle/_ref)
  cpu_baseline / --impl reference: the UNMODIFIED reference (oracle/_ref, see oracle/build_ref.py) on the box's host cores on a
             bounded sample of the same workload; the oracle port when the copy is missing (kind says which).
"""
import argparse
import json
import math
import os
import statistics
import subprocess
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

CFG = dict(C=512, L=20, cycle=4, M=128, F=1, H=256, B=64, T=2000, K=20, speedup=50, timesteps=1000, algo="dpm-solver")
CPU_SAMPLE = dict(B=1, T=2000)  # the reference arm's bounded sample of the B64 x T2000 workload (CPU throughput is flat in B)


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
                self.rows.append((time.perf_counter(), [v.strip() for v in out.strip().split(",")]))
            except Exception:
                pass
            time.sleep(0.1)

    def summary(self, t0=None, t1=None):
        sm, mx, reasons = [], 0, set()
        for ts, r in list(self.rows):
            if (t0 is not None and ts < t0) or (t1 is not None and ts > t1):
                continue
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


# ---------------------------------------------------------------------------------------------------- CPU arm
def reference_kind():
    from oracle.ref_harness import reference_available
    return "reference" if reference_available() else "port"


_CPU_CACHE = {}


def cpu_run(B, T, threads, algo="dpm-solver"):
    """One sampler call of the configs[1] network on the host cores: the unmodified reference's own
    `GaussianDiffusion(..., 'wavenet').forward(cond, infer=True)` when oracle/_ref (or /root/reference) exists, else the oracle
    port.  Returns (frames/s, seconds, kind)."""
    from oracle.wavenet_oracle import random_state_dict
    torch.set_num_threads(threads)
    c = CFG
    if "sd" not in _CPU_CACHE:
        _CPU_CACHE["sd"] = random_state_dict(c["M"], c["F"], c["L"], c["C"], c["H"], seed=1234)
    sd = _CPU_CACHE["sd"]
    g = torch.Generator().manual_seed(5678)
    cond = torch.randn(B, T, c["H"], generator=g)
    noise = torch.randn(B, c["F"], c["M"], T, generator=g)
    kind = reference_kind()
    if kind == "reference":
        from oracle import ref_runner as rr
        if ("ref", algo) not in _CPU_CACHE:
            _CPU_CACHE[("ref", algo)] = rr.build_sampler(algo, sd, M=c["M"], F=c["F"], L=c["L"], C=c["C"], cycle=c["cycle"], H=c["H"],
                                                         K=c["K"])
        _, dt = rr.run_sampler(_CPU_CACHE[("ref", algo)], cond, noise)
    else:
        from oracle.wavenet_oracle import OracleWaveNet
        from oracle import samplers_oracle as so, solvers_oracle as sv
        model = OracleWaveNet(sd, c["F"], c["cycle"])
        sch = so.Schedule(so.linear_betas(c["timesteps"]))
        t0 = time.perf_counter()
        x = sv.sample_dpm_solver_pp(model, sch.betas, noise, cond.transpose(1, 2), c["K"])
        so.finish(x, torch.tensor([-12.0]), torch.tensor([0.0]))
        dt = time.perf_counter() - t0
    return B * T / dt, dt, kind


def cpu_sample_text(kind):
    what = ("unmodified reference (modules.core.GaussianDiffusion + modules.backbones.wavenet.WaveNet from oracle/_ref)"
            if kind == "reference" else "oracle port (oracle/_ref missing)")
    return (f"B={CPU_SAMPLE['B']} x T={CPU_SAMPLE['T']} frames of the same network, {CFG['algo']} {CFG['K']} steps, fp32 torch-CPU, {what}")


def cpu_model():
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                return line.split(":", 1)[1].strip()
    except OSError:
        pass
    return "unknown"


def run_reference(args, rank, world, emit):
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    vals = []
    kind = reference_kind()
    for i in range(args.warmup + args.steps):
        v, dt, kind = cpu_run(CPU_SAMPLE["B"], CPU_SAMPLE["T"], threads)
        if i >= args.warmup:
            vals.append((v, dt))
    value = sum(v for v, _ in vals) / len(vals)
    ms = 1e3 * sum(dt for _, dt in vals) / len(vals)
    sample = cpu_sample_text(kind)
    emit({
        "impl": "reference", "metric": "denoised mel frames/sec", "value": value, "unit": "frames/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_name(), "sample": sample},
        "cpu_baseline": {"value": value, "unit": "frames/s", "cores": threads, "kind": kind, "sample": sample, "cpu": cpu_model()},
        "e2e": {"value": value, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    })


def workload_name():
    c = CFG
    return (f"acoustic GaussianDiffusion+WaveNet {c['L']}x{c['C']}ch, {c['M']} mel, batch {c['B']}x{c['T']} frames per GPU, "
            f"DPM-Solver++ {c['K']} steps (configs[1])")


LAYER_KERNELS = {0: "ds_gemm_kernel<256, EPI_GATE> + ds_gemm_kernel<256, EPI_RES>", 5: "ds_layer_grp_kernel"}  # ds_plan_layer_mode


def time_layer_kernel(session, L, min_launches=2000):
    """CUDA-event timing (on the launching stream) of the dominant kernel: ONE launch = one fused residual layer (dilated conv +
    conditioner projection + gate + residual 1x1) over the whole B x T batch.  >= `min_launches` back-to-back launches (~0.8 s)
    that walk the layers 0..L-2 like the real loop does, on the activations the last evaluation left in the session, issued
    right after the timed loop so that the chip is in the same power state.  Returns (ms per launch, launches, t0, t1)."""
    from diffsinger_b200 import _native as nat
    lib = nat.lib()
    st = torch.cuda.current_stream()
    iters = -(-min_launches // (L - 1))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    e0.record(st)
    n = 0
    for _ in range(iters):
        for l in range(L - 1):
            nat.check(lib.ds_debug_layer(session.handle, l, 0, st.cuda_stream))
            n += 1
    e1.record(st)
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n, n, t0, time.perf_counter()


def measured_traffic(kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of `kernel` from the committed ncu --set full capture
    (profiles/r02_layer_grp_kernel.txt); it is a property of the kernel and shape, not re-measured by this run."""
    path = os.path.join(ROOT, "profiles", "layer_kernel_traffic.json")
    try:
        return json.load(open(path)).get(kernel)
    except (OSError, ValueError):
        return None


def timed(fn, n, sync=True):
    """ms per call over n calls, CUDA events on the current stream."""
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


# ---------------------------------------------------------------------------------------------------- extras
def extra_variance(ds, dev, random_state_dict):
    """BASELINE configs[3]: the two sequential samplers of a variance segment (modules/toplevel.py:281,302), B = 1: pitch predictor
    (WaveNet 20 x 256, cycle 5, 64 repeat bins) and multi-variance predictor (10 x 192, 2 x 24 bins), 20 steps each; latency per
    call (median of 20 replays of the captured graph)."""
    out = {}
    H = 256
    for fam, hp in (("ddim", dict(diff_accelerator="ddim", diff_speedup=50)), ("euler", dict(sampling_algorithm="euler", sampling_steps=20))):
        ds.hparams.clear()
        ds.hparams.update(hidden_size=H, schedule_type="linear", use_shallow_diffusion=False, K_step=1000, K_step_infer=1000,
                          timesteps=1000, T_start=0.0, T_start_infer=0.0, time_scale_factor=1000, **hp)
        if fam == "ddim":
            pitch = ds.PitchDiffusion(vmin=-8.0, vmax=8.0, cmin=-12.0, cmax=12.0, repeat_bins=64, timesteps=1000, k_step=1000,
                                      backbone_type="wavenet", backbone_args=dict(num_layers=20, num_channels=256, dilation_cycle_length=5))
            multi = ds.MultiVarianceDiffusion(ranges=[(-96.0, -12.0), (-96.0, -20.0)], clamps=[(-96.0, 0.0), (-96.0, 0.0)], repeat_bins=24,
                                              timesteps=1000, k_step=1000, backbone_type="wavenet",
                                              backbone_args=dict(num_layers=10, num_channels=192, dilation_cycle_length=4))
            bp, bm = pitch.denoise_fn, multi.denoise_fn
        else:
            pitch = ds.PitchRectifiedFlow(vmin=-8.0, vmax=8.0, cmin=-12.0, cmax=12.0, repeat_bins=64, backbone_type="wavenet",
                                          backbone_args=dict(num_layers=20, num_channels=256, dilation_cycle_length=5))
            multi = ds.MultiVarianceRectifiedFlow(ranges=[(-96.0, -12.0), (-96.0, -20.0)], clamps=[(-96.0, 0.0), (-96.0, 0.0)],
                                                  repeat_bins=24, backbone_type="wavenet",
                                                  backbone_args=dict(num_layers=10, num_channels=192, dilation_cycle_length=4))
            bp, bm = pitch.velocity_fn, multi.velocity_fn
        bp.load_state_dict(random_state_dict(64, 1, 20, 256, H, seed=11))
        bm.load_state_dict(random_state_dict(24, 2, 10, 192, H, seed=12))
        pitch, multi = pitch.eval().to(dev), multi.eval().to(dev)
        for T in (128, 256, 512):
            cond = torch.randn(1, T, H, device=dev)
            with torch.no_grad():
                for m in (pitch, multi):
                    m(cond, infer=True)
                    m(cond, infer=True)
                tp = [timed(lambda: pitch(cond, infer=True), 1) for _ in range(20)]
                tm = [timed(lambda: multi(cond, infer=True), 1) for _ in range(20)]
            sp = next(iter(bp.engine().sessions.values()))
            out[f"{fam}_T{T}"] = {"pitch_ms": round(statistics.median(tp), 3), "multivar_ms": round(statistics.median(tm), 3),
                                  "pitch_path": sp.path, "pitch_launches": sp.last_launches}
    return {"workload": "PitchDiffusion/PitchRectifiedFlow (WaveNet 20x256 cycle 5, 64 bins) + MultiVariance* (10x192, 2x24 bins), "
                        "B=1, 20 steps, latency per sampler call (configs[3])", "unit": "ms", **out}


def extra_corners(model, dev, c):
    """Two corners of BASELINE configs[4] (the sweep): one short utterance (latency) and B=1024 x T=2000 (chunked through one
    session: the gated-activation store of 1024 utterances would need 84 GB)."""
    from diffsinger_b200.core.runtime import max_chunk_batch
    out = {}
    with torch.no_grad():
        cond = torch.randn(1, 500, c["H"], device=dev)
        model(cond, infer=True)
        model(cond, infer=True)
        ts = [timed(lambda: model(cond, infer=True), 1) for _ in range(10)]
        out["B1_T500"] = {"ms_per_call": round(statistics.median(ts), 3), "frames_per_s": round(500 / (statistics.median(ts) * 1e-3), 1)}
        # the 30 segment lengths of config #1 on THIS network (20 x 512): one by one (the reference's loop) vs ragged batches
        import diffsinger_b200 as ds
        lens = [504, 522, 539, 720, 1197, 550, 550, 515, 550, 539, 539, 540, 1917, 549, 550, 515, 687, 549, 549, 549, 460, 286, 578, 545,
                515, 566, 572, 542, 516, 579]
        conds = [torch.randn(1, t, c["H"], device=dev) for t in lens]
        torch.manual_seed(7)
        one = [model(cd, infer=True) for cd in conds]
        torch.manual_seed(7)
        rag = ds.sample_segments(model, conds)
        err = max(float((a - b).norm() / b.norm()) for a, b in zip(rag, one))
        ms_one = timed(lambda: [model(cd, infer=True) for cd in conds], 1)
        ds.sample_segments(model, conds)  # the one-by-one loop above walked 30 other sessions: touch the ragged ones again
        ms_rag = timed(lambda: ds.sample_segments(model, conds), 2)
        out["segments_30_ragged"] = {"frames": sum(lens), "one_by_one_ms": round(ms_one, 1), "ragged_batches_ms": round(ms_rag, 1),
                                     "ragged_frames_per_s": round(sum(lens) / (ms_rag * 1e-3), 1),
                                     "ragged_vs_one_by_one_rel_l2_max": round(err, 7)}
        del conds, one, rag
        eng = model.denoise_fn.engine()
        B, T = 1024, 2000
        bc = max_chunk_batch(eng, B, T, 3, c["K"])
        cond = torch.randn(B, T, c["H"], device=dev)
        ms = timed(lambda: model(cond, infer=True), 1)  # sessions / graphs of the chunk size already exist when bc == 64
        ms = timed(lambda: model(cond, infer=True), 1)
        out["B1024_T2000"] = {"ms_per_call": round(ms, 1), "frames_per_s": round(B * T / (ms * 1e-3), 1), "chunks": -(-B // bc),
                              "chunk_batch": bc}
        del cond
    return {"workload": f"same network and sampler as configs[1]; corners of the configs[4] sweep", **out}


def extra_config1(ds, dev, random_state_dict, threads):
    """BASELINE configs[0] on its REAL inputs: the 30 segments of samples/00_*.ds through the reference's own front end and encoder
    (oracle/_ref: preprocess_input -> tokens / mel2ph / f0), this package's condition producer (FastSpeech2Acoustic with the
    reference encoder's weights) on the GPU, then DDIM-20 over WaveNet 20 x 256, batch 1 per segment
    like inference/ds_acoustic.py:234-244.  GPU: segment by segment and several segments in flight (diffsinger_b200.sample_segments).
    CPU arm: the reference's own `model.diffusion(cond, infer=True)` on the first segments (bounded to ~10 s); those segments
    also give a final-mel parity figure on real inputs."""
    from oracle.ref_harness import reference_available
    if not reference_available():
        return {"unavailable": "oracle/_ref missing (run __graft_entry__.build() where /root/reference exists)"}
    from oracle import ref_runner as rr
    torch.set_num_threads(threads)
    ref_model, conds, batches = rr.config1_model_and_conditions()
    lens = [int(cd.shape[1]) for cd in conds]
    from oracle.ref_harness import load_reference
    ref_model_hparams = dict(load_reference().hparams)
    ds.hparams.clear()
    ds.hparams.update(hidden_size=256, schedule_type="linear", use_shallow_diffusion=False, diff_accelerator="ddim", diff_speedup=50,
                      K_step=1000, K_step_infer=1000, timesteps=1000)
    m = ds.GaussianDiffusion(128, 1, 1000, 1000, "wavenet", dict(num_layers=20, num_channels=256, dilation_cycle_length=4),
                             spec_min=ref_model.diffusion.spec_min.flatten().tolist(), spec_max=ref_model.diffusion.spec_max.flatten().tolist())
    m.load_state_dict(ref_model.diffusion.state_dict(), strict=True)  # the reference's checkpoint layout, unchanged
    m = m.eval().to(dev)
    # the condition producer on the GPU too: this package's FastSpeech2Acoustic with the reference encoder's checkpoint, fed the
    # reference front end's tensors (tokens, mel2ph, f0, key shift, speed)
    ds.hparams.update({k: ref_model_hparams[k] for k in ref_model_hparams})
    enc = ds.FastSpeech2Acoustic(ref_model.fs2.txt_embed.weight.shape[0])
    enc.load_state_dict(ref_model.fs2.state_dict(), strict=True)
    enc = enc.eval().to(dev)
    dbatches = [{k: (v.to(dev) if torch.is_tensor(v) else v) for k, v in bt.items()} for bt in batches]

    def encode(bt):
        return enc(bt['tokens'], bt['mel2ph'], bt['f0'], key_shift=bt.get('key_shift'), speed=bt.get('speed'), languages=bt.get('languages'))

    with torch.no_grad():
        dconds = [encode(bt) for bt in dbatches]
        enc_err = max(float((dc.cpu() - cd).norm() / cd.norm()) for dc, cd in zip(dconds, conds))
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for bt in dbatches:
            encode(bt)
        torch.cuda.synchronize()
        enc_ms = (time.perf_counter() - t0) * 1e3
    with torch.no_grad():
        for cd in dconds:  # sessions + graphs per distinct length
            m(cd, infer=True)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for cd in dconds:
            m(cd, infer=True)
        torch.cuda.synchronize()
        seq_ms = (time.perf_counter() - t0) * 1e3
        from diffsinger_b200 import _native as nat
        cap = int(nat.lib().ds_plan_stack_capacity(m.denoise_fn.engine().plan))
        n_streams = 6  # with ~36 tile clusters resident at once, six streams can each keep a <= 6-tile (768-frame) segment in flight
        ds.sample_segments(m, dconds, n_streams=n_streams, ragged=False)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        ds.sample_segments(m, dconds, n_streams=n_streams, ragged=False)
        torch.cuda.synchronize()
        par_ms = (time.perf_counter() - t0) * 1e3
        # ragged batches: segments of different lengths padded into one [B, T_max] launch per evaluation, every utterance with its
        # own end (ds_session_set_lengths); results are bit-identical to segment by segment (tests/test_gpu_stack.py)
        torch.manual_seed(99)
        one_by_one = [m(cd, infer=True) for cd in dconds]
        torch.manual_seed(99)
        rag = ds.sample_segments(m, dconds)
        rag_equal = all(torch.equal(a, b) for a, b in zip(one_by_one, rag))
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        ds.sample_segments(m, dconds)
        torch.cuda.synchronize()
        rag_ms = (time.perf_counter() - t0) * 1e3
        # tokens -> mel in one call: padded encoder batches + ragged sampler batches (diffsinger_b200.infer_segments)
        tensor_batches = [{k: bt[k] for k in ('tokens', 'mel2ph', 'f0', 'key_shift', 'speed', 'languages') if torch.is_tensor(bt.get(k))}
                          for bt in dbatches]
        torch.manual_seed(99)
        pipe = ds.infer_segments(enc, m, tensor_batches)
        pipe_err = max(float((a - b).norm() / b.norm()) for a, b in zip(pipe, one_by_one))
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(2):
            ds.infer_segments(enc, m, tensor_batches)
        torch.cuda.synchronize()
        pipe_ms = (time.perf_counter() - t0) * 1e3 / 2
    # CPU arm + parity on the same noise
    cpu_frames, cpu_s, errs = 0, 0.0, []
    for i, cd in enumerate(conds):
        if cpu_s > 8.0:
            break
        noise = torch.randn(1, 1, 128, lens[i], generator=torch.Generator().manual_seed(100 + i))
        y_ref, dt = rr.run_sampler(ref_model.diffusion, cd, noise)
        cpu_frames, cpu_s = cpu_frames + lens[i], cpu_s + dt
        m._noise_override = noise.to(dev)
        try:
            with torch.no_grad():
                y = m(dconds[i], infer=True)
        finally:
            m._noise_override = None
        errs.append(float((y.cpu() - y_ref).norm() / y_ref.norm()))
    frames = sum(lens)
    return {"workload": "samples/00 .ds: 30 segments (286-1917 frames, 18289 total), reference encoder -> cond, GaussianDiffusion+WaveNet "
                        "20x256, DDIM 20 steps, batch 1 per segment (configs[0])",
            "gpu_sequential_ms": round(seq_ms, 1), "gpu_sequential_frames_per_s": round(frames / (seq_ms * 1e-3), 1),
            "gpu_streams": n_streams, "stack_capacity_tiles": cap,
            "gpu_streams_ms": round(par_ms, 1), "gpu_streams_frames_per_s": round(frames / (par_ms * 1e-3), 1),
            "gpu_ragged_batches_ms": round(rag_ms, 1), "gpu_ragged_frames_per_s": round(frames / (rag_ms * 1e-3), 1),
            "gpu_ragged_equals_one_by_one": bool(rag_equal),
            "gpu_tokens_to_mel_ms": round(pipe_ms, 1), "gpu_tokens_to_mel_vs_one_by_one_rel_l2_max": round(pipe_err, 7),
            "gpu_encoder_ms_all_segments": round(enc_ms, 2), "encoder_cond_rel_l2_vs_reference_max": round(enc_err, 6),
            "cpu_reference_frames_per_s": round(cpu_frames / cpu_s, 1), "cpu_reference_segments": len(errs), "cpu_cores": threads,
            "final_mel_rel_l2_vs_reference": [round(e, 6) for e in errs]}


# ---------------------------------------------------------------------------------------------------- native arm
def run_native(args, rank, world, local_rank, emit):
    import torch.distributed as dist
    import diffsinger_b200 as ds
    from diffsinger_b200 import _native as nat
    from diffsinger_b200.synthetic import random_state_dict
    c = CFG
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)

    def hp_ddpm():
        ds.hparams.clear()
        ds.hparams.update(hidden_size=c["H"], schedule_type="linear", use_shallow_diffusion=False, diff_accelerator=c["algo"],
                          diff_speedup=c["speedup"], K_step=c["timesteps"], K_step_infer=c["timesteps"], timesteps=c["timesteps"],
                          sampling_algorithm="euler", sampling_steps=c["K"], time_scale_factor=1000, T_start=0.0, T_start_infer=0.0)

    hp_ddpm()
    bb_args = dict(num_layers=c["L"], num_channels=c["C"], dilation_cycle_length=c["cycle"])
    sd = random_state_dict(c["M"], c["F"], c["L"], c["C"], c["H"], seed=1234)
    model = ds.GaussianDiffusion(c["M"], c["F"], c["timesteps"], c["timesteps"], "wavenet", bb_args, spec_min=[-12.0], spec_max=[0.0])
    model.denoise_fn.load_state_dict(sd)
    model = model.eval().to(dev)
    B, T = c["B"], c["T"]
    g = torch.Generator().manual_seed(5678 + rank)
    cond_host = [torch.randn(B, T, c["H"], generator=g).pin_memory() for _ in range(2)]
    out_host = [torch.empty(B, T, c["M"]).pin_memory() for _ in range(2)]
    cond_dev = cond_host[0].to(dev)
    comm = torch.cuda.Stream(device=dev) if world > 1 else None
    copy = torch.cuda.Stream(device=dev)
    cur = torch.cuda.current_stream(dev)
    pending = []

    def sample(model_, cond, step):
        """The public call: this rank's shard of the global batch (its own rows of the world-size independent noise), then the
        NCCL gather of the finished mels on the side stream."""
        r = ds.sample_sharded(model_, cond, seed=1000 * step, local_rows=True, comm_stream=comm)
        if world > 1:
            pending.append(r[1])  # r = (gathered mels, event)
        return r

    def drain():
        for ev in pending:
            cur.wait_event(ev)
        pending.clear()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def resident_loop(model_, n):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        launches = 0
        barrier()
        e0.record()
        for k in range(n):
            sample(model_, cond_dev, k)
            launches += sess(model_).last_launches  # this library's kernels only (torch's Philox / copy kernels are not counted)
        drain()  # every gather has landed before the clock stops
        e1.record()
        barrier()
        return e0.elapsed_time(e1) / n, launches

    def sess(model_):
        bb = getattr(model_, "denoise_fn", None) or model_.velocity_fn
        return next(iter(bb.engine().sessions.values()))

    for k in range(max(args.warmup, 3)):
        sample(model, cond_dev, k)
    drain()
    barrier()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()
    t_loop0 = time.perf_counter()
    ms, launches = resident_loop(model, args.steps)
    t_loop1 = time.perf_counter()
    # the dominant kernel, timed right away (same power state), with its own clock window
    mode = int(nat.lib().ds_plan_layer_mode(model.denoise_fn.engine().plan))
    layer_ms, layer_n, tk0, tk1 = time_layer_kernel(sess(model), c["L"])

    # ---- end to end through the public API with HOST buffers: per step the H2D copy of that step's cond (pinned) and the D2H read of
    # its mel, double-buffered on a copy stream so that they overlap the previous / next step's kernels; timed on the device from the
    # first H2D to the last D2H.
    cond_bufs = [torch.empty_like(cond_dev) for _ in range(2)]
    h2d_done = [torch.cuda.Event() for _ in range(2)]
    step_done = [torch.cuda.Event() for _ in range(2)]

    def e2e_loop(n):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        copy.wait_stream(cur)
        e0.record(copy)
        with torch.cuda.stream(copy):
            cond_bufs[0].copy_(cond_host[0], non_blocking=True)
            h2d_done[0].record(copy)
        ys = [None, None]
        for k in range(n):
            b = k % 2
            cur.wait_event(h2d_done[b])
            r = sample(model, cond_bufs[b], k)
            y = r[0] if world > 1 else r
            ys[b] = y
            step_done[b].record(cur)
            with torch.cuda.stream(copy):
                if k + 1 < n:  # next step's conditions travel while this step computes
                    if k >= 1:
                        copy.wait_event(step_done[1 - b])  # buffer 1-b was last read by step k-1 (set_cond runs first in the call)
                    cond_bufs[1 - b].copy_(cond_host[(k + 1) % 2], non_blocking=True)
                    h2d_done[1 - b].record(copy)
                copy.wait_event(step_done[b])
                if world > 1:
                    copy.wait_event(r[1])
                    out_host[b].copy_(y[rank * B:(rank + 1) * B], non_blocking=True)
                else:
                    out_host[b].copy_(y, non_blocking=True)
                y.record_stream(copy)
        pending.clear()
        e1.record(copy)
        barrier()
        return e0.elapsed_time(e1) / n

    e2e_loop(2)
    ms_e2e = e2e_loop(args.steps)
    clocks = sampler.summary(t_loop0, t_loop1) if sampler else None
    kclocks = sampler.summary(tk0, tk1) if sampler else None

    # ---- BASELINE configs[2]: RectifiedFlow, Euler 20 steps, same backbone, 64 utterances per GPU -- at every N
    flow = ds.RectifiedFlow(c["M"], c["F"], 0.0, 1000, "wavenet", bb_args, spec_min=[-12.0], spec_max=[0.0])
    flow.velocity_fn.load_state_dict(sd)
    flow = flow.eval().to(dev)
    for k in range(2):
        sample(flow, cond_dev, k)
    drain()
    ms_flow, _ = resident_loop(flow, max(3, min(args.steps, 5)))
    flow.velocity_fn.engine().close()
    del flow

    t = torch.tensor([ms, ms_e2e, ms_flow], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, ms_e2e, ms_flow = (float(v) for v in t)
    if rank != 0:
        return
    pk = peaks()
    frames = world * B * T
    value = frames / (ms * 1e-3)
    fpe = algorithmic_flops_per_frame_eval(c["C"], c["L"], c["H"], c["F"] * c["M"])
    step_tflops = B * T * c["K"] * fpe / (ms * 1e-3) / 1e12  # per GPU
    # algorithmic FLOPs of one launch: 2 * B*T * [2C*(3C+H) + C*C]  (the skip half of output_projection runs in the deferred skip
    # GEMM, so it is NOT counted here)
    layer_flops = 2.0 * B * T * (2 * c["C"] * (3 * c["C"] + c["H"]) + c["C"] * c["C"])
    achieved = layer_flops / (layer_ms * 1e-3) / 1e12
    threads = os.cpu_count() or 1
    extra = {"configs2_reflow_euler": {
        "workload": f"acoustic RectifiedFlow+WaveNet {c['L']}x{c['C']}ch, {c['M']} mel, batch {B}x{T} frames per GPU ({world * B} utterances), "
                    f"Euler {c['K']} steps (configs[2])",
        "value": frames / (ms_flow * 1e-3), "unit": "frames/s", "ms_per_step": ms_flow, "n_gpus": world}}
    cpu_baseline = None  # rank 0 at N = 1 only: a bounded sample of the same workload on the box's host cores
    if world == 1:
        model.denoise_fn.engine()
        for name, fn in (("configs4_corners", lambda: extra_corners(model, dev, c)),
                         ("configs3_variance", lambda: extra_variance(ds, dev, random_state_dict)),
                         ("config1_real_ds", lambda: extra_config1(ds, dev, random_state_dict, threads))):
            try:
                extra[name] = fn()
            except Exception as e:  # an extra must never cost the headline line
                extra[name] = {"error": f"{type(e).__name__}: {e}"[:300]}
            hp_ddpm()
        cpu_v, cpu_dt, kind = cpu_run(CPU_SAMPLE["B"], CPU_SAMPLE["T"], threads)
        cpu_baseline = {"value": cpu_v, "unit": "frames/s", "cores": threads, "kind": kind, "sample": cpu_sample_text(kind),
                        "cpu": cpu_model()}
    emit({
        "metric": "denoised mel frames/sec", "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16", "data": "synthetic",
        "config": {"workload": workload_name(), "l2": "working set per step (gated-activation store 2.6 GB/GPU) exceeds the 126 MB L2",
                   "parallelism": f"utterance shards, {world} x (B={B}); mels gathered by one NCCL all-gather per step on a side stream",
                   "nfe_per_step": c["K"]},
        "e2e": {"value": frames / (ms_e2e * 1e-3), "unit": "frames/s", "h2d_bytes_per_step": cond_host[0].numel() * 4,
                "d2h_bytes_per_step": out_host[0].numel() * 4, "overlap": "double-buffered on a copy stream"},
        "gpu_launches": int(launches),
        "clocks": clocks,
        # the kernel is timed over ~2000 back-to-back launches (a long step: the chip sits at its power-capped clock), so the
        # SUSTAINED cuBLAS figure of MEASURED_PEAKS.json is the peak that applies; the burst figure is reported beside it
        "roofline": {"bound": "tensor", "achieved": achieved, "peak": pk["bf16_sustained"], "unit": "TFLOP/s",
                     "frac": achieved / pk["bf16_sustained"],
                     "peak_kind": "sustained (kernel timed inside a long run of back-to-back launches)",
                     "burst_peak": pk["bf16"], "frac_of_burst": achieved / pk["bf16"],
                     "frac_of_sustained": achieved / pk["bf16_sustained"],
                     "traffic": measured_traffic(LAYER_KERNELS.get(mode, "")), "traffic_source": "profiles/r02_layer_grp_kernel.txt (ncu --set full)",
                     "kernel": LAYER_KERNELS.get(mode, f"layer mode {mode}"), "kernel_ms": layer_ms, "kernel_launches_timed": layer_n,
                     "kernel_window_sm_mhz": kclocks["sm_mhz"] if kclocks else None,
                     "flops_per_launch": layer_flops, "peak_source": pk["src"],
                     "step_tflops_per_gpu": step_tflops, "step_frac_of_sustained": step_tflops / pk["bf16_sustained"]},
        "cpu_baseline": cpu_baseline,
        "extra": extra,
    })


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--batch", type=int, default=None)
    ap.add_argument("--frames", type=int, default=None)
    args = ap.parse_args()
    if args.batch:
        CFG["B"] = args.batch
    if args.frames:
        CFG["T"] = args.frames
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    # stdout carries ONE JSON line: whatever libraries print there (NCCL's version banner under NCCL_DEBUG=VERSION, the reference's
    # tqdm bars, ...) is sent to stderr by pointing fd 1 at fd 2 for the duration of the run; no environment knob is touched
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)

    def emit(line):
        sys.stdout.flush()
        os.write(real_stdout, (json.dumps(line) + "\n").encode())

    if args.impl == "reference":
        run_reference(args, rank, world, emit)
        return
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    try:
        run_native(args, rank, world, local_rank, emit)
    finally:
        if world > 1:
            import torch.distributed as dist
            dist.destroy_process_group()


if __name__ == "__main__":
    main()
