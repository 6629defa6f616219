"""NSF-HiFiGAN generator on one B200: the native path (diffsinger_b200/vocoder.py) beside the same module
as plain torch ops (fp32 / TF32 and bf16 autocast) -- mel frames per second, 44.1 kHz configuration (hop 512).

    python tools/vocoder_bench.py [B T]
"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from diffsinger_b200.vocoder import DEFAULT_CONFIG, Generator


def main():
    a = [int(v) for v in sys.argv[1:]]
    B, T = a[:2] if len(a) >= 2 else (1, 2000)
    torch.manual_seed(0)
    gen = Generator(DEFAULT_CONFIG).eval().cuda()
    mel = torch.randn(B, 128, T, device="cuda") - 4.0
    f0 = 220.0 * torch.exp2(torch.rand(B, T, device="cuda"))
    # algorithmic MACs per mel frame (unpadded channels)
    h, ch, rate, macs = DEFAULT_CONFIG, DEFAULT_CONFIG["upsample_initial_channel"], 1, 7 * 128 * 512
    for u, k in zip(h["upsample_rates"], h["upsample_kernel_sizes"]):
        cin, ch = ch, ch // 2
        macs += rate * cin * ch * k                     # transposed conv: k taps per INPUT frame
        rate *= u
        macs += rate * ch * ch * 6 * sum(h["resblock_kernel_sizes"])
    macs += rate * ch * 7
    arms = os.environ.get("ARMS", "native,torch_bf16_autocast,torch_fp32").split(",")

    @torch.no_grad()
    def run(kind):
        if kind == "native":
            return gen(mel, f0)
        har = gen.m_source(f0, gen.upp).transpose(1, 2)
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=kind == "torch_bf16_autocast"):
            return gen._forward_torch(mel, har)

    for kind in ("native", "torch_bf16_autocast", "torch_fp32"):
        if kind not in arms:
            continue
        for _ in range(2):
            run(kind)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n = int(os.environ.get("STEPS", 5))
        e0.record()
        for _ in range(n):
            run(kind)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / n
        print(json.dumps(dict(arm=kind, B=B, T=T, ms=round(ms, 3), frames_per_s=round(B * T / ms * 1e3, 1),
                              realtime_factor=round(B * T * 512 / 44100 / (ms / 1e3), 1),
                              algorithmic_tflops=round(2 * macs * B * T / ms / 1e9, 1))), flush=True)


if __name__ == "__main__":
    main()
