"""GPU "before" number: the same denoiser as plain PyTorch eager ops on one B200.

/root/reference cannot travel to the GPU box, so this times this package's torch-op branch of the backbone
(``WaveNet._forward_autograd`` under no_grad: nn.Conv1d / nn.Linear / sigmoid / tanh exactly as modules/backbones/wavenet.py:33-48,
75-107 issues them) inside a 20-evaluation DDIM-style loop (two-term update per step, ddpm.py:159-167) -- the reference's extra
~340 scalar-schedule ops per step are left out, which only flatters the eager number.  fp32 (TF32 convolutions, torch's default)
and bf16 autocast.  Prints one JSON line per arm.

    PYTHONPATH=. python tools/eager_gpu_before.py [B T [C L]]
"""
import json
import sys

import torch

import diffsinger_b200 as ds
from diffsinger_b200.synthetic import random_state_dict


def main():
    a = [int(v) for v in sys.argv[1:]]
    B, T = a[:2] if len(a) >= 2 else (64, 2000)
    C, L = a[2:4] if len(a) >= 4 else (512, 20)
    M, H, NFE = 128, 256, 20
    ds.hparams.update(hidden_size=H)
    m = ds.WaveNet(M, 1, num_layers=L, num_channels=C, dilation_cycle_length=4)
    m.load_state_dict(random_state_dict(M, 1, L, C, H, seed=3), strict=True)
    m = m.cuda().eval()
    torch.manual_seed(0)
    cond = torch.randn(B, H, T, device='cuda')
    noise = torch.randn(B, 1, M, T, device='cuda')
    ts = torch.linspace(950, 0, NFE, device='cuda')

    @torch.no_grad()
    def run(autocast):
        x = noise.clone()
        for i in range(NFE):
            with torch.autocast('cuda', dtype=torch.bfloat16, enabled=autocast):
                eps = m._forward_autograd(x, ts[i].expand(B), cond)
            x = 0.99 * x - 0.05 * eps.float()
        return x

    for name, autocast in (("eager_fp32_tf32", False), ("eager_bf16_autocast", True)):
        for _ in range(2):
            run(autocast)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n = 3
        e0.record()
        for _ in range(n):
            run(autocast)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / n
        print(json.dumps(dict(arm=name, B=B, T=T, C=C, L=L, nfe=NFE, ms_per_batch=round(ms, 2),
                              frames_per_s=round(B * T / ms * 1e3, 1))), flush=True)


if __name__ == "__main__":
    main()
