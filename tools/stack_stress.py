#!/usr/bin/env python
"""Race hunt for ds_stack_kernel: the same multi-tile / ragged sampler calls repeated many times on several streams must give
bit-identical results every time (packet halo exchange, per-block DSMEM barriers, early dependent launch)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import diffsinger_b200 as ds  # noqa: E402
from diffsinger_b200.synthetic import random_state_dict  # noqa: E402


def main():
    reps = int(os.environ.get("REPS", 200))
    H = 256
    ds.hparams.clear()
    ds.hparams.update(hidden_size=H, diff_speedup=100, diff_accelerator="ddim", timesteps=1000, K_step=1000, use_shallow_diffusion=False,
                      schedule_type="linear")
    bad = 0
    for C, L, cycle, M in ((256, 20, 5, 64), (192, 10, 4, 48)):
        m = ds.GaussianDiffusion(M, 1, 1000, 1000, "wavenet", dict(num_layers=L, num_channels=C, dilation_cycle_length=cycle),
                                 spec_min=[-12.0], spec_max=[0.0])
        m.denoise_fn.load_state_dict(random_state_dict(M, 1, L, C, H, seed=C))
        m = m.eval().cuda()
        g = torch.Generator().manual_seed(1)
        lens = [1917, 640, 515, 300, 129, 77, 1024, 2048]
        conds = [torch.randn(1, t, H, generator=g).cuda() for t in lens]
        torch.manual_seed(3)
        ref = [y.clone() for y in ds.sample_segments(m, conds)]
        torch.manual_seed(3)
        ref_s = [y.clone() for y in ds.sample_segments(m, conds, ragged=False, n_streams=4)]
        assert all(torch.equal(a, b) for a, b in zip(ref, ref_s)), "ragged != streams"
        for i in range(reps):
            torch.manual_seed(3)
            ys = ds.sample_segments(m, conds, ragged=(i % 2 == 0), n_streams=4)
            torch.cuda.synchronize()
            for t, a, b in zip(lens, ys, ref):
                if not torch.equal(a, b):
                    bad += 1
                    print(f"C={C} rep {i}: segment of {t} frames differs, max abs {float((a - b).abs().max()):.3e}", flush=True)
        print(f"C={C} L={L}: {reps} repetitions x {len(lens)} segments, mismatches so far {bad}", flush=True)
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    main()
