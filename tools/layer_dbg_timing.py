#!/usr/bin/env python
"""Times ONE fused-layer launch (ds_debug_layer) at configs[1] shapes under the kernel's debug switches
(DSB200_DBG bit0: epilogue warps idle, bit1: producer issues no TMA loads, bit2: issuer commits without MMAs) to see
which role bounds the kernel.  Results with any bit set are wrong on purpose; this is a diagnosis tool, not a bench."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import diffsinger_b200 as ds  # noqa: E402
from diffsinger_b200 import _native as nat  # noqa: E402
from diffsinger_b200.synthetic import random_state_dict  # noqa: E402


def main():
    C, L, cycle, M, H = 512, 20, 4, 128, 256
    B = int(os.environ.get("B", 64))
    T = int(os.environ.get("T", 2000))
    ds.hparams.clear()
    ds.hparams.update(hidden_size=H)
    net = ds.WaveNet(M, 1, num_layers=L, num_channels=C, dilation_cycle_length=cycle)
    net.load_state_dict(random_state_dict(M, 1, L, C, H, seed=1234))
    net = net.eval().cuda()
    spec = torch.randn(B, 1, M, T, device="cuda")
    cond = torch.randn(B, T, H, device="cuda").transpose(1, 2)
    with torch.no_grad():
        net(spec, torch.tensor([500.0], device="cuda"), cond)
    sess = next(iter(net.engine().sessions.values()))
    lib = nat.lib()
    st = torch.cuda.current_stream()
    flops = 2.0 * B * T * (2 * C * (3 * C + H) + C * C)
    print('layer mode', nat.lib().ds_plan_layer_mode(net.engine().plan), flush=True)
    for dbg in [int(v) for v in os.environ.get('DBGS', '0,1,2,3,4,5,6,7').split(',')]:
        os.environ["DSB200_DBG"] = str(dbg)
        for layer in (0, 1, 2, 3):
            for _ in range(2):
                nat.check(lib.ds_debug_layer(sess.handle, layer, 0, st.cuda_stream))
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            e0.record(st)
            n = 5
            for _ in range(n):
                nat.check(lib.ds_debug_layer(sess.handle, layer, 0, st.cuda_stream))
            e1.record(st)
            torch.cuda.synchronize()
            us = 1e3 * e0.elapsed_time(e1) / n
            print(f"dbg={dbg} (epi_off={dbg & 1} tma_off={(dbg >> 1) & 1} mma_off={(dbg >> 2) & 1}) dil={1 << layer:2d}: "
                  f"{us:8.1f} us/launch  ({flops / us / 1e6:7.1f} TFLOP/s-equivalent)", flush=True)
    os.environ["DSB200_DBG"] = "0"


if __name__ == "__main__":
    main()
