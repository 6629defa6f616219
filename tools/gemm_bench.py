"""Times one ds_gemm shape alone (CUDA events, operands larger than L2 in total): the LYNXNet C -> 2I SwiGLU product by default.
    python tools/gemm_bench.py [B T K N epi]        epi: 0 linear, 5 SwiGLU
DSB200_GEMM_PAIR=0 selects the single-CTA kernel; NOEPI=1 skips the pair kernel's epilogue (timing only: the output is not written)."""
import ctypes as C
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from diffsinger_b200 import _native as nat  # noqa: E402


def main():
    a = [int(v) for v in sys.argv[1:]]
    B, T, K, N, epi = (a + [64, 2000, 512, 2048, 5][len(a):])[:5]
    dev = "cuda"
    x = (torch.randn(B * T, K, device=dev)).to(torch.bfloat16)
    w = (torch.randn(N, K, device=dev) * K ** -0.5).to(torch.bfloat16)
    bias = torch.randn(N, device=dev)
    out = torch.empty(B * T, N // 2 if epi == 5 else N, device=dev, dtype=torch.bfloat16)
    d = nat.GemmDesc()
    d.epi, d.BN, d.B, d.T = epi, 256, B, T
    d.a0, d.a0_channels, d.a0_c0, d.kb_a0, d.n_taps, d.dil = x.data_ptr(), K, 0, K // 64, 1, 0
    d.w, d.w_rows, d.w_k, d.b_row0, d.n_tiles = w.data_ptr(), N, K, 0, N // 256
    d.bias, d.out, d.ld_out, d.out_c0, d.scale = bias.data_ptr(), out.data_ptr(), out.shape[1], 0, 1.0
    if os.environ.get("NOEPI") == "1":
        d.vec_bstride = -12345
    st = torch.cuda.current_stream().cuda_stream
    run = lambda: nat.check(nat.lib().ds_gemm(C.byref(d), st))
    for _ in range(3):
        run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        run()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 20
    print(json.dumps(dict(B=B, T=T, K=K, N=N, epi=epi, pair=os.environ.get("DSB200_GEMM_PAIR", "1"), noepi=os.environ.get("NOEPI", "0"),
                          us=round(ms * 1e3, 1), tflops=round(2.0 * B * T * K * N / ms * 1e-9, 1))))


if __name__ == "__main__":
    main()
