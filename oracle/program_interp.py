"""TEST INFRASTRUCTURE ONLY -- CPU interpreter for diffsinger_b200 sampler Programs.

Executes the EVAL / LINCOMB ops the product hands to libdsb200.so, but with the oracle WaveNet
(oracle/wavenet_oracle.py) as the backbone and torch-CPU arithmetic, so that the host-side
"sampler compiler" (diffsinger_b200/core/programs.py) can be checked against the reference's golden
vectors without a GPU.  Never imported by the product.
"""
import torch

DS_OP_EVAL, DS_OP_LINCOMB = 0, 1


@torch.no_grad()
def run_program_cpu(model, prog, x_tm, cond_bht, n_feats, extra=()):
    """model(spec[B,F,M,T], t[1], cond[B,H,T]); x_tm time-major [B,T,F*M]. Returns buffer 0."""
    B, T, FM = x_tm.shape
    M = FM // n_feats
    bufs = [torch.zeros_like(x_tm) for _ in range(prog.n_bufs)]
    bufs[0] = x_tm.clone()
    for k, e in enumerate(extra):
        bufs[2 + k] = e.clone()
    times = torch.tensor(prog.times, dtype=torch.float32)
    for kind, dst, time, srcs, coefs in prog.ops:
        if kind == DS_OP_EVAL:
            spec = bufs[srcs[0]].transpose(1, 2).reshape(B, n_feats, M, T)
            y = model(spec, times[time].reshape(1), cond_bht)
            bufs[dst] = y.reshape(B, FM, T).transpose(1, 2).contiguous()
        else:
            acc = torch.zeros_like(x_tm)
            for c, s in zip(coefs, srcs):
                acc = acc + torch.tensor(c, dtype=torch.float32) * bufs[s]
            bufs[dst] = acc
    return bufs[0]
