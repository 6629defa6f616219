"""Glue between the sampler classes and the native session (device-memory plumbing only)."""
from __future__ import annotations

from typing import List

import torch

from ..engine import Program


def to_time_major(x: torch.Tensor) -> torch.Tensor:
    """[B, F, M, T] -> contiguous time-major [B, T, F*M] fp32 (the layout every kernel uses)."""
    B, F, M, T = x.shape
    return x.reshape(B, F * M, T).transpose(1, 2).contiguous().float()


@torch.no_grad()
def run_program(backbone, prog: Program, x_tm: torch.Tensor, cond_bht: torch.Tensor,
                extra: List[torch.Tensor] = (), use_graph: bool = True) -> torch.Tensor:
    """Runs `prog` with buffer 0 = x_tm (time-major start state), buffers 2.. optionally preset from
    `extra`.  cond_bht is the [B, H, T] tensor the reference passes to the backbone.  Returns buffer 0.

    The backbone must be the native WaveNet; anything else (e.g. the reference's LYNXNet, which stays on
    the PyTorch path) is rejected rather than silently emulated."""
    if not hasattr(backbone, 'engine'):
        raise TypeError("run_program needs the diffsinger_b200 WaveNet backbone (native engine)")
    eng = backbone.engine()
    B, T, FM = x_tm.shape
    sess = eng.session(B, T, max_times=len(prog.times), per_sample=False)
    bufs = sess_buffers(sess, prog.n_bufs)
    bufs[0].copy_(x_tm)
    for k, e in enumerate(extra):
        bufs[2 + k].copy_(e)
    sess.set_cond(cond_bht.transpose(1, 2))
    sess.set_times(torch.tensor(prog.times, dtype=torch.float32))
    # 1000-step ancestral programs would make ~45k-node graphs; launch those eagerly
    sess.run(prog, bufs, use_graph=use_graph and len(prog.ops) <= 1024)
    return bufs[0].clone()


def sess_buffers(sess, n: int):
    cache = getattr(sess, '_bufs', None)
    if cache is None:
        cache = sess._bufs = []
    while len(cache) < n:
        cache.append(torch.empty(sess.B, sess.T, sess.engine.FM, dtype=torch.float32, device=sess.engine.device))
    return cache[:n]
