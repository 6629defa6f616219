"""Glue between the sampler classes and the native session (device-memory plumbing only)."""
from __future__ import annotations

import os
from typing import List

import torch

from .. import _native as nat
from ..engine import Program


def to_time_major(x: torch.Tensor) -> torch.Tensor:
    """[B, F, M, T] -> contiguous time-major [B, T, F*M] fp32 (the layout every kernel uses)."""
    B, F, M, T = x.shape
    return x.reshape(B, F * M, T).transpose(1, 2).contiguous().float()


def shallow_start(src_btm: torch.Tensor, noise_bfmt: torch.Tensor, spec_min: torch.Tensor, spec_max: torch.Tensor, ca: float,
                  cb: float) -> torch.Tensor:
    """Time-major start state of shallow diffusion, ca * norm_spec(src) + cb * noise, in one kernel (include/dsb200.h:
    ds_shallow_start; bit-identical to norm_spec -> q_sample / linear mix -> to_time_major).  src [B, T, M], noise [B, 1, M, T]."""
    B, T, M = src_btm.shape
    src = src_btm.contiguous().float()
    noise = noise_bfmt.contiguous().float()
    smin = spec_min.reshape(-1).float().expand(M).contiguous() if spec_min.numel() == 1 else spec_min.reshape(-1).float().contiguous()
    smax = spec_max.reshape(-1).float().expand(M).contiguous() if spec_max.numel() == 1 else spec_max.reshape(-1).float().contiguous()
    assert smin.numel() == M and smax.numel() == M and noise.shape == (B, 1, M, T)
    out = torch.empty(B, T, M, dtype=torch.float32, device=src.device)
    nat.check(nat.lib().ds_shallow_start(out.data_ptr(), src.data_ptr(), noise.data_ptr(), smin.data_ptr(), smax.data_ptr(), float(ca),
                                         float(cb), B, T, M, torch.cuda.current_stream().cuda_stream))
    return out


def max_chunk_batch(eng, B: int, T: int, n_bufs: int, n_times: int) -> int:
    """Largest per-launch batch whose session workspace + sampler buffers fit the device (the gated-activation store
    alone is L*C*2 bytes per frame: 20 KB at C=512, L=20, so B=1024 x T=8000 would need 168 GB).  Utterances are
    independent end to end, so a big batch is simply run as consecutive chunks through the same session and graph.
    DSB200_MAX_FRAMES (frames per chunk) overrides the memory probe (used by the tests)."""
    cap = os.environ.get("DSB200_MAX_FRAMES")
    if cap:
        return max(1, min(B, int(cap) // T))
    key = (B, T, False, torch.cuda.current_stream().cuda_stream)
    if key in eng.sessions:  # already allocated: it fits
        return B
    per_utt = nat.lib().ds_session_workspace_bytes(eng.plan, 1, T, max(n_times, 64), 0) + n_bufs * T * eng.FM * 4
    free, _total = torch.cuda.mem_get_info(eng.device)
    free += torch.cuda.memory_reserved(eng.device) - torch.cuda.memory_allocated(eng.device)
    fit = int(0.8 * free) // max(per_utt, 1)
    if fit < 1:
        raise RuntimeError(f"not enough device memory for one utterance of {T} frames ({per_utt} bytes needed)")
    if fit >= B:
        return B
    n_chunks = -(-B // fit)
    return -(-B // n_chunks)  # equal-sized chunks: one session / graph serves all but possibly the last


@torch.no_grad()
def run_program(backbone, prog: Program, x_tm: torch.Tensor, cond_bht: torch.Tensor,
                extra: List[torch.Tensor] = (), use_graph: bool = True, lengths=None) -> torch.Tensor:
    """Runs `prog` with buffer 0 = x_tm (time-major start state), buffers 2.. optionally preset from
    `extra`.  cond_bht is the [B, H, T] tensor the reference passes to the backbone.  Returns buffer 0.
    lengths: optional per-utterance frame counts of a RAGGED batch (segments.sample_segments); frames beyond them are zero
    padding for the convolutions (each utterance keeps its batch-1 result) and their outputs are undefined.

    The backbone must be one of this package's native backbones (WaveNet, LYNXNet); anything else is rejected rather
    than silently emulated in PyTorch."""
    if not hasattr(backbone, 'engine'):
        raise TypeError("run_program needs a diffsinger_b200 backbone with a native engine (WaveNet, LYNXNet)")
    if use_graph and torch.cuda.is_current_stream_capturing():
        # the caller is capturing a bigger graph (e.g. tokens -> encoder -> sampler -> vocoder in one torch.cuda.CUDAGraph): the
        # program's kernels are launched into THAT capture instead of replaying / capturing the session's own graph
        use_graph = False
    eng = backbone.engine()
    if hasattr(eng, 'run_program'):  # LYNXNet: host-composed primitive ops, one torch CUDA graph per (program, B, T)
        return eng.run_program(backbone, prog, x_tm, cond_bht, extra, use_graph, lengths)
    B, T, FM = x_tm.shape
    bc = max_chunk_batch(eng, B, T, prog.n_bufs, len(prog.times))
    if bc >= B or lengths is not None:
        return _run_chunk(eng, prog, x_tm, cond_bht, extra, use_graph, lengths).clone()
    out = torch.empty_like(x_tm)
    for lo in range(0, B, bc):
        hi = min(B, lo + bc)
        out[lo:hi] = _run_chunk(eng, prog, x_tm[lo:hi], cond_bht[lo:hi], [e[lo:hi] for e in extra], use_graph)
    return out


def _run_chunk(eng, prog: Program, x_tm, cond_bht, extra, use_graph, lengths=None) -> torch.Tensor:
    B, T, FM = x_tm.shape
    sess = eng.session(B, T, max_times=len(prog.times), per_sample=False)
    sess.set_lengths(lengths)  # None restores a full-length batch on a session that ran ragged before
    bufs = sess_buffers(sess, prog.n_bufs)
    bufs[0].copy_(x_tm)
    for k, e in enumerate(extra):
        bufs[2 + k].copy_(e)
    sess.set_cond(cond_bht.transpose(1, 2))
    sess.set_program_times(prog.times)
    # 1000-step ancestral programs would make ~45k-node graphs; launch those eagerly
    sess.run(prog, bufs, use_graph=use_graph and len(prog.ops) <= 1024)
    return bufs[0]


def sess_buffers(sess, n: int):
    cache = getattr(sess, '_bufs', None)
    if cache is None:
        cache = sess._bufs = []
    while len(cache) < n:
        cache.append(torch.empty(sess.B, sess.T, sess.engine.FM, dtype=torch.float32, device=sess.engine.device))
    return cache[:n]
