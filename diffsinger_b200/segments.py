"""Segment-parallel sampling: many independent short utterances (the `.ds` segments inference/ds_acoustic.py feeds one by
one through `forward_model`) run concurrently on several CUDA streams.

One batch-1 segment of a few hundred frames cannot fill a B200 -- each of its kernels occupies at most a few dozen of
the 148 SMs and the 20-step loop is a latency-bound chain of ~900 small kernels -- but segments are independent, so
their CUDA graphs can run side by side.  Results are identical to calling the sampler segment by segment in the same
order: noise comes from the same global generator in the same host order, and every stream owns its sessions.
"""
from __future__ import annotations

from typing import List, Sequence

import torch


@torch.no_grad()
def sample_segments(diffusion, conditions: Sequence[torch.Tensor], n_streams: int = 4, **forward_kwargs) -> List[torch.Tensor]:
    """`diffusion(cond, infer=True, **forward_kwargs)` for every `cond` ([B, T_i, H], any lengths) with up to `n_streams`
    segments in flight.  Per-segment keyword tensors (e.g. `src_spec`) are passed as lists of the same length as
    `conditions`.  Returns the outputs in order; they are safe to use on the current stream."""
    if not conditions:
        return []
    dev = conditions[0].device
    if dev.type != 'cuda':
        raise RuntimeError("sample_segments needs CUDA tensors (there is no CPU path)")
    main = torch.cuda.current_stream(dev)
    n_streams = max(1, min(n_streams, len(conditions)))
    pool = getattr(diffusion, '_segment_streams', None)
    if pool is None or len(pool) < n_streams or pool[0].device != dev:
        pool = diffusion._segment_streams = [torch.cuda.Stream(device=dev) for _ in range(n_streams)]
    backbone = getattr(diffusion, 'denoise_fn', None) or getattr(diffusion, 'velocity_fn')
    eng = backbone.engine()  # weights are packed once, on the current stream, before any side stream starts
    # Segments on the whole-network kernel (one cluster per 128-frame tile; neighbouring tiles wait for each other's edge frames)
    # need all their clusters resident at once, so the tiles in flight over all streams stay within the device's capacity:
    # a segment with more than capacity / n_streams tiles runs alone (the streams meet before and after it), the others share.
    cap = 0
    if hasattr(eng, 'plan'):
        from . import _native as nat
        cap = int(nat.lib().ds_plan_stack_capacity(eng.plan))
    share = cap // n_streams if cap else 0
    outs: List[torch.Tensor] = []
    for st in pool[:n_streams]:
        st.wait_stream(main)
    k = 0
    for i, cond in enumerate(conditions):
        tiles = cond.shape[0] * -(-cond.shape[1] // 128)
        alone = bool(cap) and 1 < tiles <= cap and tiles > share
        st = pool[0] if alone else pool[k % n_streams]
        if alone:
            for other in pool[1:n_streams]:
                st.wait_stream(other)
        else:
            k += 1
        kw = {k_: (v[i] if isinstance(v, (list, tuple)) else v) for k_, v in forward_kwargs.items()}
        with torch.cuda.stream(st):
            y = diffusion(cond, infer=True, **kw)
        if alone:
            for other in pool[1:n_streams]:
                other.wait_stream(st)
        for t in (y if isinstance(y, (list, tuple)) else [y]):
            t.record_stream(main)
        outs.append(y)
    for st in pool[:n_streams]:
        main.wait_stream(st)
    return outs
