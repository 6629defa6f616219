"""Segment-parallel sampling: many independent short utterances (the `.ds` segments inference/ds_acoustic.py feeds one by
one through `forward_model`) run together -- as RAGGED BATCHES on the whole-network kernel where the
backbone allows it (one launch per evaluation covers several segments of different lengths, each ending in its own zero
padding exactly as if it ran alone), otherwise concurrently on several CUDA streams.

One batch-1 segment of a few hundred frames cannot fill a B200 -- each of its kernels occupies at most a few dozen of
the 148 SMs and the 20-step loop is a latency-bound chain of ~900 small kernels -- but segments are independent, so
their CUDA graphs can run side by side.  Results are identical to calling the sampler segment by segment in the same
order: noise comes from the same global generator in the same host order, and every stream owns its sessions.
"""
from __future__ import annotations

from typing import List, Sequence

import torch


MAX_RAGGED_FRAMES = 1 << 16  # padded frames (B * T_max) of one ragged batch on the per-layer kernels


def _ragged_capacity(diffusion, t_longest: int, have_src: bool = False):
    """(tiles of 128 frames one ragged batch may hold, shortest segment of a batch as a fraction of its longest); (0, 0): this
    sampler / backbone / configuration cannot run ragged batches.  Plans on the whole-network kernel are bounded by the
    co-resident clusters (ds_plan_stack_capacity) as long as the longest segment fits, and padding costs them nothing (tiles beyond
    an utterance's end exit at once); everything else runs the per-layer kernels, which compute the padding too: batches of at most
    MAX_RAGGED_FRAMES padded frames whose segments are at least half as long as the longest."""
    from . import _native as nat
    from .hparams import get_hparams
    backbone = getattr(diffusion, 'denoise_fn', None) or getattr(diffusion, 'velocity_fn')
    eng = backbone.engine()
    hp = get_hparams()
    if hasattr(diffusion, 'denoise_fn') and hp.get('diff_speedup', 0) <= 1:
        return 0, 0.0  # ancestral chain draws fresh noise every step
    if getattr(diffusion, 'use_shallow_diffusion', False) and not have_src:
        return 0, 0.0  # shallow diffusion needs the per-segment source spectrograms (src_spec=[...])
    if not hasattr(eng, 'plan'):  # LYNXNet: the SwiGLU GEMM masks the depthwise conv's input (lynx_engine.LynxWorkspace.set_lengths)
        return max(MAX_RAGGED_FRAMES // 128, -(-t_longest // 128)), 0.5
    cap = int(nat.lib().ds_plan_stack_capacity(eng.plan))
    if cap >= -(-t_longest // 128):
        return cap, 0.0
    return max(MAX_RAGGED_FRAMES // 128, -(-t_longest // 128)), 0.5


def _ragged_groups(lengths: Sequence[int], cap: int, min_fill: float = 0.0) -> List[List[int]]:
    """Segment indices grouped for ragged batches: similar lengths together (longest first), B * ceil(T_max / 128) <= cap tiles,
    every segment at least min_fill * T_max frames long."""
    order = sorted(range(len(lengths)), key=lambda i: -lengths[i])
    groups: List[List[int]] = []
    cur: List[int] = []
    for i in order:
        t_max = lengths[cur[0]] if cur else lengths[i]
        if cur and ((len(cur) + 1) * -(-t_max // 128) > cap or lengths[i] < min_fill * t_max):
            groups.append(cur)
            cur = []
        cur.append(i)
    if cur:
        groups.append(cur)
    return groups


def _sample_ragged(diffusion, conditions, cap: int, min_fill: float, src_specs=None, aux_decoder=None) -> List[torch.Tensor]:
    """All batch-1 segments through ragged batches.  Noise is drawn segment by segment in the caller's order from the global
    generator -- the values a segment-by-segment loop would draw -- then padded; outputs are cut back to each segment's length."""
    dev = conditions[0].device
    lens = [int(c.shape[1]) for c in conditions]
    F, M = diffusion.num_feats, diffusion.out_dims
    noise = [torch.randn(1, F, M, t, device=dev) for t in lens]
    outs: List = [None] * len(conditions)
    for grp in _ragged_groups(lens, cap, min_fill):
        t_max = max(lens[i] for i in grp)
        cond = conditions[0].new_zeros(len(grp), t_max, conditions[0].shape[2])
        z = torch.zeros(len(grp), F, M, t_max, device=dev)
        src = None if src_specs is None else src_specs[0].new_zeros(len(grp), t_max, *src_specs[0].shape[2:])
        for b, i in enumerate(grp):
            cond[b, :lens[i]] = conditions[i][0]
            z[b, ..., :lens[i]] = noise[i][0]
            if src is not None:
                src[b, :lens[i]] = src_specs[i][0]
        if aux_decoder is not None:  # the shallow-diffusion source of the whole ragged batch in one pass
            src = aux_decoder(cond, infer=True, lengths=[lens[i] for i in grp])
        saved = getattr(diffusion, '_noise_override', None), getattr(diffusion, '_lengths_override', None)
        diffusion._noise_override, diffusion._lengths_override = z, [lens[i] for i in grp]
        try:
            y = diffusion(cond, infer=True) if src is None else diffusion(cond, src_spec=src, infer=True)
        finally:
            diffusion._noise_override, diffusion._lengths_override = saved
        for b, i in enumerate(grp):
            if isinstance(y, (list, tuple)):   # multi-variance: list of [B, T] curves
                outs[i] = [c[b:b + 1, :lens[i]].clone() for c in y]
            elif y.dim() == 4:                 # [B, F, T, M]
                outs[i] = y[b:b + 1, :, :lens[i]].clone()
            else:                              # [B, T, M] mel or [B, T] curve
                outs[i] = y[b:b + 1, :lens[i]].clone()
    return outs


@torch.no_grad()
def sample_segments(diffusion, conditions: Sequence[torch.Tensor], n_streams: int = 4, ragged: bool = True, aux_decoder=None,
                    **forward_kwargs) -> List[torch.Tensor]:
    """`diffusion(cond, infer=True, **forward_kwargs)` for every `cond` ([B, T_i, H], any lengths).  Returns the outputs in
    order; they are safe to use on the current stream.

    ragged=True (default): when every segment is batch 1 (keyword tensors: at most `src_spec=[...]`, the shallow-diffusion sources
    [1, T_i, M]), segments are padded into ragged batches (include/dsb200.h: ds_session_set_lengths; LYNXNet: the SwiGLU GEMM masks the
    depthwise conv's input): every utterance keeps its own end, so each output equals the segment-by-segment result.  Otherwise up to
    `n_streams` segments are in flight on CUDA streams;
    per-segment keyword tensors (e.g. `src_spec`) are passed as lists of the same length as `conditions`.

    aux_decoder: a diffsinger_b200.AuxDecoderAdaptor -- the shallow-diffusion sources are then computed here (for a ragged batch in
    one pass with the same lengths), i.e. the call covers `aux_decoder(condition, infer=True)` -> `diffusion(condition, src_spec=...)`
    of modules/toplevel.py:94-103."""
    if not conditions:
        return []
    dev = conditions[0].device
    if dev.type != 'cuda':
        raise RuntimeError("sample_segments needs CUDA tensors (there is no CPU path)")
    if aux_decoder is not None:
        if 'src_spec' in forward_kwargs:
            raise ValueError("pass either aux_decoder or src_spec")
        if not hasattr(aux_decoder, 'decoder') or not hasattr(aux_decoder.decoder, '_forward_native'):
            raise TypeError("aux_decoder must be a diffsinger_b200.AuxDecoderAdaptor")
    src_specs = forward_kwargs.get('src_spec')
    only_src = set(forward_kwargs) <= {'src_spec'} and (src_specs is None or (isinstance(src_specs, (list, tuple)) and
                                                                             all(v.dim() == 3 for v in src_specs)))
    if ragged and only_src and all(c.shape[0] == 1 for c in conditions) and len(conditions) > 1:
        cap, min_fill = _ragged_capacity(diffusion, max(int(c.shape[1]) for c in conditions), src_specs is not None or aux_decoder is not None)
        if cap:
            return _sample_ragged(diffusion, conditions, cap, min_fill, src_specs, aux_decoder)
    if aux_decoder is not None:  # the stream path: one source per segment, as the reference's DiffSingerAcoustic.forward computes it
        forward_kwargs = dict(forward_kwargs, src_spec=[aux_decoder(c, infer=True) for c in conditions])
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
