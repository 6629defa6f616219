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
CALL_OVERHEAD_TILES = 62     # ... and the fixed cost of one sampler call there, in 128-frame tile-times (see _ragged_groups)


def _ragged_capacity(diffusion, t_longest: int, have_src: bool = False):
    """(tiles of 128 frames one ragged batch may hold, shortest segment of a batch as a fraction of its longest, (floor, overhead)
    of the grouping cost model in tiles); (0, 0, (0, 0)): this
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
        return 0, 0.0, (0, 0)  # ancestral chain draws fresh noise every step
    if getattr(diffusion, 'use_shallow_diffusion', False) and not have_src:
        return 0, 0.0, (0, 0)  # shallow diffusion needs the per-segment source spectrograms (src_spec=[...])
    if not hasattr(eng, 'plan'):  # LYNXNet: the SwiGLU GEMM masks the depthwise conv's input (lynx_engine.LynxWorkspace.set_lengths)
        return max(MAX_RAGGED_FRAMES // 128, -(-t_longest // 128)), 0.25, (0, CALL_OVERHEAD_TILES)
    cap = int(nat.lib().ds_plan_stack_capacity(eng.plan))
    if cap >= -(-t_longest // 128):
        return cap, 0.0, (cap, 0)  # one launch per evaluation whatever its size: fewest groups
    return max(MAX_RAGGED_FRAMES // 128, -(-t_longest // 128)), 0.25, (0, CALL_OVERHEAD_TILES)


def _ragged_groups(lengths: Sequence[int], cap: int, min_fill: float = 0.0, floor: int = 0, overhead: int = 0) -> List[List[int]]:
    """Segment indices grouped for ragged batches.  Segments are taken longest first and a group is a run of that order (similar
    lengths together), B * ceil(T_max / 128) <= cap tiles (a single segment is always allowed), every segment at least
    min_fill * T_max frames long.  Among those partitions the one with the least estimated cost is chosen, a group costing
    max(B * ceil(T_max / 128), floor) + overhead tile-times.  Whole-network kernel: floor = cap (one launch per evaluation costs
    the same whatever its size, i.e. fewest groups).  Per-layer kernels: overhead = the fixed cost of a sampler call in tile-times
    (measured at 20 x 512, 20 steps: ~9 ms + 0.145 ms per 128-frame tile -> 62)."""
    order = sorted(range(len(lengths)), key=lambda i: -lengths[i])
    n = len(order)
    if n == 0:
        return []
    tiles = [-(-lengths[i] // 128) for i in order]
    best = [0.0] + [float('inf')] * n   # best[k]: least cost of the first k segments
    cut = [0] * (n + 1)
    for k in range(1, n + 1):
        for j in range(k - 1, -1, -1):  # group = order[j:k], its longest segment is order[j]
            size = k - j
            if size > 1 and (size * tiles[j] > cap or lengths[order[k - 1]] < min_fill * lengths[order[j]]):
                break
            cost = best[j] + max(size * tiles[j], floor) + overhead + 1e-6 * size * tiles[j]
            if cost < best[k]:
                best[k], cut[k] = cost, j
    groups: List[List[int]] = []
    k = n
    while k > 0:
        groups.append(order[cut[k]:k])
        k = cut[k]
    return groups[::-1]


def _sample_ragged(diffusion, conditions, cap: int, min_fill: float, src_specs=None, aux_decoder=None, cost=(0, 0)) -> List[torch.Tensor]:
    """All batch-1 segments through ragged batches.  Noise is drawn segment by segment in the caller's order from the global
    generator -- the values a segment-by-segment loop would draw -- then padded; outputs are cut back to each segment's length."""
    dev = conditions[0].device
    lens = [int(c.shape[1]) for c in conditions]
    F, M = diffusion.num_feats, diffusion.out_dims
    noise = [torch.randn(1, F, M, t, device=dev) for t in lens]
    outs: List = [None] * len(conditions)
    for grp in _ragged_groups(lens, cap, min_fill, *cost):
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
        cap, min_fill, cost = _ragged_capacity(diffusion, max(int(c.shape[1]) for c in conditions),
                                                src_specs is not None or aux_decoder is not None)
        if cap:
            return _sample_ragged(diffusion, conditions, cap, min_fill, src_specs, aux_decoder, cost)
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


TOKEN_LEVEL_KEYS = ('tokens', 'languages')  # [1, T_txt] inputs of the encoder; every other tensor of a segment is per frame ([1, T]) or broadcast


@torch.no_grad()
def encode_segments(encoder, batches: Sequence[dict], max_frames: int = MAX_RAGGED_FRAMES) -> List[torch.Tensor]:
    """Conditions of many segments from the condition producer (diffsinger_b200.FastSpeech2Acoustic) in padded batches.

    batches: one dict per segment as the reference's preprocess_input builds it (inference/ds_acoustic.py:70-207): `tokens`
    [1, T_txt], `mel2ph` [1, T], `f0` [1, T], optionally `key_shift` / `speed` / variance curves [1, T], `languages` [1, T_txt],
    `spk_mix_embed` [1, T or 1, H].  Segments of similar length are padded together -- tokens with the padding index 0, frames with
    mel2ph = 0 -- which the encoder masks exactly (padded keys take no part in the attention, padded positions are zeroed after every
    sub-layer), so every segment's condition equals its batch-1 result.  Returns [1, T_i, H] per segment, in order."""
    if not batches:
        return []
    lens = [int(bt['mel2ph'].shape[1]) for bt in batches]
    outs: List = [None] * len(batches)
    for grp in _ragged_groups(lens, max(max_frames // 128, -(-max(lens) // 128)), 0.25, 0, 8):  # the encoder's call overhead is small
        t_max = max(lens[i] for i in grp)
        tt_max = max(int(batches[i]['tokens'].shape[1]) for i in grp)
        keys = [k for k, v in batches[grp[0]].items() if torch.is_tensor(v)]
        packed = {}
        for k in keys:
            first = batches[grp[0]][k]
            if first.dim() < 2:  # one value per segment (e.g. spk_embed_id [1])
                packed[k] = torch.cat([batches[i][k].reshape(1) for i in grp])
                continue
            n = tt_max if k in TOKEN_LEVEL_KEYS else (1 if all(batches[i][k].shape[1] == 1 for i in grp) else t_max)
            buf = first.new_zeros(len(grp), n, *first.shape[2:])
            for b, i in enumerate(grp):
                v = batches[i][k]
                if v.shape[1] == n:
                    buf[b] = v[0]
                elif v.shape[1] == 1:
                    buf[b] = v[0].expand(n, *v.shape[2:])
                else:
                    buf[b, :v.shape[1]] = v[0]
            packed[k] = buf
        extra = {k: v for k, v in packed.items() if k not in ('tokens', 'mel2ph', 'f0', 'key_shift', 'speed', 'languages', 'spk_embed_id')}
        spk_id = packed.get('spk_embed_id')
        if spk_id is not None and spk_id.dim() > 1:
            spk_id = spk_id[:, 0]
        cond = encoder(packed['tokens'], packed['mel2ph'], packed['f0'], key_shift=packed.get('key_shift'), speed=packed.get('speed'),
                       spk_embed_id=spk_id, languages=packed.get('languages'), **extra)
        for b, i in enumerate(grp):
            outs[i] = cond[b:b + 1, :lens[i]].contiguous()
    return outs


@torch.no_grad()
def infer_segments(encoder, diffusion, batches: Sequence[dict], aux_decoder=None, **kw) -> List[torch.Tensor]:
    """tokens -> mel for all segments of a `.ds` file: encode_segments (padded encoder batches), then sample_segments (ragged
    batches through the optional aux decoder and the sampler) -- the reference's DiffSingerAcoustic.forward (modules/toplevel.py:
    82-122) applied to its per-segment inference loop (inference/ds_acoustic.py:234-244) at once.  Every segment's result equals
    the one-by-one loop's (same noise draw order)."""
    conds = encode_segments(encoder, batches)
    return sample_segments(diffusion, conds, aux_decoder=aux_decoder, **kw)
