"""Utterance sharding across GPUs: one process per GPU, each rank runs the full
sampling loop on a contiguous block of utterances; the only collective is the gather of finished mels.

Every utterance (batch row) is independent end to end -- WaveNet's convs are per-sample and there is no
cross-sample op (wavenet.py:33-48) -- so there is no data-path collective inside the loop.  Mirrors the
reference's length-bucketed batching (utils/training_utils.py:115-120) by sorting on length first when the
lengths differ.
"""
from __future__ import annotations

from typing import List, Sequence, Tuple

import torch
import torch.distributed as dist


def shard_range(n_items: int, world: int, rank: int) -> Tuple[int, int]:
    """Contiguous [lo, hi) block of `n_items` for `rank`; block sizes differ by at most one, earlier ranks larger."""
    if world <= 0 or not 0 <= rank < world:
        raise ValueError(f"bad rank/world: {rank}/{world}")
    base, extra = divmod(n_items, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def shard_by_length(lengths: Sequence[int], world: int) -> List[List[int]]:
    """Utterance indices per rank: sort by length (longest first), deal round-robin in a serpentine so that every
    rank gets a similar number of frames.  Deterministic and independent of the rank that calls it."""
    order = sorted(range(len(lengths)), key=lambda i: (-lengths[i], i))
    out: List[List[int]] = [[] for _ in range(world)]
    for k, idx in enumerate(order):
        r = k % (2 * world)
        out[r if r < world else 2 * world - 1 - r].append(idx)
    return out


def sliced_noise(shape: Sequence[int], seed: int, lo: int, hi: int, device=None, on_device: bool = False) -> torch.Tensor:
    """Noise for utterances [lo, hi) of a global batch `shape` = [B, ...]: utterance b always gets the stream of
    seed `seed + b`, so the result does not depend on the world size ('results are independent of world
    size').  Default: drawn by the CPU generator and copied (the same values on every machine); `on_device=True` draws
    with a generator of `device` instead (Philox on CUDA: still a function of (seed + b) only, no host work or H2D copy)."""
    gen_dev = device if (on_device and device is not None) else "cpu"
    out = torch.empty((max(hi - lo, 0),) + tuple(shape[1:]), device=gen_dev)
    g = torch.Generator(device=gen_dev)
    for b in range(lo, hi):
        g.manual_seed(seed + b)
        torch.randn(tuple(shape[1:]), generator=g, out=out[b - lo])
    return out.to(device) if device is not None else out


def gather_mels(local: torch.Tensor, n_total: int, group=None) -> torch.Tensor:
    """All ranks get the [n_total, ...] tensor of finished mels in utterance order.  `local` is this rank's
    shard_range block.  Uneven blocks are padded to the largest block for the collective (all_gather needs equal
    sizes) and trimmed afterwards.  Backend: NCCL on GPUs, gloo in the CPU tests."""
    if not (dist.is_available() and dist.is_initialized()):
        assert local.shape[0] == n_total
        return local
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    sizes = [shard_range(n_total, world, r) for r in range(world)]
    cap = max(hi - lo for lo, hi in sizes)
    lo, hi = sizes[rank]
    assert local.shape[0] == hi - lo, (local.shape, lo, hi)
    pad = local.new_zeros((cap,) + tuple(local.shape[1:]))
    pad[: hi - lo] = local
    out = local.new_empty((world * cap,) + tuple(local.shape[1:]))
    dist.all_gather_into_tensor(out, pad.contiguous(), group=group)
    parts = [out[r * cap: r * cap + (h - l)] for r, (l, h) in enumerate(sizes)]
    return torch.cat(parts, dim=0)


def async_gather_mels(local: torch.Tensor, n_total: int, comm_stream: "torch.cuda.Stream", group=None):
    """`gather_mels` enqueued on `comm_stream` behind the work already on the current stream, so that the caller's next
    sampling call does not wait for the other ranks.  Needs equal blocks (n_total % world == 0).  Returns
    (full [n_total, ...], event): the tensor is valid once `event` has completed (`event.synchronize()` /
    `torch.cuda.current_stream().wait_event(event)`)."""
    world = dist.get_world_size(group)
    assert n_total % world == 0 and local.shape[0] == n_total // world
    cur = torch.cuda.current_stream(local.device)
    src = local.contiguous()
    out = local.new_empty((n_total,) + tuple(local.shape[1:]))
    comm_stream.wait_stream(cur)
    with torch.cuda.stream(comm_stream):
        dist.all_gather_into_tensor(out, src, group=group)
        ev = torch.cuda.Event()
        ev.record(comm_stream)
    src.record_stream(comm_stream)
    out.record_stream(comm_stream)
    return out, ev


@torch.no_grad()
def sample_sharded(diffusion, condition: torch.Tensor, seed: int, *, world: int | None = None, rank: int | None = None,
                   gather: bool = True, group=None, noise_on_device: bool = True, local_rows: bool = False,
                   comm_stream: "torch.cuda.Stream | None" = None, **forward_kwargs):
    """Sampling of a GLOBAL batch `condition` [B, T, H] partitioned by utterance over the ranks: this rank
    runs the full sampler -- `diffusion(condition[lo:hi], infer=True)`, one CUDA graph -- on its `shard_range` block with the
    rows [lo, hi) of the global noise batch (`sliced_noise`, seed + utterance index), then all ranks gather the finished
    mels (the only collective; NCCL).  The result is bit-identical for every world size, including 1.

    `condition` may be the global batch on every rank (host or device tensor; only rows [lo, hi) are moved / used).
    world / rank default to the initialised torch.distributed group (1 / 0 without one); passing them explicitly runs one
    shard without any collective (gather must then be False unless world == 1).  Per-utterance keyword tensors
    (e.g. `src_spec`) are sliced the same way.  `local_rows=True`: `condition` holds only this rank's rows of an evenly split
    global batch (B = rows x world) -- no rank ever materialises the global conditions.  `comm_stream`: the gather is enqueued
    on that stream instead of the current one (`async_gather_mels`) and (full, event) is returned, so the next call's kernels
    do not wait for the slowest rank.  Returns the gathered [B, ...] mels, or (local, lo, hi) when gather=False."""
    explicit = world is not None
    if world is None:
        inited = dist.is_available() and dist.is_initialized()
        world = dist.get_world_size(group) if inited else 1
        rank = dist.get_rank(group) if inited else 0
    T = condition.shape[1]
    backbone = getattr(diffusion, 'denoise_fn', None) or getattr(diffusion, 'velocity_fn')
    dev = next(backbone.parameters()).device
    if local_rows:
        B = condition.shape[0] * world
        lo, hi = rank * condition.shape[0], (rank + 1) * condition.shape[0]
        cond = condition.to(dev, non_blocking=True)
        kw = {k: (v.to(dev) if isinstance(v, torch.Tensor) else v) for k, v in forward_kwargs.items()}
    else:
        B = condition.shape[0]
        lo, hi = shard_range(B, world, rank)
        cond = condition[lo:hi].to(dev, non_blocking=True)
        kw = {k: (v[lo:hi].to(dev) if isinstance(v, torch.Tensor) and v.shape[:1] == (B,) else v) for k, v in forward_kwargs.items()}
    diffusion._noise_override = sliced_noise((B, diffusion.num_feats, diffusion.out_dims, T), seed, lo, hi, device=dev,
                                             on_device=noise_on_device)
    try:
        local = diffusion(cond, infer=True, **kw)
    finally:
        diffusion._noise_override = None
    if not gather:
        return local, lo, hi
    if world == 1:
        return local
    if explicit and not (dist.is_available() and dist.is_initialized()):
        raise RuntimeError("gather=True with an explicit world > 1 needs an initialised torch.distributed group")
    if isinstance(local, (list, tuple)):  # multi-variance wrappers return one curve per variance
        return [gather_mels(v, B, group) for v in local]
    if comm_stream is not None:
        return async_gather_mels(local, B, comm_stream, group)
    return gather_mels(local, B, group)
