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


def sliced_noise(shape: Sequence[int], seed: int, lo: int, hi: int, device=None) -> torch.Tensor:
    """Noise for utterances [lo, hi) of a global batch `shape` = [B, ...]: utterance b always gets the stream of
    seed `seed + b`, so the result does not depend on the world size ('results are independent of world
    size')."""
    rows = []
    for b in range(lo, hi):
        g = torch.Generator().manual_seed(seed + b)
        rows.append(torch.randn(tuple(shape[1:]), generator=g))
    out = torch.stack(rows) if rows else torch.empty((0,) + tuple(shape[1:]))
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
