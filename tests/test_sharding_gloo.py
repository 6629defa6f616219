"""World-size-2 gloo tests of the multi-GPU host logic: shard ranges, world-size independent
noise, gather of the finished mels.  No GPU, no kernels."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from diffsinger_b200 import sharding


def test_shard_range_partitions():
    for n in (0, 1, 7, 64, 513):
        for world in (1, 2, 3, 8):
            blocks = [sharding.shard_range(n, world, r) for r in range(world)]
            assert blocks[0][0] == 0 and blocks[-1][1] == n
            for (a, b), (c, d) in zip(blocks, blocks[1:]):
                assert b == c
            sizes = [b - a for a, b in blocks]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        sharding.shard_range(4, 2, 2)


def test_shard_by_length_balances_frames():
    lengths = [504, 522, 539, 720, 1197, 550, 550, 515, 550, 539, 539, 540, 1917, 549, 550, 515, 687, 549, 549, 549,
               460, 286, 578, 545, 515, 566, 572, 542, 516, 579]  # samples/00 segment lengths
    for world in (2, 4, 8):
        parts = sharding.shard_by_length(lengths, world)
        assert sorted(i for p in parts for i in p) == list(range(len(lengths)))
        frames = [sum(lengths[i] for i in p) for p in parts]
        assert max(frames) - min(frames) <= max(lengths)


def test_noise_is_world_size_independent():
    full = sharding.sliced_noise((6, 1, 8, 10), 5678, 0, 6)
    for world in (2, 3):
        parts = [sharding.sliced_noise((6, 1, 8, 10), 5678, *sharding.shard_range(6, world, r)) for r in range(world)]
        assert torch.equal(torch.cat(parts), full)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n_total, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        lo, hi = sharding.shard_range(n_total, world, rank)
        # stand-in for "run the sampler on my shard": a deterministic per-utterance function of the sliced noise
        noise = sharding.sliced_noise((n_total, 4, 6), 99, lo, hi)
        local = noise * 2.0 + torch.arange(lo, hi, dtype=torch.float32)[:, None, None]
        full = sharding.gather_mels(local, n_total)
        q.put((rank, full.numpy().copy()))  # by value: a tensor would travel as a shared-memory handle that dies with this process
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n_total", [4, 5])
def test_gather_mels_world2(n_total):
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_total, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = dict(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    expect = sharding.sliced_noise((n_total, 4, 6), 99, 0, n_total) * 2.0 + torch.arange(n_total, dtype=torch.float32)[:, None, None]
    for r in range(world):
        assert torch.equal(torch.from_numpy(results[r]), expect)


class _StubDiffusion(torch.nn.Module):
    """Host-logic stand-in for a sampler: a deterministic per-utterance function of (cond row, noise row) that draws its noise
    the way GaussianDiffusion / RectifiedFlow do (`_draw_noise`, honouring the override sample_sharded installs)."""
    num_feats, out_dims = 1, 6

    def __init__(self):
        super().__init__()
        self.denoise_fn = torch.nn.Linear(1, 1)

    def forward(self, cond, infer=True):
        b, T = cond.shape[0], cond.shape[1]
        pre = self._noise_override
        assert pre is not None and tuple(pre.shape) == (b, 1, 6, T)
        return cond.sum(-1, keepdim=True) + pre.reshape(b, 6, T).transpose(1, 2)


def _sharded_worker(rank, world, port, n_total, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        cond = torch.randn(n_total, 9, 4, generator=torch.Generator().manual_seed(1))
        full = sharding.sample_sharded(_StubDiffusion(), cond, seed=321, noise_on_device=False)
        q.put((rank, full.numpy().copy()))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n_total", [4, 5])
def test_sample_sharded_world2_equals_world1(n_total):
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_sharded_worker, args=(r, world, port, n_total, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = dict(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    cond = torch.randn(n_total, 9, 4, generator=torch.Generator().manual_seed(1))
    expect = sharding.sample_sharded(_StubDiffusion(), cond, seed=321, noise_on_device=False)  # no process group: world 1
    parts = [sharding.sample_sharded(_StubDiffusion(), cond, seed=321, world=3, rank=r, gather=False, noise_on_device=False)[0]
             for r in range(3)]
    assert torch.equal(torch.cat(parts), expect)
    for r in range(world):
        assert torch.equal(torch.from_numpy(results[r]), expect)
