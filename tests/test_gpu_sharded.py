"""Utterance-sharded sampling on the GPU path: `diffsinger_b200.sample_sharded` gives the same mels,
bit for bit, whatever the world size -- checked on one device by running the shards of a 2- and 3-rank partition one
after the other, and on two devices (when the box has them) through NCCL with one process per GPU."""
import os
import socket

import pytest
import torch

import diffsinger_b200 as ds
from oracle.wavenet_oracle import random_state_dict

pytestmark = pytest.mark.gpu
M, F, L, C, H, CYCLE, B, T = 128, 1, 4, 256, 256, 4, 6, 150


def _model(dev):
    ds.hparams.clear()
    ds.hparams.update(hidden_size=H, schedule_type="linear", use_shallow_diffusion=False, diff_accelerator="dpm-solver",
                      diff_speedup=100, K_step=1000, K_step_infer=1000, timesteps=1000)
    m = ds.GaussianDiffusion(M, F, 1000, 1000, "wavenet", dict(num_layers=L, num_channels=C, dilation_cycle_length=CYCLE),
                             spec_min=[-12.0], spec_max=[0.0])
    m.denoise_fn.load_state_dict(random_state_dict(M, F, L, C, H, seed=9))
    return m.eval().to(dev)


def _cond():
    return torch.randn(B, T, H, generator=torch.Generator().manual_seed(1))


def test_shards_run_one_after_the_other_equal_the_unsharded_batch():
    m = _model("cuda")
    cond = _cond()
    full = ds.sample_sharded(m, cond, seed=77)
    assert full.shape == (B, T, M) and torch.isfinite(full).all()
    for world in (2, 3):
        parts = [ds.sample_sharded(m, cond, seed=77, world=world, rank=r, gather=False)[0] for r in range(world)]
        assert torch.equal(torch.cat(parts), full)
    # a different seed gives different noise, the same seed the same mels
    assert not torch.equal(ds.sample_sharded(m, cond, seed=78), full)
    assert torch.equal(ds.sample_sharded(m, cond, seed=77), full)
    # host-generated noise (the same values on every machine) is also world-size independent
    full_h = ds.sample_sharded(m, cond, seed=77, noise_on_device=False)
    parts = [ds.sample_sharded(m, cond, seed=77, world=2, rank=r, gather=False, noise_on_device=False)[0] for r in range(2)]
    assert torch.equal(torch.cat(parts), full_h)


def _worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        m = _model(torch.device("cuda", rank))
        full = ds.sample_sharded(m, _cond(), seed=77)
        torch.cuda.synchronize()
        q.put((rank, full.cpu().numpy().copy()))
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs (gpurun --gpus 2)")
def test_two_ranks_over_nccl_equal_one_rank():
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    results = dict(q.get(timeout=300) for _ in range(2))
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    single = ds.sample_sharded(_model("cuda"), _cond(), seed=77).cpu()
    for r in range(2):
        assert torch.equal(torch.from_numpy(results[r]), single)
