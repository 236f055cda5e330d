"""CPU, world_size 2 over gloo: the data-parallel scheme of the step (shard the batch, sum-all-reduce the flat
gradient arena, scale by 1/N inside Adam) reproduces the single-process gradients of the global batch.
The arithmetic is the oracle's (test infrastructure); the plumbing under test is imagegeneration_b200/dp.py."""
import os
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from imagegeneration_b200.dp import allreduce_sum_, broadcast_, shard_batch
    from oracle import sndcgan_oracle as O
    torch.set_num_threads(1)
    image, B = (16, 32, 3), 4
    orc = O.OracleSNDCGAN(image, dropout=0.0, seed=62 + rank)      # deliberately different init per rank ...
    flat_w = torch.cat([v.reshape(-1) for v in orc.dv])
    broadcast_(flat_w, None, 0)                                     # ... made identical by the broadcast
    off = 0
    for v in orc.dv:
        v.copy_(flat_w[off:off + v.numel()].view_as(v))
        off += v.numel()
    g = torch.Generator().manual_seed(7)
    real = torch.rand((B, *image), generator=g) * 2 - 1
    dv = [v.clone().requires_grad_(True) for v in orc.dv]
    local = shard_batch(real, rank, world)
    loss = O.d_loss_fn("bce", O.discriminator_forward(dv, local, 0.0, None), 1.0)     # mean over the LOCAL batch
    flat_g = torch.cat([x.reshape(-1) for x in torch.autograd.grad(loss, dv)])
    work = allreduce_sum_(flat_g, None, async_op=True)
    work.wait()
    flat_g *= 1.0 / world                                          # what b2g_adam's gscale does
    if rank == 0:
        dv2 = [v.clone().requires_grad_(True) for v in orc.dv]
        full = O.d_loss_fn("bce", O.discriminator_forward(dv2, real, 0.0, None), 1.0)
        ref = torch.cat([x.reshape(-1) for x in torch.autograd.grad(full, dv2)])
        q.put(float((flat_g - ref).norm() / ref.norm()))
    dist.destroy_process_group()


def test_two_rank_gradients_equal_global_batch():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=300)
        assert p.exitcode == 0
    assert q.get(timeout=10) < 1e-5


def test_shard_batch():
    from imagegeneration_b200.dp import shard_batch
    t = torch.arange(8).view(8, 1)
    assert shard_batch(t, 1, 4).flatten().tolist() == [2, 3]
    with pytest.raises(ValueError):
        shard_batch(t, 0, 3)
