"""CPU tests of the WGAN host logic (imagegeneration_b200/wgan.py): arena layout invariants the fused RMSprop / clip /
gan-optimizer launches rely on, Keras variable order, and the world_size-2 data-parallel arithmetic on gloo
(sum all-reduce of the gradient arena, 1/N inside the optimizer, BatchNorm range of the critic reduced separately)."""
import os

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import wgan_oracle as W


def test_critic_arena_layout():
    from imagegeneration_b200.wgan import Critic
    c = Critic((16, 32, 3), device="cpu")
    names = c.arena.names
    assert all(n.endswith("kernel") and n.startswith("conv") for n in names[:7])        # the clipped range comes first
    assert c.clip_n == c.arena.offsets[7] and c.clip_n % 4 == 0
    assert c.clip_n == sum(v.numel() for v in c.arena.variables[:7])                    # no padding inside it
    assert all("bn" in n for n in names[16:]) and c.bn_off == c.arena.offsets[16]       # gamma/beta form the tail
    assert c.bn_off % 4 == 0 and c.bn_n % 4 == 0 and c.bn_off + c.bn_n == c.arena.numel
    assert [tuple(v.shape) for v in c.trainable_variables] == [s for _, s in W.critic_var_shapes((16, 32, 3))]
    # Keras-order views alias the arena
    c.trainable_variables[2].fill_(3.0)                                                  # bn1/gamma
    assert float(c.arena.w[c.bn_off]) == 3.0
    # initialisers: RandomNormal(0.02) kernels, ones / zeros BatchNorm, zero biases
    assert 0.015 < float(c.arena.variables[6].std()) < 0.025
    assert float(c.arena.variables[7].abs().sum()) == 0.0


def test_generator_variables_and_rejections():
    from imagegeneration_b200.lib import B200GanError
    from imagegeneration_b200.wgan import WGANEngine, WGenerator
    g = WGenerator(128, (16, 32, 3), device="cpu")
    assert [tuple(v.shape) for v in g.trainable_variables] == [s for _, s in W.generator_var_shapes((16, 32, 3))]
    with pytest.raises(B200GanError):
        WGenerator(64, (16, 32, 3), device="cpu")
    with pytest.raises(B200GanError):
        WGenerator(128, (20, 32, 3), device="cpu")
    with pytest.raises(B200GanError):                       # no CPU fallback: a CPU batch is refused before any launch
        g.forward(torch.zeros(2, 128), training=True)
    assert WGANEngine.train_iteration.__doc__ and "WGAN.py:296-318" in WGANEngine.train_iteration.__doc__


def _dp_worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from imagegeneration_b200.dp import allreduce_sum_
    torch.manual_seed(0)
    w = torch.randn(64)
    rms = torch.zeros(64)
    g_all = torch.randn(world, 64, generator=torch.Generator().manual_seed(1))
    g_local = g_all[rank].clone()
    flat = g_local.clone()
    allreduce_sum_(flat, None, False)
    gs = 1.0 / world
    W.rmsprop_step(w, flat * gs, rms)                      # what b2g_rmsprop's gscale does
    # tail range only (the critic's gamma/beta under gan_model's optimizer)
    tail = g_local[48:].clone()
    allreduce_sum_(tail, None, False)
    out[rank] = (w.clone(), tail.clone())
    dist.destroy_process_group()


def test_data_parallel_update_equals_global_batch_update():
    world, port = 2, 29541
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_dp_worker, args=(world, port, out), nprocs=world, join=True)
    torch.manual_seed(0)
    w = torch.randn(64)
    rms = torch.zeros(64)
    g_all = torch.randn(world, 64, generator=torch.Generator().manual_seed(1))
    W.rmsprop_step(w, g_all.mean(0), rms)
    for r in range(world):
        assert torch.allclose(out[r][0], w, atol=1e-7)
        assert torch.allclose(out[r][1], g_all[:, 48:].sum(0), atol=1e-6)
