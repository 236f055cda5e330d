"""Data-parallel WGAN check, launched as
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29533 tools/wgan_dp_check.py
Every rank feeds its own shard; gradients are summed over ranks (NCCL) and scaled by 1/N inside b2g_rmsprop; BatchNorm
statistics stay per GPU (north_star item 4).  Checks that the replicas' variables stay bit-identical, and times
WGAN.train iterations (weak scaling: 64 images per GPU)."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from imagegeneration_b200.wgan import WGANEngine

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
dist.init_process_group("nccl")
B, IMAGE = 64, (144, 256, 3)
eng = WGANEngine(IMAGE, B, 5, seed=62 + rank, process_group=dist.group.WORLD)   # different seeds: broadcast must align them
g = torch.Generator(device="cuda").manual_seed(100 + rank)
reals = [torch.rand((B, *IMAGE), generator=g, device="cuda") * 2 - 1 for _ in range(5)]
zs = [torch.randn((B, 128), generator=g, device="cuda") for _ in range(6)]


def max_replica_diff(t):
    lst = [torch.empty_like(t) for _ in range(world)]
    dist.all_gather(lst, t.contiguous())
    return max(float((x - lst[0]).abs().max()) for x in lst)


out = eng.train_iteration(reals, zs[:5], zs[5])
diffs = {"critic_w": max_replica_diff(eng.critic_model.arena.w), "gen_w": max_replica_diff(eng.generator_model.arena.w),
         "gan_rms_cbn": max_replica_diff(eng.gan_model.rms_cbn)}
mm_diff = max_replica_diff(eng.critic_model.bn.moving_mean[0])     # per-GPU statistics: expected to differ
for _ in range(2):
    eng.train_iteration(reals, zs[:5], zs[5])
torch.cuda.synchronize()
dist.barrier()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
steps = 4
a.record()
for _ in range(steps):
    out = eng.train_iteration(reals, zs[:5], zs[5])
b.record()
torch.cuda.synchronize()
ms = torch.tensor([a.elapsed_time(b) / steps], device="cuda")
dist.all_reduce(ms, op=dist.ReduceOp.MAX)
if rank == 0:
    print(json.dumps({"check": "wgan data parallel", "n_gpus": world, "replica_max_abs_diff": diffs,
                      "moving_mean_diff_between_gpus": mm_diff, "ms_per_iteration": float(ms),
                      "real_images_per_s": world * 5 * B / (float(ms) * 1e-3),
                      "losses": [float(x) for x in out]}), flush=True)
    assert all(v == 0.0 for v in diffs.values()), diffs
dist.barrier()
dist.destroy_process_group()
