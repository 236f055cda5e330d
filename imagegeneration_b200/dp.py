"""Data-parallel plumbing for the SNDCGAN step (north_star item 4): pure data parallel, exact.

Every rank holds a replica and B/N samples; each apply_gradients (SNDCGAN.py:249,:257,:267) is preceded by ONE
all-reduce (sum) of the network's flat gradient arena, and the fused Adam kernel multiplies by 1/N.  The
generator has no batch statistics in the reference's train_step (inference-mode BatchNorm) and D has no BN, so
no other collective exists on the path."""
from __future__ import annotations

import torch


def shard_batch(t: torch.Tensor, rank: int, world: int) -> torch.Tensor:
    """Rank r of N takes rows [r*B/N, (r+1)*B/N) of the global batch."""
    b = t.shape[0]
    if b % world:
        raise ValueError(f"global batch {b} is not divisible by world size {world}")
    per = b // world
    return t[rank * per:(rank + 1) * per]


def allreduce_sum_(flat: torch.Tensor, group=None, async_op: bool = False):
    """In-place sum all-reduce of a flat gradient arena (NCCL on GPUs, gloo in the CPU tests)."""
    import torch.distributed as dist
    if group is None and not dist.is_initialized():
        return None
    if dist.get_world_size(group) == 1:
        return None
    return dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group, async_op=async_op)


def broadcast_(flat: torch.Tensor, group=None, src: int = 0) -> None:
    import torch.distributed as dist
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.broadcast(flat, src=src, group=group)
