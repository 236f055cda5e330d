"""Data-parallel SNDCGAN check, launched as
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29551 tools/sndcgan_dp_check.py
Runs three train_steps twice from the same state — with the discriminator's early (overlapped) tail all-reduce and with
one all-reduce after the backward (B2G_DP_EARLY_ALLREDUCE=0) — and checks that the replicas stay bit-identical and that
both schedules give the same weights up to the run-to-run noise of the split-K weight-gradient reds (fp32 atomics:
two runs of the SAME schedule differ by ~5e-5 relative after three Adam steps)."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from imagegeneration_b200.sndcgan import SNDCGANEngine

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
dist.init_process_group("nccl")
B, IMAGE = 16, (144, 256, 3)
g = torch.Generator(device="cuda").manual_seed(100 + rank)
real = torch.rand((B, *IMAGE), generator=g, device="cuda") * 2 - 1
z = torch.rand((B, 128), generator=g, device="cuda") * 2 - 1
seeds = tuple([1000 * (p + 1) + i + 7 * rank for i in range(7)] for p in range(3))


def run(early: str):
    os.environ["B2G_DP_EARLY_ALLREDUCE"] = early
    eng = SNDCGANEngine(B, dropout=0.5, image_size=IMAGE, seed=62, process_group=dist.group.WORLD, cuda_graph=False)
    for _ in range(3):
        out = eng.train_step(real, z, dropout_seeds=seeds)
    torch.cuda.synchronize()
    return eng.disc_model.arena.w.clone(), eng.gen_model.arena.w.clone(), [float(x) for x in out[:4]]


def max_replica_diff(t):
    lst = [torch.empty_like(t) for _ in range(world)]
    dist.all_gather(lst, t.contiguous())
    return max(float((x - lst[0]).abs().max()) for x in lst)


dw1, gw1, l1 = run("1")
dw0, gw0, l0 = run("0")
res = {"check": "sndcgan data parallel, early tail all-reduce", "n_gpus": world,
       "replica_max_abs_diff": {"disc": max_replica_diff(dw1), "gen": max_replica_diff(gw1)},
       "early_vs_late_rel": {"disc": float((dw1 - dw0).norm() / dw0.norm()), "gen": float((gw1 - gw0).norm() / gw0.norm())},
       "losses_early": l1, "losses_late": l0}
if rank == 0:
    print(json.dumps(res), flush=True)
    assert res["replica_max_abs_diff"]["disc"] == 0.0 and res["replica_max_abs_diff"]["gen"] == 0.0, res
    assert res["early_vs_late_rel"]["disc"] <= 1e-3 and res["early_vs_late_rel"]["gen"] <= 1e-3, res
dist.barrier()
dist.destroy_process_group()
