"""Times torch.distributed all_reduce (NCCL) of the two gradient arenas' sizes under the NCCL_* environment it was started
with: torchrun --nproc-per-node N tools/nccl_probe.py — prints one line on rank 0 (used to pick NCCL_ALGO / NCCL_PROTO for
the exposed 24.6 MB discriminator all-reduce, profiles/r2_dp_timeline.md)."""
import os
import torch
import torch.distributed as dist

local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
out = []
for mb in (24.6, 164.4):
    x = torch.ones(int(mb * 1e6 / 4), device="cuda")
    for _ in range(5):
        dist.all_reduce(x)
    torch.cuda.synchronize()
    dist.barrier()
    ts = []
    for _ in range(10):
        dist.barrier()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); dist.all_reduce(x); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    ts.sort()
    out.append(f"{mb} MB: median {ts[5]:.0f} us min {ts[0]:.0f} us")
if dist.get_rank() == 0:
    print({k: v for k, v in os.environ.items() if k.startswith("NCCL_") and k != "NCCL_DEBUG"}, "|", "; ".join(out), flush=True)
dist.destroy_process_group()
