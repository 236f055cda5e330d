"""Per-rank GPU timeline of data-parallel SNDCGAN steps (CUPTI through torch.profiler; nsys is not installed here).

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 \
      tools/dp_timeline.py [--batch 64] [--steps 4] [--graph 0|1]

For the LAST profiled step every rank writes gpurun_out/dp_timeline_n<N>_rank<r>.json: step span, busy time of our
kernels, every NCCL kernel (start offset, duration, what ran on the compute stream meanwhile) and the idle gaps of the
compute stream longer than 5 us — i.e. what fills the time a step loses when a process group exists."""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from torch.profiler import ProfilerActivity, profile

from imagegeneration_b200.sndcgan import SNDCGANEngine


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--steps", type=int, default=4)
    ap.add_argument("--graph", type=int, default=0)
    args = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    pg = None
    if world > 1:
        os.environ["B2G_CUDA_GRAPH_DP"] = str(args.graph)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        pg = dist.group.WORLD
    eng = SNDCGANEngine(args.batch, dropout=0.5, seed=62, process_group=pg, cuda_graph=bool(args.graph))
    g = torch.Generator().manual_seed(1234 + rank)
    real = (torch.rand((args.batch, 144, 256, 3), generator=g) * 2 - 1).cuda()
    z = (torch.rand((args.batch, 128), generator=g) * 2 - 1).cuda()
    for _ in range(4):
        eng.train_step(real, z)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    marks = []
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        for _ in range(args.steps):
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize()
            with torch.profiler.record_function("b2g_step"):
                eng.train_step(real, z)
                torch.cuda.synchronize()
    evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
    steps = [e for e in prof.events() if e.name == "b2g_step" and e.device_type == torch.autograd.DeviceType.CPU]
    if not steps or not evs:
        print("no CUDA events recorded (CUPTI unavailable?)")
        return
    t0, t1 = steps[-1].time_range.start, steps[-1].time_range.end
    ks = sorted([e for e in evs if e.time_range.start >= t0 and e.time_range.end <= t1 + 2000],
                key=lambda e: e.time_range.start)
    seen, uniq = set(), []
    for e in ks:                       # record_function ranges ("nccl:all_reduce") duplicate the kernels they wrap
        key = (e.time_range.start, e.time_range.end)
        if e.name.startswith("nccl:") or key in seen:
            continue
        seen.add(key)
        uniq.append(e)
    ks = uniq
    first = ks[0].time_range.start
    last = max(e.time_range.end for e in ks)
    nccl = [e for e in ks if "nccl" in e.name.lower()]
    ours = [e for e in ks if "nccl" not in e.name.lower()]
    out = {"world": world, "rank": rank, "batch": args.batch, "graph": args.graph,
           "step_span_us": last - first, "our_kernels": len(ours), "our_busy_us": sum(e.time_range.end - e.time_range.start for e in ours),
           "nccl": [], "gaps_over_5us": [],
           "kernels": [[e.name[:48], round(e.time_range.start - first, 1), round(e.time_range.end - e.time_range.start, 1)]
                       for e in ks]}
    for e in nccl:
        s, d = e.time_range.start - first, e.time_range.end - e.time_range.start
        overl = [o for o in ours if o.time_range.start < e.time_range.end and o.time_range.end > e.time_range.start]
        busy = sum(min(o.time_range.end, e.time_range.end) - max(o.time_range.start, e.time_range.start) for o in overl)
        out["nccl"].append({"name": e.name[:60], "start_us": s, "dur_us": d, "compute_busy_us_meanwhile": busy,
                            "kernels_meanwhile": len(overl)})
    prev_end, prev_name = None, ""
    for e in ours:
        if prev_end is not None and e.time_range.start - prev_end > 5:
            out["gaps_over_5us"].append({"at_us": prev_end - first, "gap_us": e.time_range.start - prev_end,
                                         "after": prev_name[:50], "before": e.name[:50]})
        if prev_end is None or e.time_range.end > prev_end:
            prev_end, prev_name = e.time_range.end, e.name
    os.makedirs("gpurun_out", exist_ok=True)
    with open(f"gpurun_out/dp_timeline_n{world}_g{args.graph}_rank{rank}.json", "w") as f:
        json.dump(out, f, indent=1)
    if rank == 0:
        print(json.dumps({k: v for k, v in out.items() if k not in ("nccl", "gaps_over_5us", "kernels")}))
        print("nccl:", json.dumps(out["nccl"]))
        print("gaps:", json.dumps(sorted(out["gaps_over_5us"], key=lambda g: -g["gap_us"])[:12]))
    eng.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
