"""Attributes the step-time difference between two dp_timeline.py records (e.g. 1 GPU vs rank 0 of 2 GPUs) kernel by
kernel: python tools/dp_timeline_diff.py gpurun_out/dp_timeline_n1_g0_rank0.json gpurun_out/dp_timeline_n2_g0_rank0.json"""
import json
import sys

a, b = (json.load(open(p)) for p in sys.argv[1:3])
ka = [k for k in a["kernels"] if "nccl" not in k[0].lower()]
kb = [k for k in b["kernels"] if "nccl" not in k[0].lower()]
print(f"step span {a['step_span_us']:.0f} -> {b['step_span_us']:.0f} us ({b['step_span_us'] - a['step_span_us']:+.0f})")
print(f"compute kernels {len(ka)} vs {len(kb)}; busy {sum(k[2] for k in ka):.0f} -> {sum(k[2] for k in kb):.0f} us")
if len(ka) == len(kb):
    rows = sorted(((kb[i][2] - ka[i][2], i, ka[i][0], ka[i][2], kb[i][2]) for i in range(len(ka))), reverse=True)
    print("largest per-kernel slow-downs (us):")
    for d, i, name, x, y in rows[:12]:
        print(f"  #{i:3d} {name:48s} {x:8.1f} -> {y:8.1f}  {d:+7.1f}")
    print(f"sum of per-kernel differences {sum(r[0] for r in rows):+.0f} us")
nc = [k for k in b["kernels"] if "nccl" in k[0].lower()]
print("NCCL kernels (start us, duration us):", [(round(k[1]), round(k[2])) for k in nc])
ends, idle = 0.0, 0.0
for k in sorted(kb, key=lambda k: k[1]):
    if ends and k[1] > ends:
        idle += k[1] - ends
    ends = max(ends, k[1] + k[2])
print(f"compute-stream idle inside the step (second record): {idle:.0f} us")
