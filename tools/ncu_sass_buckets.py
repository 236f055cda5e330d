"""Stall samples / executed instructions of one kernel launch in an .ncu-rep (captured with --import-source on), bucketed by
SASS address range (the warp roles of the warp-specialised kernels are contiguous ranges) plus the top instructions:
   python tools/ncu_sass_buckets.py file.ncu-rep <kernel regex> [launch index] [top_n]     (run here, no GPU)"""
import collections
import csv
import io
import subprocess
import sys


def main(path, kre, idx=0, top=30):
    cmd = ["ncu", "-i", path, "--page", "source", "--csv", "-k", f"regex:{kre}", "--launch-skip", str(idx), "-c", "1"]
    out = subprocess.run(cmd, capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    h = rows[hi]
    si, ns, ie = h.index("Source"), h.index("# Samples"), h.index("Instructions Executed")
    stall = [i for i, k in enumerate(h) if k.startswith("stall_") and "Not Issued" not in k]
    recs = []
    for r in rows[hi + 1:]:
        try:
            v, n = int(r[ns]), int(r[ie])
        except (ValueError, IndexError):
            continue
        why = sorted(((int(r[i] or 0), h[i][6:]) for i in stall), reverse=True)[:2]
        recs.append((int(r[0], 16), r[si][:70], v, n, why))
    base = recs[0][0]
    tot, toti = sum(r[2] for r in recs), sum(r[3] for r in recs)
    print("samples", tot, "warp instructions", toti)
    b = collections.OrderedDict()
    for a, src, v, n, why in recs:
        d = b.setdefault((a - base) // 0x400, [0, 0, collections.Counter()])
        d[0] += v
        d[1] += n
        for c, w in why:
            d[2][w] += c
    for k, d in b.items():
        if d[0] > tot * 0.005 or d[1] > toti * 0.005:
            print(f"{k * 0x400:6x} smp {100 * d[0] / tot:5.1f}% inst {100 * d[1] / toti:5.1f}% {d[2].most_common(3)}")
    print()
    for a, src, v, n, why in sorted(recs, key=lambda r: -r[2])[:top]:
        print(f"{a - base:6x} smp {100 * v / tot:5.1f}% x{n:9d} {src:70s} {why}")


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2], int(sys.argv[3]) if len(sys.argv) > 3 else 0, int(sys.argv[4]) if len(sys.argv) > 4 else 30)
