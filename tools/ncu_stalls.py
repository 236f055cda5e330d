"""Top stall-sample instructions of a kernel in an .ncu-rep captured with --import-source on (run here, no GPU):
   python tools/ncu_stalls.py file.ncu-rep [top_n]"""
import csv
import io
import subprocess
import sys


def main(path, top=40):
    out = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    h = rows[hi]
    si, ns = h.index("Source"), h.index("# Samples")
    ie = h.index("Instructions Executed")
    stall = [i for i, k in enumerate(h) if k.startswith("stall_") and "Not Issued" not in k]
    tot, recs = 0, []
    for r in rows[hi + 1:]:
        try:
            v = int(r[ns])
        except (ValueError, IndexError):
            continue
        tot += v
        why = sorted(((int(r[i] or 0), h[i][6:]) for i in stall), reverse=True)[:2]
        recs.append((v, r[0], r[si][:100], r[ie], why))
    print("total samples", tot)
    for v, addr, src, n, why in sorted(recs, reverse=True)[:top]:
        print(f"{v:6d} {100.0 * v / tot:5.1f}% {addr[-5:]} x{n:>8s} {src:100s} {why}")


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 40)
