"""Executed instructions and stall samples per SOURCE LINE of one kernel launch in an .ncu-rep captured with
--import-source on (run here, no GPU):  python tools/ncu_lines.py file.ncu-rep <kernel regex> [launch index] [top_n]"""
import collections
import csv
import io
import subprocess
import sys


def main(path, kre, idx=0, top=40):
    cmd = ["ncu", "-i", path, "--page", "source", "--csv", "--print-source", "sass,cuda", "-k", f"regex:{kre}",
           "--launch-skip", str(idx), "-c", "1"]
    out = subprocess.run(cmd, capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    tot, samp = collections.Counter(), collections.Counter()
    cur, hdr, fn = None, None, ""
    for r in rows:
        if not r:
            continue
        if r[0] == "File Path":
            cur = r[1].split("/")[-1]
            continue
        if r[0] == "Function Name":
            fn = r[1]
            continue
        if r[0] == "Line No":
            hdr = r
            continue
        if hdr is None:
            continue
        try:
            line = int(r[0])
        except ValueError:
            continue
        d = dict(zip(hdr, r))
        try:
            ie, ns = int(d.get("Instructions Executed") or 0), int(d.get("# Samples") or 0)
        except ValueError:
            continue
        key = (cur, line, r[1].strip()[:100])
        tot[key] += ie
        samp[key] += ns
    T, S = sum(tot.values()), sum(samp.values())
    print(fn[:120])
    print("instructions", T, "samples", S)
    for k, v in sorted(samp.items(), key=lambda kv: -kv[1])[:top]:
        print(f"smp {100 * v / max(S, 1):5.1f}%  inst {100 * tot[k] / max(T, 1):5.1f}%  {k[0]}:{k[1]}  {k[2]}")


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2], int(sys.argv[3]) if len(sys.argv) > 3 else 0, int(sys.argv[4]) if len(sys.argv) > 4 else 40)
