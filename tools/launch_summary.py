"""Aggregates an `ncu --metrics gpu__time_duration.sum --csv` launch list into a per-kernel share table (markdown)."""
import collections
import csv
import sys


def summarise(path):
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    agg = collections.defaultdict(lambda: [0, 0.0])
    for row in csv.DictReader(lines):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(row["Metric Value"].replace(",", ""))
        v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(row["Metric Unit"], 1.0)
        k = row["Kernel Name"].split("(")[0][:80]
        agg[k][0] += 1
        agg[k][1] += v
    tot = sum(v[1] for v in agg.values())
    out = ["| share | total us | launches | avg us | kernel |", "|---|---|---|---|---|"]
    for k, (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        out.append(f"| {100 * us / tot:.2f}% | {us:.1f} | {n} | {us / n:.1f} | `{k}` |")
    out.append(f"\nTotal {tot:.1f} us over {sum(v[0] for v in agg.values())} launches.")
    return "\n".join(out)


if __name__ == "__main__":
    print(summarise(sys.argv[1]))
