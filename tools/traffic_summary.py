"""profiles/r2_conv_traffic.json from an ncu launch list of tools/step_profile.py:
   ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none \
       -k regex:"conv_|wgrad_" -c 700 --csv --log-file gpurun_out/r2_traffic.csv python tools/step_profile.py
The LAST 66 conv / weight-gradient launches (weight packing filtered out) are one train_step at batch 64."""
import collections
import csv
import json
import sys

path, build = sys.argv[1], (sys.argv[2] if len(sys.argv) > 2 else "?")
lines = [l for l in open(path) if not l.startswith("==")]
per = collections.OrderedDict()
for row in csv.DictReader(lines):
    if "pack_" in row["Kernel Name"]:
        continue
    d = per.setdefault(row["ID"], {"name": row["Kernel Name"].split("(")[0]})
    v = float(row["Metric Value"].replace(",", ""))
    u = row["Metric Unit"]
    v *= {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1, "ms": 1e3}.get(u, 1)
    d[row["Metric Name"]] = v
launches = list(per.values())[-66:]
fam = collections.defaultdict(lambda: [0, 0.0, 0.0])
for d in launches:
    f = fam["conv_wgrad" if "wgrad" in d["name"] else "conv_igemm"]
    f[0] += 1
    f[1] += d.get("dram__bytes_read.sum", 0) + d.get("dram__bytes_write.sum", 0)
    f[2] += d.get("gpu__time_duration.sum", 0)
out = {k: {"launches": n, "dram_bytes_per_launch": b / n, "ncu_us_per_launch": t / n} for k, (n, b, t) in fam.items()}
out["source"] = ("ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k "
                 'regex:"conv_|wgrad_" python tools/step_profile.py (last 66 conv / weight-gradient launches = one train_step at '
                 f"B=64, build {build})")
json.dump(out, open("profiles/r2_conv_traffic.json", "w"), indent=1)
print(json.dumps(out, indent=1))
