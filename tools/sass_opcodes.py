"""Per-kernel histogram of the Blackwell tensor-core / TMA opcodes in libb200gan.so (run here, no GPU):
   python tools/sass_opcodes.py > profiles/r2_sass_opcodes.md
UTCHMMA = tcgen05.mma, UTCBAR = tcgen05.commit, LDTM = tcgen05.ld, UTMALDG / UTMASTG / UTMAPF = TMA load / store / prefetch."""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "imagegeneration_b200", "libb200gan.so")
OPS = ["UTCHMMA", "UTCQMMA", "UTCBAR", "LDTM", "UTMALDG", "UTMASTG", "UTMAPF", "UTCATOMSWS", "SYNCS", "HMMA", "FFMA"]
out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
demangle = lambda s: subprocess.run(["c++filt", s], capture_output=True, text=True).stdout.strip()
cur, hist = None, collections.OrderedDict()
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1)
        hist[cur] = collections.Counter()
        continue
    if cur is None:
        continue
    m = re.search(r"^\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
    if m:
        op = m.group(1)
        hist[cur]["_total"] += 1
        for o in OPS:
            if op.startswith(o):
                hist[cur][o] += 1
print("# SASS opcode histogram of libb200gan.so (sm_100a)\n")
print("`cuobjdump -sass imagegeneration_b200/libb200gan.so`, counted by tools/sass_opcodes.py. UTCHMMA = tcgen05.mma, UTCBAR = "
      "tcgen05.commit, LDTM = tcgen05.ld, UTMALDG / UTMASTG / UTMAPF = TMA tensor load / store / prefetch, SYNCS = mbarrier ops.\n")
print("| kernel | instr | " + " | ".join(OPS) + " |")
print("|---|---|" + "---|" * len(OPS))
for k, c in hist.items():
    name = demangle(k)
    name = re.sub(r"\(.*", "", name).replace("b2g::", "")[:90]
    print(f"| `{name}` | {c['_total']} | " + " | ".join(str(c[o]) if c[o] else "" for o in OPS) + " |")
