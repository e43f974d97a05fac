"""SASS opcode census of librpde_b200.so per kernel: python tools/sass_census.py > profiles/<stage>_sass_census.txt
Counts the instructions that prove which hardware path a kernel uses (B200_PROFILING.md): tcgen05 (UTCHMMA, LDTM, STTM),
TMA (UTMALDG, UTMASTG, UTMAREDG, UTMAPF), the legacy warp-level tensor path (HMMA), cp.async (LDGSTS), packed fp32 (FFMA2)."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(ROOT, "resolution_pde_b200", "librpde_b200.so")
ops = ["UTCHMMA", "UTCQMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTMAREDG", "UTMAPF", "HMMA", "LDGSTS", "FFMA2", "FFMA", "MUFU"]
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
names = subprocess.run(["c++filt", "-p"], input="\n".join(re.findall(r"Function : (\S+)", out)), capture_output=True, text=True).stdout.split("\n")
counts, order, cur, k = {}, [], None, 0
for line in out.split("\n"):
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = names[k]
        k += 1
        counts[cur] = collections.Counter()
        order.append(cur)
        continue
    if cur is None:
        continue
    m = re.search(r"^\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
    if m:
        op = m.group(1)
        counts[cur]["total"] += 1
        for o in ops:
            if op == o:
                counts[cur][o] += 1
print("kernel".ljust(70), " ".join(o.rjust(8) for o in ["total"] + ops))
for kname in sorted(order):
    c = counts[kname]
    print(kname[:70].ljust(70), " ".join(str(c[o]).rjust(8) for o in ["total"] + ops))
