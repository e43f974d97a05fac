"""Executed-instruction mix per opcode of an ncu report's SASS page: python tools/ncu_mix.py REPORT.ncu-rep [kernel-index]"""
import collections
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
which = int(sys.argv[2]) if len(sys.argv) > 2 else 0
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
blocks, cur = [], None
for r in rows:
    if r and r[0] == "Kernel Name":
        cur = {"name": r[1], "hdr": None, "data": []}
        blocks.append(cur)
    elif cur is not None and cur["hdr"] is None:
        cur["hdr"] = r
    elif cur is not None:
        cur["data"].append(r)
b = blocks[which]
hdr = b["hdr"]
si, src, ie = hdr.index("# Samples"), hdr.index("Source"), hdr.index("Instructions Executed")
mix, smp = collections.Counter(), collections.Counter()
for r in b["data"]:
    toks = r[src].strip().split()
    op = toks[1] if toks and toks[0].startswith("@") else (toks[0] if toks else "?")
    op = op.split(".")[0]
    mix[op] += int(r[ie])
    smp[op] += int(r[si])
tot, ts = sum(mix.values()), sum(smp.values())
print(b["name"][:100], "executed", tot)
for op, n in mix.most_common(25):
    print(f"{op:12s} {100.0 * n / tot:5.1f}% of instructions  {100.0 * smp[op] / max(ts, 1):5.1f}% of samples")
