"""Hot instructions of an ncu report's source page: python tools/ncu_hot.py REPORT.ncu-rep [N] [kernel-index]
(runs `ncu -i REPORT --page source --csv --print-source sass` and prints the N most sampled SASS lines in program order)."""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
n = int(sys.argv[2]) if len(sys.argv) > 2 else 40
which = int(sys.argv[3]) if len(sys.argv) > 3 else 0
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
# the page repeats (kernel name row, header row, data...) per kernel
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
hdr, data = b["hdr"], b["data"]
si, src, ie = hdr.index("# Samples"), hdr.index("Source"), hdr.index("Instructions Executed")
tot = sum(int(r[si]) for r in data)
print(b["name"][:100])
print("total samples", tot, "instructions", len(data), "executed", sum(int(r[ie]) for r in data))
top = sorted(range(len(data)), key=lambda i: -int(data[i][si]))[:n]
for i in sorted(top):
    r = data[i]
    print(f"{i:5d} {100.0 * int(r[si]) / tot:5.1f}% exec {r[ie]:>9s}  {r[src].strip()[:120]}")
