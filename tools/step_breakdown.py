#!/usr/bin/env python
"""Joins an ncu launch list (gpu__time_duration.sum CSV) of `bench.py --steps 1 --no-graph` with the engine's shape
log (RPDE_TC_DBG=9 stderr of the same command) and prints the last step aggregated per (kernel, shape, epilogue)."""
import collections
import csv
import re
import sys


def main(csv_path, shapes_path, first_kernel="concat_grid_kernel"):
    rows = list(csv.DictReader([l for l in open(csv_path) if l.startswith('"')]))
    idx = [i for i, r in enumerate(rows) if first_kernel in r["Kernel Name"]]
    rows = rows[idx[-1]:]
    shapes = [l.strip() for l in open(shapes_path) if l.startswith("[rpde")]
    # drop the trailing roofline-probe GEMMs (same shape repeated after the step) by aligning from the step's start
    n_gemm = sum(1 for r in rows if "gemm" in r["Kernel Name"])
    shapes = shapes[-n_gemm:]
    # bench.py's roofline probe runs after the step: randn fills + the same plain GEMM repeated; cut it off
    probe = shapes[-1] if shapes else None
    while rows and (("gemm" in rows[-1]["Kernel Name"] and shapes and shapes[-1] == probe)
                    or "distribution_elementwise" in rows[-1]["Kernel Name"]):
        if "gemm" in rows[-1]["Kernel Name"]:
            shapes.pop()
        rows.pop()
    agg = collections.OrderedDict()
    gi = 0
    total = 0.0
    for r in rows:
        name = re.sub(r"\(.*", "", r["Kernel Name"]).replace("void ", "").replace("rpde::", "")
        us = float(r["Metric Value"].replace(",", "")) / 1e3
        key = name
        if "gemm" in name:
            s = shapes[gi]
            gi += 1
            m = re.search(r"M=(\d+) N=(\d+) K=(\d+) batch=(\d+)", s)
            fl = " ".join(f"{k}" for k in ("beta=1", "act=1", "dact=1", "dact=3", "Z=1", "red=1", "epi_tma=0") if k in s)
            sk = re.search(r"splitk=(\d+)", s)
            key = f"{name[:26]:26s} M={m.group(1)} N={m.group(2)} K={m.group(3)} b={m.group(4)} sk={sk.group(1) if sk else '-'} {fl}"
        a = agg.setdefault(key, [0, 0.0])
        a[0] += 1
        a[1] += us
        total += us
    print(f"last step: {len(rows)} launches, {total / 1e3:.2f} ms device time (ncu, serialised)")
    for k, (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        if us / total < 0.003:
            continue
        print(f"{us / total * 100:5.1f}%  n={n:3d}  avg={us / n:8.1f} us  {k}")


if __name__ == "__main__":
    main(*sys.argv[1:])
