#!/usr/bin/env python
"""Summarise an ncu launch list (gpu__time_duration.sum CSV): share, count, mean time per kernel (+ grid size buckets)."""
import collections
import csv
import re
import sys


def main(path, skip_frac=0.0, by_grid=False):
    lines = [l for l in open(path) if l.startswith('"')]
    rows = list(csv.DictReader(lines))
    rows = rows[int(len(rows) * skip_frac):]
    agg = collections.defaultdict(lambda: [0, 0.0])
    for row in rows:
        name = re.sub(r"\(.*", "", row["Kernel Name"])
        name = re.sub(r"^void ", "", name)
        if by_grid:
            name += " grid=" + row["Grid Size"]
        v = float(row["Metric Value"].replace(",", ""))
        agg[name][0] += 1
        agg[name][1] += v
    tot = sum(v[1] for v in agg.values())
    print(f"{len(rows)} launches, {tot / 1e6:.3f} ms total device time")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:40]:
        print(f"{v[1] / tot * 100:6.2f}%  n={v[0]:5d}  avg={v[1] / v[0] / 1e3:9.1f} us  {k[:110]}")


if __name__ == "__main__":
    main(sys.argv[1], float(sys.argv[2]) if len(sys.argv) > 2 else 0.0, len(sys.argv) > 3)
