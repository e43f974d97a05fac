"""Print the step time and the in-situ kernel table of bench JSON lines: python tools/show_bench.py FILE..."""
import json
import sys

for f in sys.argv[1:]:
    for l in open(f):
        l = l.strip()
        if not l.startswith("{"):
            continue
        d = json.loads(l)
        rs = d.get("roofline_step", {})
        print(f, d["config"]["workload"], "ms/step", round(d["ms_per_step"], 4), "launches", d.get("gpu_launches"), "frac_hbm",
              round(rs.get("frac_hbm", 0), 4), "frac", round(rs.get("frac", 0), 4), "clk", d.get("clocks", {}).get("sm_mhz"))
        ks = d.get("kernels")
        if ks:
            print("  launches", ks["launches"], "sum_ms", round(ks["sum_ms"], 3))
            for k in ks["top"]:
                print("   ", k["name"], k["shape"], "n", k["n"], "avg_us", round(k["avg_us"], 1), "share", k["share"])
