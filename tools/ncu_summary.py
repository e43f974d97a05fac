"""Transpose `ncu -i X.ncu-rep --page raw --csv` into metric,unit,value[,value...] (one column per captured launch),
keeping the metrics that matter for the roofline reading.  usage: ncu_summary.py X.ncu-rep > profiles/<name>.csv"""
import csv, io, subprocess, sys
KEEP = ("Kernel Name", "Block Size", "Grid Size", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_bytes.sum", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__shared_mem_per_block",
        "launch__occupancy_limit", "smsp__inst_executed.sum", "sm__inst_executed_pipe_fma", "smsp__cycles_active.avg",
        "smsp__average_warp", "launch__waves_per_multiprocessor", "sm__cycles_elapsed.max")
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, units, data = rows[0], rows[1], rows[2:]
w = csv.writer(sys.stdout)
w.writerow(["metric", "unit"] + [f"launch{i}" for i in range(len(data))])
for c, name in enumerate(hdr):
    if any(k in name for k in KEEP):
        w.writerow([name, units[c]] + [r[c] for r in data])
