#!/usr/bin/env python
"""Summarise an .ncu-rep (read here, no GPU needed): one CSV row per profiled launch with the metrics the profiles/
summaries quote.   python tools/ncu_summary.py gpurun_out/x.ncu-rep > profiles/x_summary.csv"""
import csv
import subprocess
import sys

WANT = ["gpu__time_duration.sum", "sm__cycles_elapsed.avg.per_second", "launch__registers_per_thread",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__grid_size", "launch__block_size", "launch__shared_mem_per_block_dynamic"]
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr, units, data = rows[0], rows[1], rows[2:]
cols = [hdr.index("Kernel Name")] + [hdr.index(w) for w in WANT if w in hdr]
w = csv.writer(sys.stdout)
w.writerow([hdr[c] + (" [%s]" % units[c] if units[c] else "") for c in cols])
for r in data:
    w.writerow([r[c].split("(")[0] if c == cols[0] else r[c] for c in cols])
