"""Per-launch summary of an `ncu --csv` metrics log: time, tensor pipe, DRAM bytes (usage: show_launches.py file.csv [first_id] [min_us])."""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
d = collections.OrderedDict()
for r in rows[hdr + 1:]:
    if len(r) < 15:
        continue
    k = (int(r[0]), r[4].split("(")[0][-30:])
    try:
        d.setdefault(k, {})[r[12]] = float(r[14].replace(",", ""))
    except ValueError:
        pass
lo = int(sys.argv[2]) if len(sys.argv) > 2 else 0
min_us = float(sys.argv[3]) if len(sys.argv) > 3 else 50.0
for (i, k), v in d.items():
    t = v.get("gpu__time_duration.sum", 0) / 1e3
    if t >= min_us and i >= lo:
        print(i, k, "%.1f us" % t, "tensor %.1f %%" % v.get("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", 0),
              "dram r/w %.2f / %.2f GB" % (v.get("dram__bytes_read.sum", 0) / 1e9, v.get("dram__bytes_write.sum", 0) / 1e9))
