"""Summarise an ncu --csv launch list (gpurun_out/*.csv): per-launch time, DRAM bytes, instructions."""
import csv
import sys
from collections import OrderedDict

path = sys.argv[1]
first = int(sys.argv[2]) if len(sys.argv) > 2 else 0
with open(path) as f:
    lines = [l for l in f if not l.startswith("==")]
d = OrderedDict()
for row in csv.DictReader(lines):
    d.setdefault((int(row["ID"]), row["Kernel Name"][:38]), {})[row["Metric Name"]] = row["Metric Value"].replace(",", "")
tot = 0.0
for (i, name), v in d.items():
    if i < first or name.startswith("void at::"):
        continue
    us = float(v["gpu__time_duration.sum"]) / 1e3
    tot += us
    extra = ""
    if "dram__bytes_read.sum" in v:
        extra = " rd=%.0fMB wr=%.0fMB" % (float(v["dram__bytes_read.sum"]) / 1e6, float(v["dram__bytes_write.sum"]) / 1e6)
    if "smsp__inst_executed.sum" in v:
        extra += " inst=%.1fM warps=%s%%" % (float(v["smsp__inst_executed.sum"]) / 1e6, v.get("sm__warps_active.avg.pct_of_peak_sustained_active", "?"))
    print("%3d %-38s %8.1f us%s" % (i, name, us, extra))
print("total %.1f us" % tot)
