"""Turn ncu output brought back in gpurun_out/ into the tracked summaries under profiles/.

  python tools/make_profile_summary.py launches <ncu --csv launch list> <out.csv> "<command that was profiled>"
  python tools/make_profile_summary.py full <ncu -i rep --page raw --csv> <out.csv> "<command>" [conv_traffic.json]

`launches`: per kernel name -> launches, average microseconds, share of the summed kernel time.
`full`:     one row per captured launch with the metrics DESIGN.md quotes; with a fifth argument also writes the
            conv kernel's per-launch DRAM traffic (bench.py's roofline.traffic).
"""
import csv
import json
import sys
from collections import OrderedDict

mode, src, dst, cmd = sys.argv[1:5]

if mode == "launches":
    with open(src) as f:
        lines = [l for l in f if not l.startswith("==")]
    per = OrderedDict()
    for row in csv.DictReader(lines):
        if row["Metric Name"] != "gpu__time_duration.sum":
            continue
        name = row["Kernel Name"].split("(")[0].replace("gnd::", "")
        if name.startswith("void at::") or "at::native" in name:
            continue
        per.setdefault(name, []).append(float(row["Metric Value"].replace(",", "")) / 1e3)
    tot = sum(sum(v) for v in per.values())
    with open(dst, "w") as f:
        f.write("# ncu launch list: %s\n" % cmd)
        f.write("# per-launch times are cold-cache and serialised (each kernel alone at boost clocks): compare SHARES, not absolutes\n")
        f.write("kernel,launches,avg_us,share_pct\n")
        for k, v in per.items():
            f.write("%s,%d,%.1f,%.1f\n" % (k, len(v), sum(v) / len(v), 100 * sum(v) / tot))
    print(open(dst).read())
else:
    rows = list(csv.reader(open(src)))
    hdr, units = rows[0], rows[1]
    ix = {h: i for i, h in enumerate(hdr)}
    cols = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
            "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
            "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
            "launch__registers_per_thread", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
            "lts__throughput.avg.pct_of_peak_sustained_elapsed", "launch__grid_size", "launch__block_size"]
    cols = [c for c in cols if c in ix]
    conv_bytes = []
    with open(dst, "w") as f:
        f.write("# ncu --set full --clock-control none --import-source on: %s\n" % cmd)
        f.write("# units: " + "; ".join("%s [%s]" % (c, units[ix[c]]) for c in cols[:3]) + "; percentages as named\n")
        f.write("Kernel Name," + ",".join(cols) + "\n")
        for r in rows[2:]:
            name = r[ix["Kernel Name"]].split("(")[0].replace("gnd::", "")
            f.write(name + "," + ",".join(r[ix[c]].replace(",", "") for c in cols) + "\n")
            if "conv3x3" in name:
                scale = {"Mbyte": 1e6, "Kbyte": 1e3, "Gbyte": 1e9, "byte": 1.0}
                rd = float(r[ix["dram__bytes_read.sum"]].replace(",", "")) * scale[units[ix["dram__bytes_read.sum"]]]
                wr = float(r[ix["dram__bytes_write.sum"]].replace(",", "")) * scale[units[ix["dram__bytes_write.sum"]]]
                conv_bytes.append(rd + wr)
    print(open(dst).read())
    if len(sys.argv) > 5 and conv_bytes:
        with open(sys.argv[5], "w") as f:
            json.dump({"bytes_per_launch": sum(conv_bytes) / len(conv_bytes), "launches": len(conv_bytes),
                       "per_launch_bytes": conv_bytes, "source": dst + " (" + cmd + ")"}, f, indent=1)
        print(open(sys.argv[5]).read())
