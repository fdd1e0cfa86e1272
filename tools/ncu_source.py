"""Hot spots of one kernel from `ncu -i rep --page source --csv --kernel-name regex:X` output:
instructions executed and stall samples per SASS instruction, largest first (plus a linear listing window)."""
import csv
import sys

path = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
rows = list(csv.reader(open(path)))
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
body = rows[2:]
tot_i = sum(int(r[ix["Instructions Executed"]] or 0) for r in body)
tot_s = sum(int(r[ix["Warp Stall Sampling (All Samples)"]] or 0) for r in body)
print("instructions %.1fM, stall samples %d, SASS lines %d" % (tot_i / 1e6, tot_s, len(body)))
mode = sys.argv[3] if len(sys.argv) > 3 else "samples"
key = "Warp Stall Sampling (All Samples)" if mode == "samples" else "Instructions Executed"
order = sorted(range(len(body)), key=lambda j: -int(body[j][ix[key]] or 0))[:top]
for j in sorted(order):
    r = body[j]
    print("%5d  inst %6.2f%%  samp %6.2f%%  %s" % (j, 100.0 * int(r[ix["Instructions Executed"]] or 0) / max(tot_i, 1),
                                                  100.0 * int(r[ix["Warp Stall Sampling (All Samples)"]] or 0) / max(tot_s, 1), r[ix["Source"]].strip()[:90]))
