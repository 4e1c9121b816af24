"""Aggregate `ncu --page source --csv --print-source cuda,sass` output by source line.
Usage: python profiles/top_lines.py <csv> [topN]"""
import csv
import sys
from collections import defaultdict

path = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 50
cur = None
agg = defaultdict(lambda: [0, 0, ""])
for row in csv.reader(open(path)):
    if not row:
        continue
    if row[0] == "File Path":
        cur = row[1].split("/")[-1]
        continue
    if row[0] in ("Function Name", "Line No") or cur is None:
        continue
    if row[0].isdigit():
        try:
            samples, inst = int(row[6]), int(row[7])
        except ValueError:
            continue
        a = agg[(cur, int(row[0]))]
        a[0] += samples; a[1] += inst; a[2] = row[1]
tot_s = sum(a[0] for a in agg.values()); tot_i = sum(a[1] for a in agg.values())
print(f"total samples {tot_s}  total warp-instructions {tot_i}")
byfile = defaultdict(lambda: [0, 0])
for (f, l), a in agg.items():
    byfile[f][0] += a[0]; byfile[f][1] += a[1]
for f, v in sorted(byfile.items(), key=lambda kv: -kv[1][0]):
    print(f"  {f:24s} samples {100 * v[0] / tot_s:5.1f}%  inst {100 * v[1] / tot_i:5.1f}%")
print("top lines by stall samples:")
for (f, l), a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"{100 * a[0] / tot_s:5.2f}% s {100 * a[1] / tot_i:5.2f}% i  {f}:{l}  {a[2][:110]}")
