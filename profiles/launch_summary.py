"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list per (kernel, grid, block).
Usage: python profiles/launch_summary.py <launches.csv> "<command that was profiled>" > summary.txt"""
import csv
import re
import sys
from collections import defaultdict

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 14 and r[0].isdigit()]
cmd = sys.argv[2] if len(sys.argv) > 2 else "?"
agg = defaultdict(lambda: [0, 0.0])
for r in rows:
    name = re.sub(r"\(.*", "", r[4])[:50]
    a = agg[(name, r[8], r[7])]
    a[0] += 1; a[1] += float(r[14]) * 1e-6
tot = sum(a[1] for a in agg.values())
print(f"# ncu launch list of `{cmd}` (gpu__time_duration.sum, serialised, cold cache)")
print("# columns: launches, kernel, grid, block, total ms, share of all GPU time in the process (calibration included)\n")
for (name, grid, block), (n, ms) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:24]:
    print(f"{n:4d} x {name:50s} grid {grid:16s} block {block:14s} {ms:10.3f} ms  {100 * ms / tot:4.1f}%")
big = {k: v for k, v in agg.items() if k[0].startswith("void k_") and int(re.findall(r"\d+", k[1])[0]) >= 16}
bt = sum(v[1] for v in big.values())
print("\n# kernels of the env step itself (full-size launches), share of the step:")
for (name, grid, block), (n, ms) in sorted(big.items(), key=lambda kv: -kv[1][1]):
    print(f"{name:50s} {grid:16s} {n:4d} launches {ms:10.3f} ms  {100 * ms / bt:5.2f}%")
