"""Static SASS footprint per source line, counting only instructions that executed (touched).
Usage: python profiles/static_lines.py <cuda,sass csv> [topN]"""
import csv, sys
from collections import defaultdict
path = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
cur_file = cur_line = None; src = {}
agg = defaultdict(lambda: [0, 0, 0])
for row in csv.reader(open(path)):
    if not row: continue
    if row[0] == "File Path": cur_file = row[1].split("/")[-1]; continue
    if row[0] in ("Function Name", "Line No"): continue
    if row[0].isdigit(): cur_line = int(row[0]); src[(cur_file, cur_line)] = row[1]; continue
    if row[0] == "" and len(row) > 7 and row[2].startswith("0x"):
        try: ex = int(row[7])
        except ValueError: continue
        a = agg[(cur_file, cur_line)]; a[0] += 1; a[1] += 1 if ex > 0 else 0; a[2] += ex
tt = sum(a[1] for a in agg.values())
print("touched static instructions:", tt, " total static:", sum(a[0] for a in agg.values()))
byf = defaultdict(int)
for (f, l), a in agg.items(): byf[f] += a[1]
print(dict(byf))
for (f, l), a in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
    print(f"{a[1]:5d} touched / {a[0]:5d} static  x{a[2] // max(a[1], 1):8d}  {f}:{l}  {src[(f, l)][:100]}")
