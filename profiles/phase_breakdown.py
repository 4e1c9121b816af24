"""Attribute every SASS instruction of k_step (ncu source page, cuda,sass view) to a phase of the
substep. Phases are line ranges of gmj_physics.cuh / gmj_tail.cuh / gmj_api.cu; instructions that
come from inlined helpers (gmj_device.cuh, intrinsics, gmj_collide.cuh) inherit the phase of the
nearest preceding phase-bearing instruction in address order.
Usage: python profiles/phase_breakdown.py <csv>"""
import csv
import re
import sys
from collections import defaultdict

import os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def phase_table():
    src = open(f"{ROOT}/gripper_mujoco_b200/csrc/gmj_physics.cuh").read().split("\n")
    marks = []
    infn = None
    for i, line in enumerate(src, 1):
        m = re.match(r"\s*// -{10,} (.*)", line)
        if m and infn == "physics_substep":
            marks.append((i, m.group(1).strip()))
        m2 = re.match(r"__device__ .*? (\w+)\(", line)
        if m2:
            infn = m2.group(1)
            if infn != "physics_substep":
                marks.append((i - 1, "fn:" + infn))
    return marks


marks = phase_table()


def phase_of(fname, line):
    if fname == "gmj_physics.cuh":
        cur = "physics:pre"
        for ln, nm in marks:
            if line >= ln:
                cur = nm
        return cur
    if fname == "gmj_primal.cuh":
        return "primal solve (gmj_primal.cuh)"
    if fname == "gmj_tail.cuh":
        return "tail (lane 0: stepper, sensors, events, obs)"
    if fname == "gmj_api.cu":
        return "k_step shell (state load/store)"
    return None


rows = []          # (addr, file, line, samples, inst)
cur_file = cur_line = None
for row in csv.reader(open(sys.argv[1])):
    if not row:
        continue
    if row[0] == "File Path":
        cur_file = row[1].split("/")[-1]; continue
    if row[0] in ("Function Name", "Line No"):
        continue
    if row[0].isdigit():
        cur_line = int(row[0]); continue
    if row[0] == "" and len(row) > 7 and row[2].startswith("0x"):
        try:
            rows.append((int(row[2], 16), cur_file, cur_line, int(row[6]), int(row[7])))
        except ValueError:
            pass
rows.sort()
agg = defaultdict(lambda: [0, 0, 0, 0])
last = "k_step shell (state load/store)"
for addr, f, l, s, i in rows:
    p = phase_of(f, l)
    if p is None:
        p = last
    else:
        last = p
    agg[p][0] += s; agg[p][1] += i; agg[p][2] += 1; agg[p][3] += (1 if i > 0 else 0)
ts = sum(v[0] for v in agg.values()); ti = sum(v[1] for v in agg.values())
print(f"total stall samples {ts}, warp instructions {ti}")
for p, v in sorted(agg.items(), key=lambda kv: -kv[1][0]):
    print(f"{100 * v[0] / ts:6.2f}% time  {100 * v[1] / ti:6.2f}% inst  static {v[2]:6d} touched {v[3]:6d}  {p}")
