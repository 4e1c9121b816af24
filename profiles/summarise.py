"""Turn an .ncu-rep of k_step into the committed text summary + traffic json.
Usage: python profiles/summarise.py gpurun_out/<name>.ncu-rep profiles/<out-prefix>"""
import csv
import json
import os
import subprocess
import sys

rep, out = sys.argv[1], sys.argv[2]
HERE = os.path.dirname(os.path.abspath(__file__))
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units, vals = rows[0], rows[1], rows[2]
get = {h: (v, u) for h, u, v in zip(hdr, units, vals)}
keys = ["Kernel Name", "Grid Size", "Block Size", "gpu__time_duration.sum", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__inst_executed_op_local_ld.sum",
        "smsp__inst_executed_op_local_st.sum", "smsp__warps_eligible.avg.per_cycle_active", "smsp__warps_active.avg.per_cycle_active",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "launch__waves_per_multiprocessor"]
lines = [f"# ncu --set full summary of {os.path.basename(rep)} (one k_step launch, profiles/profile_step.py workload)", ""]
for k in keys:
    if k in get:
        lines.append(f"{k:70s} {get[k][0]} {get[k][1]}")


def to_bytes(v, u):
    m = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    return float(v) * m.get(u, 1)


if "dram__bytes_read.sum" in get:
    tr = to_bytes(*get["dram__bytes_read.sum"]) + to_bytes(*get["dram__bytes_write.sum"])
    json.dump({"dram_bytes_per_launch": tr, "source": os.path.basename(rep)}, open(out + "_traffic.json", "w"))
    lines.append(f"{'dram bytes per launch (read+write)':70s} {tr:.0f} byte")
sass = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(sass.splitlines()))
h = rows[1]
idx = {n: i for i, n in enumerate(h)}
stalls = [n for n in h if n.startswith("stall_") and "Not Issued" not in n]
tot = {n: 0 for n in stalls}
for r in rows[2:]:
    for n in stalls:
        try:
            tot[n] += int(r[idx[n]])
        except (ValueError, IndexError):
            pass
s = sum(tot.values()) or 1
lines += ["", f"# warp stall samples ({len(rows) - 2} SASS instructions in the kernel)"]
for n, v in sorted(tot.items(), key=lambda kv: -kv[1])[:10]:
    lines.append(f"{n:30s} {100 * v / s:6.2f} %")
cs = out + "_cs.tmp.csv"
open(cs, "w").write(subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                                   capture_output=True, text=True).stdout)
lines += ["", "# time / instruction share per phase of the substep (profiles/phase_breakdown.py)"]
lines += subprocess.run([sys.executable, os.path.join(HERE, "phase_breakdown.py"), cs], capture_output=True, text=True).stdout.splitlines()
os.remove(cs)
open(out + "_summary.txt", "w").write("\n".join(lines) + "\n")
print("\n".join(lines[:40]))
