"""Summarise an ncu --set full report of one kernel: headline counters, then the SASS grouped into regions of equal execution count
(instructions per warp-iteration, share of executed instructions, share of stall samples) and the hottest stall lines.
usage: python tools/ncu_segments.py report.ncu-rep [units_per_launch]"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
units = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
det = subprocess.run(["ncu", "-i", rep, "--page", "details", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(det)))
h = rows[0]
iN, iV, iU = h.index("Metric Name"), h.index("Metric Value"), h.index("Metric Unit")
want = {"Duration", "Executed Ipc Active", "Issue Slots Busy", "Registers Per Thread", "Achieved Active Warps Per SM", "Executed Instructions", "L1/TEX Hit Rate", "L2 Cache Throughput",
        "DRAM Throughput", "Eligible Warps Per Scheduler", "Active Warps Per Scheduler", "Dynamic Shared Memory Per Block", "Mem Pipes Busy", "SM Frequency"}
for r in rows[1:]:
    if r[iN] in want:
        print(f"{r[iN]:38s} {r[iV]} {r[iU]}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
h = rows[1]
iS, iE, iSm = h.index("Source"), h.index("Instructions Executed"), h.index("# Samples")
data = [(r[iS].strip(), int(r[iE]), int(r[iSm])) for r in rows[2:] if len(r) > iE and r[iE].isdigit()]
tot, tsm = sum(d[1] for d in data), sum(d[2] for d in data)
print(f"warp-instructions {tot}  ({tot / units:.1f} per unit)  stall samples {tsm}")
segs, cur = [], None
for i, (s, e, sm) in enumerate(data):
    key = e / units
    if cur is None or abs(key - cur[0]) > 0.06 * max(cur[0], 0.05):
        cur = [key, i, i, 0, 0]
        segs.append(cur)
    cur[2] = i
    cur[3] += e
    cur[4] += sm
for k, a, b, e, sm in segs:
    if e / tot > 0.004 or sm / tsm > 0.01:
        print(f"  lines {a:4d}-{b:4d} n={b - a + 1:4d} exec/unit={k:6.2f} instr={e / tot:.3f} samples={sm / tsm:.3f}  {data[a][0][:60]}")
print("hottest stall lines:")
for i in sorted(range(len(data)), key=lambda i: -data[i][2])[:12]:
    print(f"  {i:4d} {data[i][2]:5d} ({data[i][2] / tsm:.3f})  {data[i][0][:90]}")
