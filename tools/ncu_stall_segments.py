#!/usr/bin/env python
"""Fold the warp-stall samples of one kernel of an ncu report per barrier-delimited code segment.

    ncu -i <report>.ncu-rep --page source --csv --kernel-name regex:<kernel> > src.csv
    python tools/ncu_stall_segments.py src.csv > profiles/<round>_<kernel>_stall_segments.txt

A "segment" is the SASS between two BAR / EXIT instructions (in address order): for the shared-memory FFT kernels that is one pass.  Per segment:
share of all stall samples, executed warp-instructions, the most executed opcodes and the leading stall reasons.
"""
import collections
import csv
import sys


def main(path):
    rows = list(csv.reader(open(path)))
    hdr = rows[1]
    ix = {}
    for i, h in enumerate(hdr):
        ix.setdefault(h, i)
    stall = {h: i for h, i in ix.items() if h.startswith("stall_")}
    data = rows[2:]
    tot = sum(int(r[ix["# Samples"]]) for r in data)
    print(rows[0][1] if len(rows[0]) > 1 else rows[0][0])
    print(f"stall samples {tot}, SASS instructions {len(data)}")
    seg, segs = 0, collections.defaultdict(lambda: [0, 0, collections.Counter(), collections.Counter()])
    for r in data:
        src = r[ix["Source"]].strip()
        op = src.split()[1] if src.startswith("@") else src.split()[0]
        n, ex = int(r[ix["# Samples"]]), int(r[ix["Instructions Executed"]])
        s = segs[seg]
        s[0] += n
        s[1] += ex
        s[2][op.split(".")[0]] += ex
        for h, i in stall.items():
            try:
                s[3][h.replace("stall_", "")] += int(r[i])
            except ValueError:
                pass
        if op.startswith("BAR") or op.startswith("EXIT"):
            seg += 1
    for k, v in segs.items():
        if v[0] < tot * 0.01:
            continue
        print(f"segment {k}: {100 * v[0] / tot:.1f} % of the samples, {v[1]} warp-instructions")
        print("   opcodes:", ", ".join(f"{o} {c}" for o, c in v[2].most_common(10)))
        print("   stalls :", ", ".join(f"{o} {c}" for o, c in v[3].most_common(6) if "Not Issued" not in o))


if __name__ == "__main__":
    main(sys.argv[1])
