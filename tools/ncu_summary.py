"""Condense the ncu outputs of tools/gpu_profiles.sh (gpurun_out/<tag>_*) into what profiles/ keeps:

  profiles/<tag>_{trk,acq,trkc4}_pipes.csv   the pipe / issue / DRAM metric rows as ncu wrote them (one kernel launch each)
  profiles/<tag>_{trk,acq}_details.csv       the `--set full` details page
  profiles/<tag>_bench_launches_summary.csv  the launch list of the bench command folded per kernel (launches, total / mean us, share)
  profiles/<tag>_summary.json                the figures DESIGN.md quotes, by kernel
  profiles/ncu_traffic.json                  dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant kernels (bench.py's roofline.traffic)

usage: python tools/ncu_summary.py <tag>
"""
import csv
import io
import json
import os
import re
import shutil
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "gpurun_out")
PROF = os.path.join(ROOT, "profiles")


def ncu_csv(path):
    """rows of an ncu --csv log (the log starts with ==PROF== lines)"""
    txt = open(path, errors="replace").read()
    start = txt.find('"ID"')
    if start < 0:
        return []
    return list(csv.DictReader(io.StringIO(txt[start:])))


def to_num(s):
    try:
        return float(s.replace(",", ""))
    except Exception:
        return None


def short(name):
    base = re.split(r"[<(]", name.replace("void ", "", 1).strip(), 1)[0]
    return (base.split("::")[-1] or name)[:48]


def pipes(path):
    rows = ncu_csv(path)
    out = {}
    for r in rows:
        v = to_num(r.get("Metric Value", ""))
        if v is None:
            continue
        unit = r.get("Metric Unit", "")
        name = r["Metric Name"]
        # ncu scales byte and time units: bring them back to bytes / ns
        scale = {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1.0, "us": 1e3, "ms": 1e6, "ns": 1.0, "usecond": 1e3, "msecond": 1e6, "nsecond": 1.0}.get(unit, 1.0)
        out[name] = v * scale
        out["kernel"] = r.get("Kernel Name", "")
    return out


def launches(path):
    rows = ncu_csv(path)
    agg = {}
    for r in rows:
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = to_num(r["Metric Value"])
        unit = r.get("Metric Unit", "ns")
        ns = v * {"us": 1e3, "ms": 1e6, "ns": 1.0, "usecond": 1e3, "msecond": 1e6, "nsecond": 1.0, "second": 1e9}.get(unit, 1.0)
        k = short(r["Kernel Name"])
        a = agg.setdefault(k, [0, 0.0])
        a[0] += 1
        a[1] += ns
    return agg


def main():
    tag = sys.argv[1]
    summary = {}
    traffic = {}
    for what, key in (("trk", "trk_run_kernel"), ("acq", "acq_inv_kernel"), ("trkc4", "trk_correlate_kernel")):
        src = os.path.join(OUT, f"{tag}_{what}_pipes.csv")
        if not os.path.exists(src):
            continue
        shutil.copy(src, os.path.join(PROF, f"{tag}_{what}_pipes.csv"))
        p = pipes(src)
        if not p:
            continue
        summary[key] = p
        if "dram__bytes_read.sum" in p and what != "trkc4":
            traffic[key] = int(p["dram__bytes_read.sum"] + p.get("dram__bytes_write.sum", 0.0))
    for what in ("trk", "acq"):
        src = os.path.join(OUT, f"{tag}_{what}_details.csv")
        if os.path.exists(src) and os.path.getsize(src) > 0:
            shutil.copy(src, os.path.join(PROF, f"{tag}_{what}_details.csv"))
    src = os.path.join(OUT, f"{tag}_bench_launches.csv")
    if os.path.exists(src):
        agg = launches(src)
        total = sum(a[1] for a in agg.values()) or 1.0
        with open(os.path.join(PROF, f"{tag}_bench_launches_summary.csv"), "w") as f:
            f.write("kernel,launches,total_us,mean_us,share_of_all_kernel_time\n")
            for k, (n, ns) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
                f.write(f"{k},{n},{ns / 1e3:.1f},{ns / 1e3 / n:.2f},{ns / total:.4f}\n")
        summary["bench_launches"] = {k: {"launches": n, "mean_us": ns / 1e3 / n} for k, (n, ns) in agg.items()}
    for f in (f"{tag}_bench_quick.json", f"{tag}_bench_full.json", f"{tag}_bench_ref.json", f"{tag}_prof.log"):
        src = os.path.join(OUT, f)
        if os.path.exists(src) and os.path.getsize(src) > 0:
            shutil.copy(src, os.path.join(PROF, f))
    json.dump(summary, open(os.path.join(PROF, f"{tag}_summary.json"), "w"), indent=1, sort_keys=True)
    if traffic:
        traffic["_source"] = f"profiles/{tag}_*_pipes.csv (ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum, one launch, --clock-control none)"
        json.dump(traffic, open(os.path.join(PROF, "ncu_traffic.json"), "w"), indent=1, sort_keys=True)
    print(json.dumps({k: (v if k == "bench_launches" else {m: v[m] for m in v if "pct" in m or "dram" in m or m == "gpu__time_duration.sum"}) for k, v in summary.items()}, indent=1)[:4000])


if __name__ == "__main__":
    main()
