#!/usr/bin/env python
"""One acquisition configuration of bench.py on its own (tuning switches of the library are read from the environment):
    GB200_ACQ_CLUSTER=1 python tools/prof_acq_cfg.py C2
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402
import bench  # noqa: E402

which = sys.argv[1] if len(sys.argv) > 1 else "C2"
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda:0")
shapes = {
    "C1": ("C1 GPS L1 4 MS/s, fft 4000 x 80 bins x 32 PRN", 4e6, 4000, 4000, 32, 10000, 250, 1, 0, False),
    "C2": ("C2 Galileo E1 4 MS/s cshort, bit_transition, fft 32000 x 80 bins x 36 PRN", 4e6, 32000, 32000, 36, 5000, 125, 1, 1, True),
    "C4": ("C4 GPS L5 / Galileo E5a 20 MS/s, fft 20000 x 80 bins x 32 PRN x 20 dwells", 20e6, 20000, 20000, 32, 10000, 250, 20, 0, False),
}
if which.startswith("N:"):  # a plain cf32 sweep of that transform length: 32 PRN x 80 bins
    n_ = int(which[2:])
    shapes[which] = (f"fft {n_} x 80 bins x 32 PRN", n_ * 1000.0, n_, n_, 32, 10000, 250, 1, 0, False)
name, fs, n, consumed, prns, dmax, dstep, dwells, iq, bt = shapes[which]
r = bench.acq_config_bench(0, name, fs, n, consumed, prns, dmax, dstep, dwells, iq, 5, flush, bit_transition=bt)
print({k: r[k] for k in ("workload", "ms_per_sweep_one_dwell", "value")})
