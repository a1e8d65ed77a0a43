#!/usr/bin/env python
"""One tracking configuration of bench.py on its own, for ncu captures of the general kernel:
    ncu --set full --import-source on --kernel-name regex:trk_correlate_kernel --launch-skip 3 --launch-count 1 -o <out> python tools/prof_c4.py [C1|C2|C4]
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402
import bench  # noqa: E402

which = sys.argv[1] if len(sys.argv) > 1 else "C4"
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda:0")
shapes = {
    "C1": ("C1 GPS L1 4 MS/s gr_complex, 3 taps", 4e6, 4000, 1023, 1.023e6, 3, 0, 8192),
    "C2": ("C2 Galileo E1 4 MS/s ishort, 11 taps", 4e6, 16000, 8184, 2.046e6, 11, 1, 4096),
    "C4": ("C4 GPS L5 / Galileo E5a 20 MS/s gr_complex, 11 taps", 20e6, 20000, 10230, 10.23e6, 11, 0, 4096),
}
name, fs, n_blk, L, chip_rate, taps, iq, nch = shapes[which]
r = bench.trk_config_bench(0, name, fs, n_blk, L, chip_rate, taps, iq, nch, 5, flush)
print({k: r[k] for k in ("workload", "ms_per_step", "value", "kernel", "fp32_frac")})
