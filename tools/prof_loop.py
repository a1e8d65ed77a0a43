"""Closed-loop step for ncu: 10240 channels, GPS L1 25 MS/s, 11 taps, noise input (the loop's work per period does not depend on lock)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from gnss_sim_receiver_b200.tracking import TrackingEngine, TrackingLoop, loop_config  # noqa: E402

nch = int(sys.argv[1]) if len(sys.argv) > 1 else 10240
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 30
n = bench.N_BLK
codes = bench.gps_codes()
eng = TrackingEngine(device=0, max_channels=nch, max_code_len=1023, ring_samples=16 * n, max_block=n + 8, max_cmds=nch)
cfg = loop_config("1C", bench.FS, pull_in_time_s=1, cn0_samples=20, cn0_min=25, max_code_lock_fail=50, max_carrier_lock_fail=5000, carrier_lock_th=0.5,
                  early_late_space_chips=0.5, n_correlators=11)
for c in range(nch):
    eng.set_code(c, codes[c % 32])
loop = TrackingLoop(eng, cfg, nch)
rng = np.random.default_rng(1)
for c in range(nch):
    loop.start(c, c % 32 + 1, float(rng.uniform(0, n)), float(rng.uniform(-4000, 4000)), 0, 0)
x = bench.iq_stream(n, seed=2)
ms = 0.0
for i in range(steps):
    loop.step(x)
    if i >= steps // 2:
        ms += loop.last_step["device_ms"]
print(f"{loop.kernel} channels={nch} device ms per period={ms / (steps - steps // 2):.4f}")
