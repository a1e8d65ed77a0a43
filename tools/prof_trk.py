"""Small fixed tracking workload for ncu (same shape as bench.py's step, fewer channels by default)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from gnss_sim_receiver_b200.tracking import TrackingEngine  # noqa: E402

nch = int(sys.argv[1]) if len(sys.argv) > 1 else 10240
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 3
w = bench.tracking_workload(0, nch)
codes = bench.gps_codes()
eng = TrackingEngine(device=0, max_channels=nch, max_code_len=bench.L_CODE, ring_samples=16 * bench.N_BLK, max_block=bench.N_BLK + 8, max_cmds=nch)
for c in range(nch):
    eng.set_code(c, codes[w["prn"][c] - 1])
eng.push_iq(bench.iq_stream(3 * bench.N_BLK))
cmds = eng.make_cmds(np.arange(nch), w["off"], w["len"], w["shifts"], w["rem_carr"], w["phase_step"], w["rem_code"], w["code_step"])
eng.enqueue(cmds)
eng.collect()
ms = eng.replay(iters)
print(f"channels={nch} ms_per_step={ms:.4f} ch-Msps/s={np.sum(w['len']) / 1e6 / (ms / 1e3):.0f}")
