"""Small fixed tracking workload for ncu: bench.py's headline step by default, or any BASELINE shape.
usage: python tools/prof_trk.py [channels] [iters] [shape]      shape in {c3 (default), c1, c2, c4, c5, c3e1}"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from gnss_sim_receiver_b200.tracking import TrackingEngine  # noqa: E402

SHAPES = {  # fs, samples per period, table entries, entries per second, taps, iq_dtype, tap spacing
    "c3": (25e6, 25000, 1023, 1.023e6, 11, 0, 0.1), "c1": (4e6, 4000, 1023, 1.023e6, 3, 0, 0.5), "c2": (4e6, 16000, 8184, 2.046e6, 11, 1, 0.15),
    "c4": (20e6, 20000, 10230, 10.23e6, 11, 0, 0.1), "c5": (100e6, 100000, 1023, 1.023e6, 11, 2, 0.05), "c3e1": (25e6, 100000, 8184, 2.046e6, 11, 0, 0.15),
}
nch = int(sys.argv[1]) if len(sys.argv) > 1 else 10240
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 3
fs, n_blk, L, rate, taps, iq, spacing = SHAPES[sys.argv[3] if len(sys.argv) > 3 else "c3"]
w = bench.tracking_workload(0, nch, fs=fs, n_blk=n_blk, L=L, chip_rate=rate, taps=taps, spacing=spacing)
codes = bench.gps_codes() if L == 1023 else bench.random_codes(32, L)
eng = TrackingEngine(device=0, max_channels=nch, max_code_len=L, iq_dtype=iq, ring_samples=4 * n_blk, max_block=n_blk + 8, max_cmds=nch)
for c in range(nch):
    eng.set_code(c, codes[w["prn"][c] - 1])
x = bench.iq_stream(3 * n_blk)
eng.push_iq(x if iq == 0 else bench.quantise(x, np.int16 if iq == 1 else np.int8))
cmds = eng.make_cmds(np.arange(nch), w["off"], w["len"], w["shifts"], w["rem_carr"], w["phase_step"], w["rem_code"], w["code_step"])
eng.enqueue(cmds)
eng.collect()
ms = eng.replay(iters)
print(f"{eng.last_kernel} channels={nch} ms_per_step={ms:.4f} ch-Msps/s={np.sum(w['len']) / 1e6 / (ms / 1e3):.0f}")
