"""Fixed C3 acquisition sweep for ncu (32 PRN x 80 bins, fft 25000)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from gnss_sim_receiver_b200.acquisition import AcquisitionEngine  # noqa: E402
import numpy as np  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else bench.N_BLK
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 2
fs = n * 1000.0
eng = AcquisitionEngine(fs_in=fs, fft_size=n, samples_per_code=float(n), samples_per_chip=int(np.ceil(fs / 1.023e6)), max_prns=32)
eng.set_doppler_grid(10000, 250)
codes = bench.gps_codes()
for p in range(32):
    eng.set_local_code(p, bench.gps_code_sampled(codes[p], fs, n))
eng.run(bench.iq_stream(n, seed=3))
ms = eng.replay(iters)
print(f"fft={n} ms_per_sweep={ms:.3f} cells/s={32 * eng.num_doppler_bins / (ms / 1e3):.0f}")
