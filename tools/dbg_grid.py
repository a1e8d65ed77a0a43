import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "tools"))
import numpy as np, _oracle as O
from gen_if import gen_if
from gnss_sim_receiver_b200.acquisition import AcquisitionEngine
fs, spc = 4e6, 4000
codes = {p: O.gps_l1_ca_code(p) for p in range(1, 33)}
sats = [dict(prn=1, doppler=1300.0, delay_chips=200.3, cn0=50.0)]
sig = gen_if(codes, fs, spc, sats, seed=9)
local = O.gps_l1_ca_code_sampled(1, fs, spc)
eng = AcquisitionEngine(fs_in=fs, fft_size=spc, samples_per_code=float(spc), samples_per_chip=4, max_prns=1, keep_grid=True)
eng.set_doppler_grid(5000, 250)
eng.set_local_code(0, local)
r = eng.run(sig)[0]
g = eng.read_grid(0)
for f64 in (0, 1):
    conf = O.acq_conf(fs, spc, spc, 5000, 250, float(spc), 4, fft_f64=bool(f64))
    fc = O.acq_set_local_code(conf, local)
    go = O.acq_dwell(conf, sig, fc)
    print("oracle f64=%d: max|d|/max = %.3e, at peak rel = %.3e" % (f64, np.max(np.abs(g - go)) / go.max(), abs(g.max() - go.max()) / go.max()))
# exact double: wipe-off exact, numpy FFT
n = np.arange(spc)
C = np.conj(np.fft.fft(local.astype(np.complex128)))
gd = np.zeros_like(g, dtype=np.float64)
gw = np.zeros_like(g, dtype=np.float64)
for k in range(40):
    f = -5000 + 250 * k
    w = np.exp(-2j * np.pi * f * n / fs)
    gd[k] = np.abs(np.fft.ifft(np.fft.fft(sig.astype(np.complex128) * w) * C) * spc) ** 2
    w32 = O.acq_wipeoff(spc, f, fs).astype(np.complex128)
    gw[k] = np.abs(np.fft.ifft(np.fft.fft(sig.astype(np.complex128) * w32) * C) * spc) ** 2
print("gpu vs exact-double: %.3e" % (np.max(np.abs(g - gd)) / gd.max()))
print("oracle-wipeoff(double fft) vs exact-double: %.3e" % (np.max(np.abs(gw - gd)) / gd.max()))
print("oracle fp32 vs oracle-wipeoff(double fft): %.3e" % (np.max(np.abs(go - gw)) / gd.max()))
