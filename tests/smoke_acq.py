"""Acquisition half of __graft_entry__.smoke(): one small PCPS sweep on cuda:0 checked against the oracle."""
import os
import sys

import numpy as np

import _oracle as O

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))


def run():
    from gen_if import gen_if
    from gnss_sim_receiver_b200.acquisition import AcquisitionEngine
    fs, n = 4e6, 4000
    codes = {p: O.gps_l1_ca_code(p) for p in (1, 2)}
    sig = gen_if(codes, fs, n, [dict(prn=1, doppler=1680.0, delay_chips=889.0, cn0=48.0)], seed=2)
    eng = AcquisitionEngine(fs_in=fs, fft_size=n, samples_per_code=float(n), samples_per_chip=4, max_prns=2)
    eng.set_doppler_grid(5000, 250)
    local = {p: O.gps_l1_ca_code_sampled(p, fs, n) for p in (1, 2)}
    for i, p in enumerate((1, 2)):
        eng.set_local_code(i, local[p])
    res = eng.run(sig)
    conf = O.acq_conf(fs, n, n, 5000, 250, float(n), 4)
    ro = O.acq_statistic(conf, O.acq_dwell(conf, sig, O.acq_set_local_code(conf, local[1])))
    r = res[0]
    assert (r.indext, r.doppler_index) == (ro.indext, ro.doppler_index), (r.indext, r.doppler_index, ro.indext, ro.doppler_index)
    assert abs(r.test_statistics - ro.test_statistics) < 1e-3 * ro.test_statistics
    assert r.test_statistics > eng.threshold(0.001) > res[1].test_statistics
    eng.close()
    print(f"smoke ok: acquisition PRN 1 at code phase {r.indext}, Doppler {r.doppler_hz} Hz, statistic {r.test_statistics:.1f}")
