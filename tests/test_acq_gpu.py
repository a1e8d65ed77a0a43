"""GPU parity of the PCPS acquisition engine (C ABI gb200_acq_*) against the oracle restatement of
pcps_acquisition::acquisition_core (oracle/oracle_acq.c).

Bar (north_star / SURVEY §8d): code-phase index and Doppler bin IDENTICAL to the oracle for every PRN present (against both
the fp32-FFT and the float64-FFT oracle), Acq_delay_samples / Acq_doppler_hz identical, test statistic within 1e-3
relative; for absent PRNs the accept/reject decision against the pfa threshold is identical.
"""
import os
import sys

import numpy as np
import pytest

import _oracle as O

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
from gen_if import dequantise, gen_if, quantise  # noqa: E402

pytestmark = pytest.mark.gpu


def _engine(**kw):
    from gnss_sim_receiver_b200.acquisition import AcquisitionEngine
    return AcquisitionEngine(**kw)


@pytest.mark.parametrize("n", [8, 20, 60, 1400, 4000, 16000, 20000, 25000, 32000, 50000, 100000, 200000,
                               6625, 5456, 2600, 4092, 13250, 3312, 2650])  # second row: prime factors 11 ... 53 (generic prime-radix stage)
def test_fft_hook_vs_numpy(n):
    rng = np.random.default_rng(n)
    x = (rng.standard_normal(n) + 1j * rng.standard_normal(n)).astype(np.complex64)
    eng = _engine(fs_in=4e6, fft_size=n, max_prns=1)
    X = eng.fft(x)
    ref = np.fft.fft(x.astype(np.complex128))
    err = np.linalg.norm(X - ref) / np.linalg.norm(ref)
    assert err < 2e-6, err
    y = eng.fft(ref.astype(np.complex64), inverse=True)
    err = np.linalg.norm(y - n * x.astype(np.complex128)) / (n * np.linalg.norm(x))
    assert err < 2e-6, err
    eng.close()


def _gps_setup(fs, n_code_periods=1, seed=1, sats=None, prns=range(1, 9), bit_transition=False, noise=True):
    spc = int(round(fs * 1e-3))
    consumed = spc * n_code_periods
    codes = {p: O.gps_l1_ca_code(p) for p in range(1, 33)}
    if sats is None:
        rng = np.random.default_rng(seed)
        sats = [dict(prn=p, doppler=float(rng.uniform(-4800, 4800)), delay_chips=float(rng.uniform(0, 1023)), cn0=float(rng.uniform(46, 52)),
                     phase=float(rng.uniform(0, 6.28))) for p in (1, 3, 6)]
    sig = gen_if(codes, fs, consumed, sats, seed=seed, noise=noise)
    fft_size = consumed * 2 if bit_transition else consumed
    n_code = fft_size // 2 if bit_transition else consumed
    local = {p: np.tile(O.gps_l1_ca_code_sampled(p, fs, spc), n_code_periods * 2)[:max(n_code, consumed)] for p in prns}
    return sig, sats, local, spc, consumed, fft_size


def _oracle_run(conf, sig_dwells, code, pfa=None):
    fc = O.acq_set_local_code(conf, code)
    grid = None
    r = None
    for d, s in enumerate(sig_dwells):
        grid = O.acq_dwell(conf, s, fc, grid, d)
        r = O.acq_statistic(conf, grid, d + 1)
    return r, grid


def _check(r, ro, present, thr=None, stat_tol=1e-3):
    if present:
        assert (r.indext, r.doppler_index) == (ro.indext, ro.doppler_index)
        assert r.doppler_hz == ro.doppler_hz
        assert r.acq_delay_samples == ro.acq_delay_samples and r.acq_doppler_hz == ro.acq_doppler_hz
        assert abs(r.test_statistics - ro.test_statistics) <= stat_tol * ro.test_statistics
        assert abs(r.peak - ro.peak) <= stat_tol * ro.peak
    if thr is not None:
        assert (r.test_statistics > thr) == (ro.test_statistics > thr)


@pytest.mark.parametrize("fs,dmax,dstep", [(4e6, 10000, 250), (25e6, 5000, 250), (6.625e6, 5000, 250)])  # 6625 = 5^3 * 53
def test_gps_l1_cfar_vs_oracle(fs, dmax, dstep):
    """C1 / C3 shape: GPS L1 C/A, one code period, +-dmax in dstep bins, CFAR statistic (pfa threshold)."""
    prns = list(range(1, 9))
    sig, sats, local, spc, consumed, fft_size = _gps_setup(fs, prns=prns)
    spchip = int(np.ceil(fs / 1.023e6))
    eng = _engine(fs_in=fs, fft_size=fft_size, samples_per_code=float(spc), samples_per_chip=spchip, max_prns=len(prns))
    eng.set_doppler_grid(dmax, dstep)
    for i, p in enumerate(prns):
        eng.set_local_code(i, local[p])
    res = eng.run(sig)
    present = {s["prn"] for s in sats}
    thr = eng.threshold(0.001)
    for f64 in (False, True):
        conf = O.acq_conf(fs, consumed, fft_size, dmax, dstep, float(spc), spchip, fft_f64=f64)
        assert abs(O.acq_threshold(conf, 0.001) - thr) <= 1e-6 * thr
        assert eng.num_doppler_bins == conf.num_doppler_bins and eng.neff == O.acq_neff(conf)
        for i, p in enumerate(prns):
            ro, _ = _oracle_run(conf, [sig], local[p])
            _check(res[i], ro, p in present, thr)
            if p in present:
                assert res[i].test_statistics > thr
                s = next(s for s in sats if s["prn"] == p)
                assert abs(res[i].doppler_hz - s["doppler"]) <= dstep
    eng.close()


def test_bit_transition_and_two_peak_statistic():
    fs = 4e6
    prns = [1, 2, 3, 6]
    sig, sats, local, spc, consumed, fft_size = _gps_setup(fs, n_code_periods=1, prns=prns, bit_transition=True, seed=4)
    spchip = 4
    for use_cfar in (True, False):
        eng = _engine(fs_in=fs, fft_size=fft_size, consumed_samples=consumed, samples_per_code=float(spc), samples_per_chip=spchip, bit_transition=True,
                      use_cfar=use_cfar, max_prns=len(prns))
        eng.set_doppler_grid(5000, 250)
        for i, p in enumerate(prns):
            eng.set_local_code(i, local[p])
        res = eng.run(sig)
        conf = O.acq_conf(fs, consumed, fft_size, 5000, 250, float(spc), spchip, bit_transition=True, use_cfar=use_cfar)
        for i, p in enumerate(prns):
            ro, grid = _oracle_run(conf, [sig], local[p])
            _check(res[i], ro, p in {1, 3, 6})
            if not use_cfar:
                g = eng.read_grid(i)
                assert np.max(np.abs(g - grid)) <= 2e-5 * np.max(grid)
        eng.close()


def test_noncoherent_dwells_and_int16_input():
    """C4 shape in small: max_dwells > 1 accumulates |.|^2 over fresh blocks (pcps_acquisition.cc:657-665); ishort input."""
    fs = 4e6
    prns = [1, 3, 5, 6]
    nd = 4
    spc = 4000
    codes = {p: O.gps_l1_ca_code(p) for p in range(1, 33)}
    sats = [dict(prn=1, doppler=1300.0, delay_chips=200.3, cn0=40.0), dict(prn=3, doppler=-2750.0, delay_chips=900.7, cn0=41.0),
            dict(prn=6, doppler=100.0, delay_chips=10.1, cn0=42.0)]
    stream = gen_if(codes, fs, nd * spc, sats, seed=9)
    q = quantise(stream, "ishort")
    sigq = dequantise(q)
    eng = _engine(fs_in=fs, fft_size=spc, samples_per_code=float(spc), samples_per_chip=4, max_dwells=nd, max_prns=len(prns), iq_dtype=1)
    eng.set_doppler_grid(5000, 250)
    local = {p: O.gps_l1_ca_code_sampled(p, fs, spc) for p in prns}
    for i, p in enumerate(prns):
        eng.set_local_code(i, local[p])
    conf = O.acq_conf(fs, spc, spc, 5000, 250, float(spc), 4, max_dwells=nd)
    fcs = {p: O.acq_set_local_code(conf, local[p]) for p in prns}
    grids = {p: None for p in prns}
    for d in range(nd):
        res = eng.run(q[2 * d * spc:2 * (d + 1) * spc], dwell_index=d)
        for i, p in enumerate(prns):
            grids[p] = O.acq_dwell(conf, sigq[d * spc:(d + 1) * spc], fcs[p], grids[p], d)
            ro = O.acq_statistic(conf, grids[p], d + 1)
            _check(res[i], ro, p in {1, 3, 6} and d == nd - 1)
    g = eng.read_grid(0)
    assert np.max(np.abs(g - grids[1])) <= 2e-5 * np.max(grids[1])
    eng.close()


@pytest.mark.parametrize("fs,periods", [(16e6, 2), (25e6, 4)])
def test_split_transform_sizes(fs, periods):
    """fft sizes beyond one SM's shared memory (32000, 100000): first/last stage through global memory."""
    prns = [1, 3, 4]
    sig, sats, local, spc, consumed, fft_size = _gps_setup(fs, n_code_periods=periods, prns=prns, seed=12)
    spchip = int(np.ceil(fs / 1.023e6))
    eng = _engine(fs_in=fs, fft_size=fft_size, samples_per_code=float(spc), samples_per_chip=spchip, max_prns=len(prns))
    eng.set_doppler_grid(1000, 125)
    for i, p in enumerate(prns):
        eng.set_local_code(i, local[p])
    res = eng.run(sig)
    conf = O.acq_conf(fs, consumed, fft_size, 1000, 125, float(spc), spchip)
    for i, p in enumerate(prns):
        ro, _ = _oracle_run(conf, [sig], local[p])
        if p in {1, 3}:
            # the local code spans `periods` code periods, so the circular correlation has `periods` equal peaks (exact
            # ties in exact arithmetic): the index is pinned modulo one code period, which is what Acq_delay_samples reports
            assert res[i].indext % spc == ro.indext % spc and res[i].doppler_index == ro.doppler_index
            assert res[i].acq_delay_samples == ro.acq_delay_samples and res[i].acq_doppler_hz == ro.acq_doppler_hz
            assert abs(res[i].test_statistics - ro.test_statistics) <= 1e-3 * ro.test_statistics
    eng.close()


def test_errors_are_loud():
    from gnss_sim_receiver_b200._capi import GB200Error
    with pytest.raises(GB200Error):
        _engine(fs_in=4e6, fft_size=4000, consumed_samples=4001)  # more samples than the transform holds
    with pytest.raises(GB200Error):
        _engine(fs_in=4e6, fft_size=4001, bit_transition=True)     # bit_transition needs an even block
    e = _engine(fs_in=4e6, fft_size=2222)  # 2 * 11 * 101: not factorable by the plan — served by a padded transform (tests/test_acq_anysize_gpu.py)
    e.close()
    eng = _engine(fs_in=4e6, fft_size=4000, max_prns=2)
    with pytest.raises(GB200Error):
        eng.run(np.zeros(4000, np.complex64))  # no doppler grid
    eng.set_doppler_grid(5000, 250)
    with pytest.raises(GB200Error):
        eng.run(np.zeros(4000, np.complex64))  # no local code
    eng.close()
