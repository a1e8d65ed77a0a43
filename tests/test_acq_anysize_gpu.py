"""Acquisition block lengths the mixed-radix plan cannot factor (a prime factor above 61): FFTW behind gr::fft takes any size
(gnss_sdr_fft.h:27-61), and so does the engine — it computes the SAME circular correlation of length fft_size on a padded transform
(csrc/acq_api.inc: gb200_acq_create / gb200_acq_set_local_code): M >= 2 * fft_size - 1 smooth, input zero-padded, replica laid out wrapped,
lags 0 .. fft_size - 1 searched.

Checker: the restated acquisition_core with numpy's pocketfft (any length, float64) as the DFT and the reference's wipe-off law
(O.acq_wipeoff = volk_gnsssdr_s32f_sincos_32fc_generic restated and pinned), statistics by the oracle (O.acq_statistic).  The whole search
grid within 2e-5 of its peak, indices identical, statistic within 1e-4.
"""
import numpy as np
import pytest

import _oracle as O
import _signals as S

pytestmark = pytest.mark.gpu


def _oracle_grid(conf, sig, local, bit_transition):
    """pcps_acquisition.cc:175-208 + :600-672 with numpy's FFT: grid[bin][neff]"""
    n, consumed = conf.fft_size, conf.consumed
    code = np.zeros(n, np.complex128)
    if bit_transition:
        code[n // 2:] = local[:n // 2]
    else:
        code[:consumed] = local[:consumed]
    fc = np.conj(np.fft.fft(code))
    neff = n // 2 if bit_transition else n
    off = n // 2 if bit_transition else 0
    grid = np.zeros((conf.num_doppler_bins, neff), np.float32)
    for b in range(conf.num_doppler_bins):
        f = -conf.doppler_max + conf.doppler_step * b
        x = np.zeros(n, np.complex128)
        x[:consumed] = (sig[:consumed].astype(np.complex64) * O.acq_wipeoff(consumed, f, conf.fs_in)).astype(np.complex64)
        y = np.fft.ifft(np.fft.fft(x) * fc) * n  # FFTW's inverse is unnormalised
        grid[b] = (np.abs(y[off:off + neff]) ** 2).astype(np.float32)
    return grid


@pytest.mark.parametrize("fs,n,bit_transition,item", [
    (4.001e6, 4001, False, "gr_complex"),       # 4001 is prime
    (6.826e6, 6826, False, "gr_complex"),       # 2 x 3413
    (6.826e6, 13652, True, "cshort"),           # 4 x 3413, two code periods with bit_transition_flag, int16 items
])
def test_unfactorable_block_lengths_through_a_padded_transform(fs, n, bit_transition, item):
    from gnss_sim_receiver_b200.acquisition import AcquisitionEngine
    rng = np.random.default_rng(n)
    prns, present = [3, 7, 12, 20], [7, 20]
    dmax, dstep = 5000, 250
    spc_n = n // 2 if bit_transition else n
    comps = [dict(table=S.gps_l1_table(p), f_chip=1.023e6, doppler=float(rng.uniform(-4000, 4000)), delay=float(rng.uniform(0, 1000)), cn0=50.0, phase=float(rng.uniform(0, 6))) for p in present]
    x = S.synth_if(n, fs, comps, seed=n % 97)
    if item == "cshort":
        q = np.empty(2 * n, np.int16)
        q[0::2], q[1::2] = np.clip(np.rint(x.real * 500), -32767, 32767), np.clip(np.rint(x.imag * 500), -32767, 32767)
        items, sig, iq = q, (q[0::2].astype(np.float32) + 1j * q[1::2].astype(np.float32)).astype(np.complex64), 1
    else:
        items, sig, iq = x, x, 0
    # replica of one code period at this rate (gps_l1_ca_code_gen_complex_sampled's law through the host library)
    locals_ = {}
    for p in prns:
        t = S.gps_l1_table(p)
        locals_[p] = t[(np.floor(np.arange(spc_n) * (1.023e6 / fs)).astype(np.int64)) % 1023].astype(np.complex64)
    spchip = int(np.ceil(fs / 1.023e6))
    eng = AcquisitionEngine(fs_in=fs, fft_size=n, consumed_samples=n, samples_per_code=float(spc_n), samples_per_chip=spchip, bit_transition=bit_transition,
                            use_cfar=True, max_prns=len(prns), iq_dtype=iq, keep_grid=True)
    eng.set_doppler_grid(dmax, dstep)
    for i, p in enumerate(prns):
        eng.set_local_code(i, locals_[p])
    res = eng.run(items)
    conf = O.acq_conf(fs, n, n, dmax, dstep, float(spc_n), spchip, bit_transition=bit_transition)
    assert eng.threshold(0.001) > 0
    for i, p in enumerate(prns):
        want_grid = _oracle_grid(conf, sig, locals_[p], bit_transition)
        got_grid = eng.read_grid(i)
        peak = float(want_grid.max())
        assert got_grid.shape == want_grid.shape
        assert np.max(np.abs(got_grid - want_grid)) <= 2e-5 * peak, (p, float(np.max(np.abs(got_grid - want_grid)) / peak))
        ro = O.acq_statistic(conf, want_grid, 1)
        r = res[i]
        assert (r.indext, r.doppler_index, r.doppler_hz) == (ro.indext, ro.doppler_index, ro.doppler_hz), (p, p in present)
        assert r.acq_delay_samples == ro.acq_delay_samples
        assert abs(r.test_statistics - ro.test_statistics) <= 1e-4 * ro.test_statistics
        if p in present:
            assert r.test_statistics > eng.threshold(0.001)
    # the fft door has no fft_size-point transform on such a handle and says so
    with pytest.raises(Exception, match="padded transform"):
        eng.fft(np.zeros(n, np.complex64))
    eng.close()
