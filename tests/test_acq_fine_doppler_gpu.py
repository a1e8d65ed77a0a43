"""the fine-Doppler acquisition strategy (`GPS_L1_CA_PCPS_Acquisition_Fine_Doppler`) on the GPU engines.

Referee: the reference's OWN block — src/algorithms/acquisition/gnuradio_blocks/pcps_acquisition_fine_doppler_cc.cc compiled unmodified into
oracle/_ref (oracle/ref_acqfd.cc, FFTW replaced by oracle/oracle_fft.c) — filled and driven like the reference's adapter does.  On the same
stream, with the same scheduler calls, the product block (host/b200_pcps_acquisition_fine_doppler.cc over gb200_acq_run and
gb200_acq_spectrum_peak) must consume the same items, raise the same event and leave the same Gnss_Synchro fields: code phase, coarse /
fine Doppler (exactly: both are table values), sample stamp, Doppler step; the first-to-second-peak statistic within 1e-4.

The block's quirk is part of the contract: pcps_acquisition_fine_doppler_cc.cc:278 computes the coarse Doppler in unsigned arithmetic, so a
satellite below zero Doppler is reported at 2^32 - |f| and the fine estimate never replaces it (:393).  The reference block shows it
(tests/test_acq_fine_doppler_refblock_cpu.py) and the product reproduces it.
"""
import numpy as np
import pytest

import _oracle as O
import _signals as S

pytestmark = pytest.mark.gpu


def _conf(fs, dwells, dmax=5000):
    return (f"GNSS-SDR.internal_fs_sps={int(fs)}\nAcquisition_1C.implementation=GPS_L1_CA_PCPS_Acquisition_Fine_Doppler_B200\nAcquisition_1C.item_type=gr_complex\n"
            f"Acquisition_1C.doppler_max={dmax}\nAcquisition_1C.max_dwells={dwells}\n")


@pytest.mark.parametrize("m,fold,pad", [(20000, 1, 8), (20000, 2, 8), (40000, 1, 8), (25000, 5, 4)])
def test_spectrum_peak_equals_the_padded_transform(m, fold, pad):
    """gb200_acq_spectrum_peak against numpy's float64 transform of the zero-padded product (estimate_Doppler's :366-374), one shared-memory
    block (20000), a split transform (40000) and folded inputs"""
    from gnss_sim_receiver_b200.acquisition import AcquisitionEngine
    rng = np.random.default_rng(m + fold)
    n = m * fold
    rep_len = n // 10
    rep = (1j * rng.choice(np.array([-1.0, 1.0]), rep_len)).astype(np.complex64)
    eng = AcquisitionEngine(fs_in=4e6, fft_size=m, consumed_samples=m, samples_per_code=float(m), samples_per_chip=4, use_cfar=False, max_prns=1)
    eng.set_doppler_bins(np.array([0.0], np.float32))
    for k_true in (137.3, pad * n - 2211.6):  # a positive and a negative frequency, off the bin centres
        t = np.arange(n)
        x = (np.exp(2j * np.pi * k_true * t / (pad * n)) * np.conj(np.tile(rep, 10)) / 1.0 + 0.5 * (rng.standard_normal(n) + 1j * rng.standard_normal(n))).astype(np.complex64)
        idx, val = eng.spectrum_peak(x, rep, pad)
        y = x.astype(np.complex128) * np.tile(rep, 10).astype(np.complex128)
        p = np.abs(np.fft.fft(y, pad * n)) ** 2
        want = int(np.argmax(p))
        assert idx == want == int(round(k_true)) % (pad * n), (idx, want, k_true)
        assert abs(val - p[want]) <= 1e-4 * p[want]
    # an all-zero block: volk_gnsssdr_32f_index_max_32u answers index 0
    assert eng.spectrum_peak(np.zeros(n, np.complex64), rep, pad) == (0, 0.0)
    with pytest.raises(Exception, match="multiple"):
        eng.spectrum_peak(np.zeros(n + 1, np.complex64), rep, pad)
    eng.close()


def _both(fs, prn, comps, dwells, threshold, noutput, n_ms, seed, dmax=5000, dstep=250):
    from gnss_sim_receiver_b200 import _hostapi as H
    n_per_ms = int(fs / 1000)
    x = S.synth_if(n_ms * n_per_ms, fs, comps, seed=seed)
    a = H.FineDopplerAcquisitionBlock(_conf(fs, dwells, dmax), "Acquisition_1C", prn)
    a.arm(dmax, dstep, threshold)
    ca = a.work(x, noutput)
    sa = a.status()
    a.close()
    b = O.RefAcqFd(fs, dmax, dwells, prn)
    b.arm(dmax, dstep, threshold)
    cb = b.work(x, noutput)
    sb = b.status()
    b.close()
    assert ca == cb and sa.sample_counter == sb.sample_counter == cb
    assert [sa.last_event] == list(sb.events[:sb.n_events]) and sa.positive_acq == sb.positive_acq
    assert (sa.state, sa.active) == (sb.state, sb.active) == (0, 0)
    assert (sa.num_doppler_points, sa.fft_size) == (sb.num_doppler_points, sb.fft_size) == (2 * dmax // dstep, n_per_ms)
    assert sa.acq_delay_samples == sb.acq_delay_samples and sa.acq_samplestamp_samples == sb.acq_samplestamp_samples == dwells * n_per_ms
    assert sa.acq_doppler_step == sb.acq_doppler_step == dstep
    assert sa.acq_doppler_hz == sb.acq_doppler_hz, (sa.acq_doppler_hz, sb.acq_doppler_hz, sa.fine_index)
    assert abs(sa.test_statistics - sb.test_statistics) <= 1e-4 * sb.test_statistics, (sa.test_statistics, sb.test_statistics)
    return sa, sb


@pytest.mark.skipif(not O.ref_available(), reason="oracle/_ref not built")
@pytest.mark.parametrize("fs,prn,doppler,dwells,noutput", [
    (4e6, 7, 1830.0, 2, 4000), (4e6, 19, 3216.0, 1, 1000), (4e6, 3, 405.0, 4, 4000), (2e6, 11, 2718.0, 2, 2000), (2.5e6, 23, 1111.0, 3, 2500)])
def test_positive_acquisition_equals_the_reference_block(fs, prn, doppler, dwells, noutput):
    comps = [dict(table=S.gps_l1_table(prn), f_chip=1.023e6, doppler=doppler, delay=311.0 + prn, cn0=50.0, phase=1.0)]
    sa, sb = _both(fs, prn, comps, dwells, 2.5, noutput, 16 + dwells, seed=prn)
    assert sa.last_event == 1
    # the fine estimate replaced the grid value (250 Hz bins) by the padded spectrum's bin: fs / (80 * samples per ms) = 12.5 Hz apart
    assert abs(sa.acq_doppler_hz - doppler) <= 12.5 and sa.acq_doppler_hz % 12.5 == 0.0 and sa.fine_index >= 0
    n_per_ms = int(fs / 1000)
    assert sa.fine_index == int(round(sa.acq_doppler_hz / (fs / (80 * n_per_ms))))


@pytest.mark.skipif(not O.ref_available(), reason="oracle/_ref not built")
def test_negative_doppler_keeps_the_reference_blocks_unsigned_value():
    prn, doppler = 12, -2490.0
    comps = [dict(table=S.gps_l1_table(prn), f_chip=1.023e6, doppler=doppler, delay=77.0, cn0=50.0, phase=0.3)]
    sa, sb = _both(4e6, prn, comps, 2, 2.5, 4000, 18, seed=5)
    assert sa.last_event == 1 and sa.acq_doppler_hz == float(2 ** 32 - 2500)
    # the fine estimate ran and found the satellite's frequency in the upper half of the padded spectrum (negative frequencies), :384-391
    n_ext = 80 * 4000
    assert abs((sa.fine_index - n_ext) * (4e6 / n_ext) - doppler) <= 12.5


@pytest.mark.skipif(not O.ref_available(), reason="oracle/_ref not built")
def test_negative_acquisition_equals_the_reference_block():
    """a satellite that is not there: both blocks take every dwell, publish event 2 and consume one more scheduler call (state 5, :533-550)"""
    comps = [dict(table=S.gps_l1_table(9), f_chip=1.023e6, doppler=1200.0, delay=500.0, cn0=48.0, phase=0.0)]
    sa, sb = _both(4e6, 25, comps, 3, 2.5, 4000, 12, seed=8)
    assert sa.last_event == 2 and sa.positive_acq == 0 and sa.fine_index == -1
    assert sa.sample_counter == 3 * 4000 + 4000


def test_errors_are_loud():
    from gnss_sim_receiver_b200 import _hostapi as H
    with pytest.raises(Exception, match="gr_complex"):
        H.FineDopplerAcquisitionBlock(_conf(4e6, 1).replace("item_type=gr_complex", "item_type=cshort"), "Acquisition_1C", 1)
    with pytest.raises(Exception, match="max_dwells"):
        H.FineDopplerAcquisitionBlock(_conf(4e6, 60), "Acquisition_1C", 1)
    a = H.FineDopplerAcquisitionBlock(_conf(4e6, 1), "Acquisition_1C", 1)
    with pytest.raises(Exception, match="doppler_step"):
        a.arm(5000, 0, 2.5)
    a.close()
