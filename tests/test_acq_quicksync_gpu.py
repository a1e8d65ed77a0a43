"""the QuickSync acquisition strategy (`GPS_L1_CA_PCPS_QuickSync_Acquisition`) on the GPU engine.

Referee: the reference's OWN block — pcps_quicksync_acquisition_cc.cc compiled unmodified into oracle/_ref (oracle/ref_acqqs.cc, FFTW replaced by
oracle/oracle_fft.c), made and driven like the reference's adapter does.  On the same stream, cut into the same vector items and handed over in the
same scheduler calls, the product block (host/b200_pcps_quicksync_acquisition.cc over gb200_acq_set_input_folding, gb200_acq_run,
gb200_acq_read_cells and gb200_acq_folded_candidates) must consume the same items, raise the same event and leave the same Gnss_Synchro fields —
resolved code phase, Doppler and sample stamp exactly, input power exactly (same summation order), peak / statistic within 1e-4.
"""
import numpy as np
import pytest

import _oracle as O
from test_acq_quicksync_refblock_cpu import FS, gps_stream

pytestmark = pytest.mark.gpu


def _conf(folding=None, sampled_ms=4, max_dwells=1, bit_transition=False, dmax=5000, fs=FS, extra=""):
    r = "Acquisition_1C"
    return (f"GNSS-SDR.internal_fs_sps={int(fs)}\n{r}.implementation=GPS_L1_CA_PCPS_QuickSync_Acquisition_B200\n{r}.item_type=gr_complex\n{r}.doppler_max={dmax}\n"
            f"{r}.coherent_integration_time_ms={sampled_ms}\n{r}.max_dwells={max_dwells}\n{r}.bit_transition_flag={'true' if bit_transition else 'false'}\n"
            + (f"{r}.folding_factor={folding}\n" if folding else "") + extra)


def _both(x, prn, threshold, max_items, folding=None, sampled_ms=4, max_dwells=1, bit_transition=False, dstep=250, dmax=5000, fs=FS):
    from gnss_sim_receiver_b200 import _hostapi as H
    a = H.QuickSyncAcquisitionBlock(_conf(folding, sampled_ms, max_dwells, bit_transition, dmax, fs), "Acquisition_1C", prn)
    a.arm(dmax, dstep, threshold)
    ca = a.work(x, max_items)
    sa = a.status()
    a.close()
    b = O.RefAcqQs(fs, folding or 0, sampled_ms, max_dwells, dmax, prn, bit_transition=bit_transition)
    b.arm(dmax, dstep, threshold)
    cb = b.work(x, max_items)
    sb = b.status()
    b.close()
    assert ca == cb and sa.sample_counter == sb.sample_counter
    assert (sa.folding_factor, sa.sampled_ms, sa.item_length, sa.fft_size, sa.num_doppler_bins) == (sb.folding_factor, sb.sampled_ms, sb.item_length, sb.fft_size, sb.num_doppler_bins)
    assert [sa.last_event] == list(sb.events[:sb.n_events]) and (sa.state, sa.active, sa.well_count) == (sb.state, sb.active, sb.well_count)
    assert sa.input_power == sb.input_power
    assert sa.acq_delay_samples == sb.acq_delay_samples and sa.acq_doppler_hz == sb.acq_doppler_hz, (sa.acq_delay_samples, sb.acq_delay_samples, sa.acq_doppler_hz, sb.acq_doppler_hz)
    assert sa.acq_samplestamp_samples == sb.acq_samplestamp_samples and sa.acq_doppler_step == sb.acq_doppler_step == dstep
    assert abs(sa.mag - sb.mag) <= 1e-4 * sb.mag and abs(sa.test_statistics - sb.test_statistics) <= 1e-4 * sb.test_statistics, (sa.mag, sb.mag)
    return sa, sb


@pytest.mark.skipif(not O.ref_available(), reason="oracle/_ref not built")
@pytest.mark.parametrize("folding,sampled_ms,doppler,delay_chips,fs", [
    (None, 4, 1830.0, 311.0, 4e6),   # the adapter's default folding factor: ceil(sqrt(log2(4000))) = 4
    (2, 2, -2440.0, 833.0, 4e6),
    (4, 4, 600.0, 977.0, 4e6),
    (3, 4, 1830.0, 120.0, 4e6),      # 4000 / 3 = 1333 = 31 x 43 points: the engine's padded transform; coherent time rounded down to 3 ms
    (2, 2, 3300.0, 500.0, 2e6),
    (4, 4, -900.0, 44.0, 8e6),
])
def test_positive_acquisition_equals_the_reference_block(folding, sampled_ms, doppler, delay_chips, fs):
    prn = 7
    x = gps_stream(prn, doppler, delay_chips, 64, fs=fs)
    sa, sb = _both(x, prn, 0.02, 2, folding=folding, sampled_ms=sampled_ms, fs=fs)
    assert sa.last_event == 1
    assert abs(sa.acq_delay_samples - delay_chips * (fs / 1000) / 1023) <= 2.0 and abs(sa.acq_doppler_hz - doppler) <= 125.0


@pytest.mark.skipif(not O.ref_available(), reason="oracle/_ref not built")
def test_dwells_negative_acquisition_bit_transition_and_pfa():
    from gnss_sim_receiver_b200 import _hostapi as H
    x = gps_stream(7, 1830.0, 311.0, 64)
    sa, sb = _both(x, 19, 1.0, 1, folding=2, sampled_ms=2, max_dwells=3)  # PRN 19 is not in the air
    assert sa.last_event == 2 and sa.well_count == 3
    sa, sb = _both(x, 7, 0.02, 1, folding=4, sampled_ms=4, bit_transition=True)
    assert sa.last_event == 1 and sa.well_count == 2
    # pfa in the configuration wins over the threshold handed in (adapters/gps_l1_ca_pcps_quicksync_acquisition.cc:154-176, :274-291)
    a = H.QuickSyncAcquisitionBlock(_conf(4, extra="Acquisition_1C.pfa=0.001\n"), "Acquisition_1C", 3)
    a.arm(5000, 250, 123.0)
    want = -np.log1p(-((1.0 - float(np.float32(0.001))) ** (1.0 / (1000 * 41)))) / 1000.0
    assert abs(a.status().threshold - want) <= 1e-6 * want
    a.close()


@pytest.mark.skipif(not O.ref_available(), reason="oracle/_ref not built")
@pytest.mark.parametrize("folding,sampled_ms,signal,cboc", [(None, 8, "1B", False), (2, 16, "1B", True), (4, 16, "1C", False)])
def test_galileo_e1_adapter_equals_the_reference_block(folding, sampled_ms, signal, cboc):
    """`Galileo_E1_PCPS_QuickSync_Ambiguous_Acquisition_B200`: the same block over 4 ms E1 codes (E1-B or the E1-C pilot, sinBOC or CBOC replica), vector
    items of sampled_ms x samples per ms — longer than the searched part when the coherent time exceeds folding_factor code periods"""
    import _signals as S
    from gnss_sim_receiver_b200 import _hostapi as H
    prn, doppler, delay_chips, dmax, dstep = 11, 1830.0, 1500.0, 5000, 125
    comps = [dict(table=S.e1_table(prn, signal), f_chip=2 * 1.023e6, doppler=doppler, delay=2 * delay_chips, cn0=50.0, phase=1.0)]
    x = S.synth_if(160 * 4000, FS, comps, seed=3)
    r = "Acquisition_1B"
    conf = (f"GNSS-SDR.internal_fs_sps={int(FS)}\n{r}.implementation=Galileo_E1_PCPS_QuickSync_Ambiguous_Acquisition_B200\n{r}.item_type=gr_complex\n{r}.doppler_max={dmax}\n"
            f"{r}.coherent_integration_time_ms={sampled_ms}\nAcquisition0.cboc={'true' if cboc else 'false'}\n" + (f"{r}.folding_factor={folding}\n" if folding else ""))
    a = H.QuickSyncAcquisitionBlock(conf, r, prn, "E", signal)
    a.arm(dmax, dstep, 0.001)
    ca = a.work(x, 2)
    sa = a.status()
    a.close()
    b = O.RefAcqQs(FS, folding or 0, sampled_ms, 1, dmax, prn, galileo_signal=signal, cboc=cboc)
    b.arm(dmax, dstep, 0.001)
    cb = b.work(x, 2)
    sb = b.status()
    b.close()
    assert ca == cb == 3 and sa.sample_counter == sb.sample_counter and [sa.last_event] == list(sb.events[:sb.n_events]) == [1]
    assert (sa.folding_factor, sa.sampled_ms, sa.item_length, sa.fft_size, sa.num_doppler_bins) == (sb.folding_factor, sb.sampled_ms, sb.item_length, sb.fft_size, sb.num_doppler_bins)
    assert sa.item_length == sampled_ms * 4000 and sa.fft_size == 16000 // (folding or 2)
    assert sa.input_power == sb.input_power and sa.acq_delay_samples == sb.acq_delay_samples and sa.acq_doppler_hz == sb.acq_doppler_hz
    assert sa.acq_samplestamp_samples == sb.acq_samplestamp_samples == sampled_ms * 4000 and sa.acq_doppler_step == sb.acq_doppler_step == dstep
    assert abs(sa.mag - sb.mag) <= 1e-4 * sb.mag and abs(sa.test_statistics - sb.test_statistics) <= 1e-4 * sb.test_statistics
    assert abs(sa.acq_delay_samples - delay_chips * 16000 / 4092) <= 2.0 and abs(sa.acq_doppler_hz - doppler) <= 62.5


def test_errors_are_loud():
    from gnss_sim_receiver_b200 import _hostapi as H
    with pytest.raises(Exception, match="gr_complex"):
        H.QuickSyncAcquisitionBlock(_conf().replace("item_type=gr_complex", "item_type=cshort"), "Acquisition_1C", 1)
    with pytest.raises(Exception, match="folding_factor"):
        H.QuickSyncAcquisitionBlock(_conf(1), "Acquisition_1C", 1)  # the reference reads past its wiped block with folding factor 1


def test_folded_run_against_numpy():
    """gb200_acq_set_input_folding on its own: the grid of a folded run against numpy (wipe-off of the whole block, chunks added up, circular
    correlation with the folded replica)"""
    from gnss_sim_receiver_b200.acquisition import AcquisitionEngine
    rng = np.random.default_rng(5)
    m, k, length, nb = 1000, 16, 16000 + 24, 5
    eng = AcquisitionEngine(fs_in=4e6, fft_size=m, consumed_samples=m, samples_per_code=float(m), samples_per_chip=1, use_cfar=False, max_prns=1, keep_grid=True)
    freqs = np.linspace(-1000, 1000, nb).astype(np.float32)
    eng.set_doppler_bins(freqs)
    code = rng.choice(np.array([-1.0, 1.0]), m).astype(np.complex64)
    eng.set_local_code(0, code)
    eng.set_input_folding(k, length)
    x = (rng.standard_normal(length) + 1j * rng.standard_normal(length)).astype(np.complex64)
    eng.run(x, slots=[0])
    g = eng.read_grid(0)
    fc = np.conj(np.fft.fft(code.astype(np.complex128)))
    for b in range(nb):
        w = O.acq_wipeoff(length, float(freqs[b]), 4e6).astype(np.complex128)
        folded = (x.astype(np.complex128) * w)[:k * m].reshape(k, m).sum(axis=0)
        want = np.abs(np.fft.ifft(np.fft.fft(folded) * fc) * m) ** 2
        assert np.max(np.abs(g[b] - want)) <= 2e-5 * want.max(), (b, float(np.max(np.abs(g[b] - want)) / want.max()))
    # gb200_acq_folded_candidates on the same block: acc[i] = sum_j x[d_i + j] * wipeoff_bin[d_i + j] * code[j], d_i = index + i * fft_size
    long_code = (1j * rng.choice(np.array([-1.0, 1.0]), 4000)).astype(np.complex64)
    b, index, ncand = 3, 137, 4
    got = eng.folded_candidates(b, index, long_code, ncand)
    w = O.acq_wipeoff(length, float(freqs[b]), 4e6).astype(np.complex128)
    for i in range(ncand):
        d = index + i * m
        want_i = np.sum(x[d:d + 4000].astype(np.complex128) * w[d:d + 4000] * long_code.astype(np.complex128))
        assert abs(got[i] - want_i) <= 2e-5 * np.sqrt(4000.0) * 2, (i, got[i], want_i)
    with pytest.raises(Exception, match="outside the block"):
        eng.folded_candidates(b, index, long_code, 14)  # 137 + 13 * 1000 + 4000 > 16024
    eng.close()
