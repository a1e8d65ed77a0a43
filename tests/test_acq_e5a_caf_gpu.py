"""the Galileo E5a non-coherent I/Q acquisition strategy with CAF filtering (`Galileo_E5a_Noncoherent_IQ_Acquisition_CAF`) on the GPU engine.

Referee: the reference's OWN block — galileo_e5a_noncoherent_iq_acquisition_caf_cc.cc compiled unmodified into oracle/_ref (oracle/ref_acqcaf.cc,
FFTW replaced by oracle/oracle_fft.c), made and fed its replicas like the reference's adapter does.  On the same stream and the same scheduler
calls the product block (host/b200_e5a_noncoherent_iq_acquisition_caf.cc: one engine sweep over the I_A / Q_A / I_B / Q_B code-spectrum
slots, gb200_acq_read_cells, gb200_acq_grid_probe, gb200_acq_grid_pair_max) must consume the same items, raise the same event and leave the
same Gnss_Synchro fields — code phase, Doppler (of the peak, or of the CAF maximum) and sample stamp exactly, input power exactly (same
summation order), peak / statistic within 1e-4, the filtered CAF vector within 1e-4 of its maximum.
"""
import numpy as np
import pytest

import _oracle as O
from test_acq_e5a_caf_refblock_cpu import FS, e5a_stream

pytestmark = pytest.mark.gpu


def _conf(sampled_ms=1, max_dwells=1, bit_transition=False, caf=0, zero_padding=0, channel_signal="5X", dmax=5000, fs=FS, extra=""):
    r = "Acquisition_5X"
    return (f"GNSS-SDR.internal_fs_sps={int(fs)}\nChannel.signal={channel_signal}\n{r}.implementation=Galileo_E5a_Noncoherent_IQ_Acquisition_CAF_B200\n{r}.item_type=gr_complex\n"
            f"{r}.doppler_max={dmax}\n{r}.coherent_integration_time_ms={sampled_ms}\n{r}.max_dwells={max_dwells}\n{r}.bit_transition_flag={'true' if bit_transition else 'false'}\n"
            f"{r}.CAF_window_hz={caf}\n{r}.Zero_padding={zero_padding}\n" + extra)


def _both(x, prn, threshold, chunk, signal="5X", dstep=250, dmax=5000, **kw):
    from gnss_sim_receiver_b200 import _hostapi as H
    a = H.E5aCafAcquisitionBlock(_conf(dmax=dmax, **kw), "Acquisition_5X", prn, signal)
    a.arm(dmax, dstep, threshold)
    ca = a.work(x, chunk)
    sa, va = a.status(), a.caf_vector()
    a.close()
    b = O.RefAcqCaf(FS, kw.get("sampled_ms", 1), kw.get("max_dwells", 1), dmax, prn, signal=signal, channel_signal=kw.get("channel_signal", "5X"),
                    bit_transition=kw.get("bit_transition", False), caf_window_hz=kw.get("caf", 0), zero_padding=kw.get("zero_padding", 0))
    b.arm(dmax, dstep, threshold)
    cb = b.work(x, chunk)
    sb, vb = b.status(), b.caf_vector()
    b.close()
    assert ca == cb and sa.sample_counter == sb.sample_counter == cb
    assert [sa.last_event] == list(sb.events[:sb.n_events]) and (sa.state, sa.active) == (sb.state, sb.active) == (0, 0)
    assert (sa.num_doppler_bins, sa.fft_size, sa.well_count) == (sb.num_doppler_bins, sb.fft_size, sb.well_count)
    assert sa.input_power == sb.input_power
    assert sa.acq_delay_samples == sb.acq_delay_samples and sa.acq_doppler_hz == sb.acq_doppler_hz, (sa.acq_delay_samples, sb.acq_delay_samples, sa.acq_doppler_hz, sb.acq_doppler_hz)
    assert sa.acq_samplestamp_samples == sb.acq_samplestamp_samples and sa.acq_doppler_step == sb.acq_doppler_step == dstep
    assert abs(sa.mag - sb.mag) <= 1e-4 * sb.mag and abs(sa.test_statistics - sb.test_statistics) <= 1e-4 * sb.test_statistics, (sa.mag, sb.mag, sa.test_statistics, sb.test_statistics)
    assert va.shape == vb.shape
    if len(vb):
        assert np.max(np.abs(va - vb)) <= 1e-4 * np.max(vb), float(np.max(np.abs(va - vb)) / np.max(vb))
    return sa, sb


PRN, DOPPLER, DELAY = 11, 1830.0, 3333.0
FLIP_I, FLIP_Q = (1, -1, -1, 1, 1, -1), (1, 1, -1, -1, 1, -1)


@pytest.mark.skipif(not O.ref_available(), reason="oracle/_ref not built")
@pytest.mark.parametrize("name,kw,signal,sym,chunk", [
    ("1 ms, both components", dict(), "5X", ((1,), (1,)), 5000),
    ("1 ms, one 5X-composed replica", dict(channel_signal="5I"), "5I", ((1,), (1,)), 12000),
    ("3 ms, both components, sign flips between periods: A / B replicas chosen per bin", dict(sampled_ms=3), "5X", (FLIP_I, FLIP_Q), 5000),
    ("3 ms, both, CAF 3 kHz", dict(sampled_ms=3, caf=3000), "5X", (FLIP_I, FLIP_Q), 7000),
    ("1 ms, both, CAF 2 kHz", dict(caf=2000), "5X", ((1,), (1,)), 4000),
    ("zero padding (1 ms code + 1 ms zeros), CAF", dict(zero_padding=1, caf=2000), "5X", ((1,), (1,)), 5000),
    ("2 ms, single replica, bit transition flag, 2 dwells", dict(sampled_ms=2, max_dwells=2, bit_transition=True, channel_signal="5I"), "5I", ((1, -1, 1, 1), (1, -1, 1, 1)), 5000),
    ("2 ms, single replica, CAF", dict(sampled_ms=2, caf=1500, channel_signal="5Q"), "5Q", ((1, -1, 1, 1), (1, 1, -1, 1)), 24000),
])
def test_search_equals_the_reference_block(name, kw, signal, sym, chunk):
    x = e5a_stream(PRN, DOPPLER, DELAY, 14, sym_i=sym[0], sym_q=sym[1], seed=len(name))
    sa, sb = _both(x, PRN, 0.0, chunk, signal=signal, **kw)
    assert sa.last_event == 1, name
    if not kw.get("caf"):
        assert abs(sa.acq_delay_samples - DELAY * 12000 / 10230) <= 2.0 and abs(sa.acq_doppler_hz - DOPPLER) <= 250.0, name


@pytest.mark.skipif(not O.ref_available(), reason="oracle/_ref not built")
def test_negative_acquisition_and_pfa_threshold():
    from gnss_sim_receiver_b200 import _hostapi as H
    x = e5a_stream(PRN, DOPPLER, DELAY, 40)
    sa, sb = _both(x, 19, 4e-3, 6000, max_dwells=3)  # PRN 19 is not in the air
    assert sa.last_event == 2 and sa.well_count == 3
    # pfa in the configuration wins over the threshold handed in (adapters/...caf.cc:134-154): the (1 - pfa)^(1 / cells) quantile of an
    # exponential distribution of rate vector_length (:272-289)
    a = H.E5aCafAcquisitionBlock(_conf(extra="Acquisition_5X.pfa=0.001\n"), "Acquisition_5X", 3)
    a.arm(5000, 250, 123.0)
    want = -np.log1p(-((1.0 - float(np.float32(0.001))) ** (1.0 / (12000 * 41)))) / 12000.0
    assert abs(a.status().threshold - want) <= 1e-6 * want
    a.close()


def test_errors_are_loud():
    from gnss_sim_receiver_b200 import _hostapi as H
    with pytest.raises(Exception, match="gr_complex"):
        H.E5aCafAcquisitionBlock(_conf().replace("item_type=gr_complex", "item_type=cshort"), "Acquisition_5X", 1)
    a = H.E5aCafAcquisitionBlock(_conf(caf=300), "Acquisition_5X", 1)
    with pytest.raises(Exception, match="CAF_window_hz"):
        a.arm(5000, 250, 0.0)  # a window below two Doppler steps divides by zero in the reference
    a.close()


def test_grid_pair_max_and_probe_against_the_kept_grid():
    """gb200_acq_grid_pair_max / gb200_acq_grid_probe on their own: against the grid read back from the device"""
    from gnss_sim_receiver_b200.acquisition import AcquisitionEngine
    rng = np.random.default_rng(7)
    n, nb = 6000, 9
    eng = AcquisitionEngine(fs_in=6e6, fft_size=n, consumed_samples=n, samples_per_code=float(n), samples_per_chip=1, use_cfar=False, max_prns=3, keep_grid=True)
    eng.set_doppler_bins(np.linspace(-1000, 1000, nb).astype(np.float32))
    for s in range(3):
        eng.set_local_code(s, (rng.choice(np.array([-1.0, 1.0]), n) + 1j * rng.choice(np.array([-1.0, 1.0]), n)).astype(np.complex64))
    eng.run((rng.standard_normal(n) + 1j * rng.standard_normal(n)).astype(np.complex64), slots=[0, 1, 2])
    g = [eng.read_grid(s) for s in range(3)]
    sa = rng.integers(0, 3, nb).astype(np.int32)
    sb = rng.integers(-1, 3, nb).astype(np.int32)
    v, i = eng.grid_pair_max(sa, sb)
    for b in range(nb):
        row = g[sa[b]][b] + g[sb[b]][b] if sb[b] >= 0 else g[sa[b]][b]
        assert i[b] == int(np.argmax(row)) and v[b] == row[i[b]]
    pts = (rng.integers(0, 3, 20), rng.integers(0, nb, 20), rng.integers(0, n, 20))
    got = eng.grid_probe(*pts)
    assert np.array_equal(got, np.array([g[s][b][k] for s, b, k in zip(*pts)], np.float32))
    with pytest.raises(Exception, match="outside"):
        eng.grid_probe([3], [0], [0])
    eng.close()
