"""The reference's OWN Galileo E5a non-coherent I/Q acquisition block with CAF filtering (galileo_e5a_noncoherent_iq_acquisition_caf_cc.cc compiled
unmodified into oracle/_ref, oracle/ref_acqcaf.cc) on synthetic E5a streams: what the referee of tests/test_acq_e5a_caf_gpu.py does on its own —
buffering state machine and item accounting, code phase and Doppler of a plain 1 ms search, the dwell logic.
"""
import numpy as np
import pytest

import _oracle as O
import _signals as S

pytestmark = pytest.mark.skipif(not O.ref_available(), reason="oracle/_ref not built")
FS = 12e6


def e5a_stream(prn, doppler, delay_chips, n_ms, cn0=53.0, seed=1, sym_i=(1,), sym_q=(1,), fs=FS):
    """E5a-I + j E5a-Q of one satellite (half the power each), each with its own sign per code period"""
    comps = [dict(table=S.e5a_primary(prn, "5I"), f_chip=10.23e6, doppler=doppler, delay=delay_chips, cn0=cn0 - 3, phase=0.7, symbols=sym_i),
             dict(table=1j * S.e5a_primary(prn, "5Q"), f_chip=10.23e6, doppler=doppler, delay=delay_chips, cn0=cn0 - 3, phase=0.7, symbols=sym_q)]
    return S.synth_if(int(n_ms * fs / 1000), fs, comps, seed=seed, f_carrier=1176.45e6)


@pytest.mark.parametrize("chunk", [5000, 12000, 36000])
def test_plain_search_and_item_accounting(chunk):
    prn, doppler, delay = 11, 1830.0, 3333.0
    x = e5a_stream(prn, doppler, delay, 12)
    b = O.RefAcqCaf(FS, 1, 1, 5000, prn)
    b.arm(5000, 250, 1e-4)
    c = b.work(x, chunk)
    s = b.status()
    b.close()
    # state 1 fills 12000 items (the search starts with the call after the one that filled the buffer, consuming nothing), state 3
    # consumes one more scheduler call (:722-727)
    assert c == 12000 + chunk and s.sample_counter == c and s.acq_samplestamp_samples == 12000
    assert list(s.events[:s.n_events]) == [1] and (s.state, s.active) == (0, 0)
    assert (s.num_doppler_bins, s.fft_size, s.well_count, s.acq_doppler_step) == (41, 12000, 1, 250)
    assert abs(s.acq_delay_samples - delay * 12000 / 10230) <= 2.0 and s.acq_doppler_hz == 1750.0
    assert s.test_statistics > 4e-3 and abs(s.input_power - 1.0) < 0.05


def test_negative_acquisition_takes_every_dwell():
    x = e5a_stream(11, 1830.0, 3333.0, 40)
    b = O.RefAcqCaf(FS, 1, 3, 5000, 19)  # PRN 19 is not in the air
    b.arm(5000, 250, 4e-3)  # (a noise-only search reaches about 2e-3: the statistic is the peak over fft_size^4 and the mean input power)
    c = b.work(x, 6000)
    s = b.status()
    b.close()
    assert list(s.events[:s.n_events]) == [2] and s.well_count == 3 and c == 3 * 12000 + 6000
