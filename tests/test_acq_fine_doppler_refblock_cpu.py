"""The reference's OWN fine-Doppler acquisition block (pcps_acquisition_fine_doppler_cc.cc compiled unmodified into oracle/_ref,
oracle/ref_acqfd.cc) on synthetic streams: what the referee of tests/test_acq_fine_doppler_gpu.py does on its own, so that the GPU test's
"equal to the reference block" has a stated meaning — state machine and item accounting, the 12.5 Hz fine estimate, and the unsigned
coarse Doppler of :278 for satellites below zero Doppler.
"""
import numpy as np
import pytest

import _oracle as O
import _signals as S

pytestmark = pytest.mark.skipif(not O.ref_available(), reason="oracle/_ref not built")
FS = 4e6


def _run(prn, comps, dwells, noutput, n_ms=18, threshold=2.5, seed=3):
    x = S.synth_if(n_ms * 4000, FS, comps, seed=seed)
    b = O.RefAcqFd(FS, 5000, dwells, prn)
    b.arm(5000, 250, threshold)
    c = b.work(x, noutput)
    s = b.status()
    b.close()
    return c, s


@pytest.mark.parametrize("noutput", [4000, 1000])
def test_positive_acquisition_refines_the_doppler_to_the_padded_bin(noutput):
    prn, doppler, delay_chips = 7, 1830.0, 311.0
    c, s = _run(prn, [dict(table=S.gps_l1_table(prn), f_chip=1.023e6, doppler=doppler, delay=delay_chips, cn0=50.0, phase=1.0)], 2, noutput)
    # two dwells of 4000 items (state 1), the decision (state 2, nothing consumed), 32000 more items into the 10 ms buffer (state 3), then the
    # positive state consumes one scheduler call's noutput_items (state 4, :519-523)
    assert c == 8000 + 32000 + noutput and s.sample_counter == c
    assert list(s.events[:s.n_events]) == [1] and s.positive_acq == 1 and (s.state, s.active) == (0, 0)
    assert s.num_doppler_points == 40 and s.fft_size == 4000 and s.acq_doppler_step == 250
    assert s.acq_samplestamp_samples == 8000  # :279: the counter when compute_CAF ran
    assert abs(s.acq_delay_samples - delay_chips * 4000 / 1023) <= 2.0
    assert s.acq_doppler_hz == 1825.0  # fs / (8 * 10 * 4000) = 12.5 Hz bins; the coarse grid said 1750
    assert s.test_statistics > 2.5


def test_unsigned_coarse_doppler_below_zero():
    prn, doppler = 12, -2490.0
    c, s = _run(prn, [dict(table=S.gps_l1_table(prn), f_chip=1.023e6, doppler=doppler, delay=77.0, cn0=50.0, phase=0.3)], 2, 4000)
    assert list(s.events[:s.n_events]) == [1]
    # :278: static_cast<double>(index_doppler * d_doppler_step - d_config_doppler_max) with an unsigned d_doppler_step: bin 10 (-2500 Hz)
    # comes out as 2^32 - 2500, which the 1 kHz plausibility test of :393 then keeps
    assert s.acq_doppler_hz == float(2 ** 32 - 2500)


def test_negative_acquisition():
    c, s = _run(25, [dict(table=S.gps_l1_table(9), f_chip=1.023e6, doppler=1200.0, delay=500.0, cn0=48.0, phase=0.0)], 3, 4000, n_ms=12)
    assert list(s.events[:s.n_events]) == [2] and s.positive_acq == 0 and c == 3 * 4000 + 4000
    assert 1.0 <= s.test_statistics < 2.5
