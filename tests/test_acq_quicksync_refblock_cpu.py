"""The reference's OWN QuickSync acquisition block (pcps_quicksync_acquisition_cc.cc compiled unmodified into oracle/_ref, oracle/ref_acqqs.cc) on
synthetic GPS L1 C/A streams: what the referee of tests/test_acq_quicksync_gpu.py does on its own — vector items and their accounting, the
folded search and the resolution of the folding ambiguity, the dwell logic with and without the bit-transition flag.
"""
import numpy as np
import pytest

import _oracle as O
import _signals as S

pytestmark = pytest.mark.skipif(not O.ref_available(), reason="oracle/_ref not built")
FS = 4e6


def gps_stream(prn, doppler, delay_chips, n_ms, cn0=50.0, seed=3, fs=FS):
    comps = [dict(table=S.gps_l1_table(prn), f_chip=1.023e6, doppler=doppler, delay=delay_chips, cn0=cn0, phase=1.0)]
    return S.synth_if(int(n_ms * fs / 1000), fs, comps, seed=seed)


@pytest.mark.parametrize("folding,sampled_ms,delay_chips", [(0, 4, 311.0), (2, 2, 833.0), (4, 4, 977.0), (3, 4, 120.0)])
def test_folded_search_resolves_the_delay(folding, sampled_ms, delay_chips):
    prn, doppler = 7, 1830.0
    x = gps_stream(prn, doppler, delay_chips, 64)
    b = O.RefAcqQs(FS, folding, sampled_ms, 1, 5000, prn)
    b.arm(5000, 250, 0.02)
    items = b.work(x, 2)
    s = b.status()
    b.close()
    f = s.folding_factor
    assert f == (folding or 4) and s.item_length == 4000 * f and s.fft_size == 4000 // f and s.num_doppler_bins == 41
    assert s.sampled_ms == (sampled_ms if sampled_ms % f == 0 else (sampled_ms // f) * f)  # (3, 4) -> 3 ms
    # one item searched (state 1), then the positive state swallows the scheduler call's items (2 of them, :447-452)
    assert items == 3 and list(s.events[:s.n_events]) == [1] and (s.state, s.active, s.well_count) == (0, 0, 1)
    # the counters run in sampled_ms * samples_per_ms per item, whatever the item holds (:246,:281,:449)
    assert s.acq_samplestamp_samples == s.sampled_ms * 4000 and s.sample_counter == 3 * s.sampled_ms * 4000
    # the delay is found modulo the folded length and resolved among the folding_factor candidates by correlation with the unfolded replica
    assert abs(s.acq_delay_samples - delay_chips * 4000 / 1023) <= 2.0 and s.acq_doppler_hz == 1750.0 and s.acq_doppler_step == 250


def test_dwells_and_negative_acquisition():
    x = gps_stream(7, 1830.0, 311.0, 64)
    b = O.RefAcqQs(FS, 2, 2, 3, 5000, 19)  # PRN 19 is not in the air
    b.arm(5000, 250, 1.0)
    items = b.work(x, 1)
    s = b.status()
    b.close()
    assert list(s.events[:s.n_events]) == [2] and s.well_count == 3 and items == 3 + 1
    # bit-transition flag: max_dwells is 2 and the decision waits for the last dwell (:412-424)
    b = O.RefAcqQs(FS, 4, 4, 1, 5000, 7, bit_transition=True)
    b.arm(5000, 250, 0.02)
    items = b.work(x, 1)
    s = b.status()
    b.close()
    assert list(s.events[:s.n_events]) == [1] and s.well_count == 2 and items == 2 + 1 and s.acq_samplestamp_samples == 2 * 16000
