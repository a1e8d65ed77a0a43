"""Raw-sample ingest through the C ABI: gb200_trk_push_raw against the oracle, bit-exact.

Every variant of the 2-bit packed stream (sample_type x byte order x item order / size) and every file item type, into every
ring type, with pushes that cross the end of the ring; then a correlation over a window that straddles the wrap (served from
the mirrored stretch) against the oracle multicorrelator fed with the oracle-converted samples.
"""
import itertools

import numpy as np
import pytest

import _oracle as O
from gnss_sim_receiver_b200 import tracking as T
from gnss_sim_receiver_b200._capi import IQ_CF32, IQ_CS16, IQ_CS8, GB200Error

pytestmark = pytest.mark.gpu

RING, BLOCK = 1 << 15, 4096


def ring_view(eng, first, n):
    got = eng.read_iq(first, n)
    return got if got.dtype == np.complex64 else (got[:, 0].astype(np.float32) + 1j * got[:, 1].astype(np.float32)).astype(np.complex64)


def expect_2bit(raw, sample_type, item_size, be_items, be_bytes):
    r = O.unpack_2bit(raw, item_size, be_items, be_bytes, sample_type == T.TWOBIT_QI).astype(np.float32)
    if sample_type == T.TWOBIT_REAL:
        return r.astype(np.complex64)
    return (r[0::2] + 1j * r[1::2]).astype(np.complex64)


@pytest.mark.parametrize("ring_dtype", [IQ_CF32, IQ_CS16, IQ_CS8])
def test_two_bit_all_variants_bit_exact(ring_dtype):
    rng = np.random.default_rng(11 + ring_dtype)
    eng = T.TrackingEngine(max_channels=2, max_code_len=1023, iq_dtype=ring_dtype, ring_samples=RING, max_block=BLOCK)
    for sample_type, item_size, be_items, be_bytes in itertools.product((T.TWOBIT_REAL, T.TWOBIT_IQ, T.TWOBIT_QI), (1, 2, 4, 8), (False, True), (False, True)):
        n_bytes = int(rng.integers(1, 700)) * 8  # whole items of every size; ragged against the ring on purpose
        raw = rng.integers(0, 256, n_bytes, dtype=np.uint8)
        first, n = eng.push_raw(raw, T.SRC_2BIT, sample_type, item_size, be_items, be_bytes)
        want = expect_2bit(raw, sample_type, item_size, be_items, be_bytes)
        assert n == want.size == (4 if sample_type == T.TWOBIT_REAL else 2) * n_bytes
        assert np.array_equal(ring_view(eng, first, n), want), (sample_type, item_size, be_items, be_bytes)
    assert eng.head > 2 * RING  # the pushes wrapped the ring more than once
    eng.close()


@pytest.mark.parametrize("ring_dtype", [IQ_CF32, IQ_CS16, IQ_CS8])
def test_item_types_bit_exact(ring_dtype):
    rng = np.random.default_rng(3)
    eng = T.TrackingEngine(max_channels=2, max_code_len=1023, iq_dtype=ring_dtype, ring_samples=RING, max_block=BLOCK)
    s8 = rng.integers(-128, 128, 20002, dtype=np.int8)
    s16 = rng.integers(-32768, 32768, 20002, dtype=np.int16)
    cf = (rng.standard_normal(9001) + 1j * rng.standard_normal(9001)).astype(np.complex64)
    cases = [(T.SRC_CS8, s8), (T.SRC_S8, s8)]
    if ring_dtype != IQ_CS8:
        cases += [(T.SRC_CS16, s16), (T.SRC_S16, s16)]
    if ring_dtype == IQ_CF32:
        cases += [(T.SRC_CF32, cf)]
    for _ in range(3):  # three rounds: crosses the end of the ring
        for kind, raw in cases:
            first, n = eng.push_raw(raw, kind)
            want = O.items_to_complex(raw, kind)
            assert n == want.size
            assert np.array_equal(ring_view(eng, first, n), want), kind
    # refused: a source wider than the ring type, a ragged byte count, an unknown item size
    if ring_dtype == IQ_CS8:
        with pytest.raises(GB200Error):
            eng.push_raw(s16, T.SRC_CS16)
    if ring_dtype != IQ_CF32:
        with pytest.raises(GB200Error):
            eng.push_raw(cf, T.SRC_CF32)
    with pytest.raises(GB200Error):
        eng.push_raw(s8[:3], T.SRC_CS8)
    with pytest.raises(GB200Error):
        eng.push_raw(s8[:6], T.SRC_2BIT, T.TWOBIT_IQ, 3)
    with pytest.raises(GB200Error):
        eng.push_raw(s8[:6], T.SRC_2BIT, T.TWOBIT_IQ, 4)  # 6 bytes: not whole 4-byte items
    first, n = eng.push_raw(s8[:0], T.SRC_CS8)  # empty push: nothing appended
    assert n == 0 and first == eng.head
    eng.close()


@pytest.mark.parametrize("ring_dtype", [IQ_CF32, IQ_CS16, IQ_CS8])
def test_labsat_single_channel_bit_exact(ring_dtype):
    """LabSat 2/3 single-channel payload (1+1 and 2+2 bits per complex sample), every 16-bit word value, ragged pushes over the wrap."""
    rng = np.random.default_rng(23 + ring_dtype)
    eng = T.TrackingEngine(max_channels=2, max_code_len=1023, iq_dtype=ring_dtype, ring_samples=RING, max_block=BLOCK)
    every = np.arange(65536, dtype=np.uint16)
    for kind, bits, per in ((T.SRC_LABSAT2, 2, 8), (T.SRC_LABSAT4, 4, 4)):
        for lo in range(0, 65536, 3000):  # all 65536 words in pushes that fit the ring
            raw = every[lo:lo + 3000]
            first, n = eng.push_raw(raw, kind)
            want = O.labsat_one_channel(raw, bits)
            assert n == want.size == per * raw.size
            assert np.array_equal(ring_view(eng, first, n), want), (kind, lo)
        for _ in range(6):
            raw = rng.integers(0, 65536, int(rng.integers(1, 3500)), dtype=np.uint16)
            first, n = eng.push_raw(raw, kind)
            assert np.array_equal(ring_view(eng, first, n), O.labsat_one_channel(raw, bits)), kind
    assert eng.head > 2 * RING
    with pytest.raises(GB200Error):
        eng.push_raw(np.zeros(3, np.uint8), T.SRC_LABSAT2)  # not whole int16 words
    eng.close()


def test_correlation_over_ingested_window_across_the_wrap():
    """2-bit "iq" stream -> byte ring -> multicorrelator, window straddling the end of the ring (mirror written by the kernel)."""
    rng = np.random.default_rng(8)
    eng = T.TrackingEngine(max_channels=1, max_code_len=1023, iq_dtype=IQ_CS8, ring_samples=RING, max_block=BLOCK)
    code = rng.choice([-1.0, 1.0], 1023).astype(np.float32)
    eng.set_code(0, code)
    raw1 = rng.integers(0, 256, (RING - 1500) // 2, dtype=np.uint8)
    raw2 = rng.integers(0, 256, 3000, dtype=np.uint8)
    f1, n1 = eng.push_raw(raw1, T.SRC_2BIT, T.TWOBIT_IQ)
    f2, n2 = eng.push_raw(raw2, T.SRC_2BIT, T.TWOBIT_IQ)
    assert f2 == n1 and f2 + n2 > RING
    sig = np.concatenate([expect_2bit(raw1, T.TWOBIT_IQ, 1, False, False), expect_2bit(raw2, T.TWOBIT_IQ, 1, False, False)])
    start, n = RING - 2000, 4000
    shifts = np.array([-0.5, 0.0, 0.5], np.float32)
    cm = eng.make_cmds([0], [start], [n], shifts[None, :], [0.3], [0.002], [0.25], [0.2557])
    got = eng.correlate(cm)[0]
    want = O.mc_real_codes(sig[start:start + n], code, shifts, 0.3, 0.002, 0.25, 0.2557)
    assert np.max(np.abs(got - want)) / np.max(np.abs(want)) < 2e-5
    eng.close()
