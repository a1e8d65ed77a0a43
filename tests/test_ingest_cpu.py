"""Raw-sample conversions in front of the hot paths: the oracle against hand-worked known answers.

The rules are the reference's unpack_2bit_samples (gnuradio_blocks/unpack_2bit_samples.cc:121-208): four two's-complement
2-bit fields per byte, least significant first, each mapped s -> 2 s + 1.
"""
import numpy as np
import pytest

import _oracle as O


def test_two_bit_known_answers():
    # 0xE4 = 11 10 01 00 : fields (lsb first) 0, 1, -2, -1 -> 1, 3, -3, -1
    b = np.array([0xE4], np.uint8)
    assert O.unpack_2bit(b).tolist() == [1, 3, -3, -1]
    assert O.unpack_2bit(b, big_endian_bytes=True).tolist() == [-1, -3, 3, 1]
    assert O.unpack_2bit(b, reverse_interleaving=True).tolist() == [3, 1, -1, -3]
    assert O.unpack_2bit(b, big_endian_bytes=True, reverse_interleaving=True).tolist() == [-3, -1, 1, 3]
    # 0x1B = 00 01 10 11 : fields -1, -2, 1, 0 -> -1, -3, 3, 1
    assert O.unpack_2bit(np.array([0x1B], np.uint8)).tolist() == [-1, -3, 3, 1]


def test_two_bit_item_byte_order():
    raw = np.array([0xE4, 0x1B, 0x00, 0xFF], np.uint8)
    plain = O.unpack_2bit(raw).reshape(4, 4)
    # items of one byte are never reversed
    assert np.array_equal(O.unpack_2bit(raw, item_size=1, big_endian_items=True).reshape(4, 4), plain)
    # big-endian 2-byte items on this little-endian machine: bytes swap in pairs
    assert np.array_equal(O.unpack_2bit(raw, item_size=2, big_endian_items=True).reshape(4, 4), plain[[1, 0, 3, 2]])
    assert np.array_equal(O.unpack_2bit(raw, item_size=4, big_endian_items=True).reshape(4, 4), plain[[3, 2, 1, 0]])
    assert np.array_equal(O.unpack_2bit(raw, item_size=4, big_endian_items=False).reshape(4, 4), plain)


def test_two_bit_every_byte_value():
    raw = np.arange(256, dtype=np.uint8)
    out = O.unpack_2bit(raw).reshape(256, 4).astype(np.int32)
    for k in range(4):
        f = (raw.astype(np.int32) >> (2 * k)) & 3
        s = np.where(f >= 2, f - 4, f)
        assert np.array_equal(out[:, k], 2 * s + 1)
    assert set(np.unique(out)) == {-3, -1, 1, 3}


def test_items_to_complex():
    rng = np.random.default_rng(5)
    s16 = rng.integers(-32768, 32767, 64, dtype=np.int16)
    s8 = rng.integers(-128, 127, 64, dtype=np.int8)
    assert np.array_equal(O.items_to_complex(s16, 1), (s16[0::2] + 1j * s16[1::2]).astype(np.complex64))
    assert np.array_equal(O.items_to_complex(s8, 2), (s8[0::2] + 1j * s8[1::2]).astype(np.complex64))
    assert np.array_equal(O.items_to_complex(s16, 3), s16.astype(np.complex64))
    assert np.array_equal(O.items_to_complex(s8, 4), s8.astype(np.complex64))
    c = (rng.standard_normal(16) + 1j * rng.standard_normal(16)).astype(np.complex64)
    assert np.array_equal(O.items_to_complex(c, 0), c)


def test_labsat_known_answers():
    """labsat23_source::decode_samples_one_channel (labsat23_source.cc:364-433), hand-worked words."""
    # 2 bits per complex sample: 0xB400 = 10 11 01 00 | 00.. : (I,Q) pairs msb first -> (+1,-1) (+1,+1) (-1,+1) (-1,-1) then (-1,-1) x4
    got = O.labsat_one_channel(np.array([0xB400], np.uint16), 2)
    assert got.tolist() == [1 - 1j, 1 + 1j, -1 + 1j, -1 - 1j] + [-1 - 1j] * 4
    # 4 bits per complex sample, nibble = (I_hi, Q_hi, I_lo, Q_lo): 0x0 -> (1,1); 0xF -> (-1,-1); 0xA = 1010 -> I=(1,1)=-1, Q=(0,0)=+1;
    # 0x6 = 0110 -> I=(0,1)=+2, Q=(1,0)=-2
    got = O.labsat_one_channel(np.array([0x0FA6], np.uint16), 4)
    assert got.tolist() == [1 + 1j, -1 - 1j, -1 + 1j, 2 - 2j]
    # every word, against an independent numpy formulation
    w = np.arange(65536, dtype=np.uint16).astype(np.int64)
    two = O.labsat_one_channel(w.astype(np.uint16), 2).reshape(65536, 8)
    four = O.labsat_one_channel(w.astype(np.uint16), 4).reshape(65536, 4)
    lvl = np.array([[1.0, 2.0], [-2.0, -1.0]])
    for k in range(8):
        assert np.array_equal(two[:, k], (2 * ((w >> (15 - 2 * k)) & 1) - 1) + 1j * (2 * ((w >> (14 - 2 * k)) & 1) - 1))
    for k in range(4):
        re = lvl[(w >> (15 - 4 * k)) & 1, (w >> (13 - 4 * k)) & 1]
        im = lvl[(w >> (14 - 4 * k)) & 1, (w >> (12 - 4 * k)) & 1]
        assert np.array_equal(four[:, k], (re + 1j * im).astype(np.complex64))


@pytest.mark.skipif(not O.ref_available(), reason="oracle/_ref not built")
@pytest.mark.parametrize("item_size", [1, 2, 4])
@pytest.mark.parametrize("big_endian_items", [False, True])
@pytest.mark.parametrize("big_endian_bytes", [False, True])
@pytest.mark.parametrize("reverse_interleaving", [False, True])
def test_two_bit_oracle_is_pinned_to_the_reference_block(item_size, big_endian_items, big_endian_bytes, reverse_interleaving):
    """The oracle's 2-bit unpacker against the reference's OWN unpack_2bit_samples::work (gnuradio_blocks/unpack_2bit_samples.cc:121-208,
    compiled unmodified into oracle/_ref against stub declarations of gr::sync_interpolator): every byte value, every option combination,
    bit for bit — the GPU ingest kernel is tested bit-exact against this oracle in tests/test_ingest_gpu.py."""
    rng = np.random.default_rng(item_size + 2 * big_endian_items + 4 * big_endian_bytes + 8 * reverse_interleaving)
    raw = np.concatenate([np.arange(256, dtype=np.uint8), rng.integers(0, 256, 4096 - 256, dtype=np.uint8)])
    want = np.zeros(4 * len(raw), np.int8)
    f = O.ref().ref_unpack_2bit
    f.restype = O.C.c_int
    f.argtypes = [O.C.c_void_p, O.C.c_long, O.C.c_int, O.C.c_int, O.C.c_int, O.C.c_int, O.C.c_void_p]
    assert f(raw.ctypes.data, len(raw), item_size, int(big_endian_items), int(big_endian_bytes), int(reverse_interleaving), want.ctypes.data) == 4 * len(raw)
    got = O.unpack_2bit(raw, item_size=item_size, big_endian_items=big_endian_items, big_endian_bytes=big_endian_bytes, reverse_interleaving=reverse_interleaving)
    assert np.array_equal(got, want)
