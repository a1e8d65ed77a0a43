"""Raw-sample conversions in front of the hot paths: the oracle against hand-worked known answers.

The rules are the reference's unpack_2bit_samples (gnuradio_blocks/unpack_2bit_samples.cc:121-208): four two's-complement
2-bit fields per byte, least significant first, each mapped s -> 2 s + 1.
"""
import numpy as np

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
