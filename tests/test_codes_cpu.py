"""Replica generators of the host layer against the reference's own generators (oracle/_ref), bit-exact.

Galileo E1 (sinBOC(1,1) tracking table, sampled sinBOC / CBOC acquisition replicas with chip shift and the E1-C secondary
code), GPS L5 I/Q (two 13-stage registers), Galileo E5a (memory codes): every PRN, several sampling rates.
The Galileo memory codes embedded in the host library come from tools/make_code_tables.py; a digest of the table file is
checked here so that the committed file and the library agree even where oracle/_ref is absent.
"""
import ctypes as C
import hashlib
import os

import numpy as np
import pytest

import _oracle as O
from gnss_sim_receiver_b200 import _hostapi as H

needs_ref = pytest.mark.skipif(not O.ref_available(), reason="oracle/_ref not built")
HERE = os.path.dirname(os.path.abspath(__file__))


def gen(what, n_floats, sig=b"", prn=1, fs=0, chip_shift=0, cboc=0, secondary=0, fill=7.0):
    buf = np.full(n_floats, fill, np.float32)
    n = H.check(H.lib().gb200h_code_gen(what, buf.ctypes.data, sig, prn, fs, chip_shift, cboc, secondary))
    return buf, n


def vp(a):
    return a.ctypes.data_as(C.c_void_p)


def test_table_file_shape_and_digest():
    path = os.path.join(HERE, "..", "gnss_sim_receiver_b200", "data", "gnss_memory_codes.bin")
    raw = open(path, "rb").read()
    assert raw[:4] == b"GB2C" and len(raw) == 12 + 40 + 2 * 50 * 512 + 2 * 50 * 1279 + 47 * 13
    assert hashlib.sha256(raw).hexdigest() == open(os.path.join(HERE, "golden", "gnss_memory_codes.sha256")).read().strip()
    # balanced codes: a memory code has about as many +1 as -1 chips
    e1b = np.unpackbits(np.frombuffer(raw[52:52 + 50 * 512], np.uint8).reshape(50, 512), axis=1, bitorder="little")[:, :4092]
    assert np.all(np.abs(e1b.mean(axis=1) - 0.5) < 0.03)
    # the library serves exactly these chips
    code, n = gen(0, 8184, b"1B", 11)
    assert n == 8184 and np.array_equal(code[0::2], 1.0 - 2.0 * e1b[10]) and np.array_equal(code[1::2], -code[0::2])


@needs_ref
def test_galileo_e1_tables_and_replicas_bit_exact():
    r = O.ref()
    for sig in (b"1B", b"1C"):
        for prn in range(1, 51):
            want = np.zeros(8184, np.float32)
            r.ref_galileo_e1_code_gen_sinboc11_float(vp(want), 8184, sig, prn)
            got, _ = gen(0, 8184, sig, prn)
            assert np.array_equal(got, want), (sig, prn)
    for sig, prn, fs, shift, cboc, sec in [(b"1B", 11, 4000000, 0, 0, 0), (b"1B", 19, 4000000, 1500, 1, 0), (b"1C", 4, 6625000, 0, 0, 0),
                                           (b"1C", 30, 4000000, 7, 0, 1), (b"1B", 2, 2046000, 0, 0, 0), (b"1C", 50, 12276000, 100, 1, 0), (b"1B", 36, 25000000, 0, 1, 0)]:
        n = int(fs / (1.023e6 / 4092)) * (25 if (sig == b"1C" and sec) else 1)
        want = np.zeros(2 * n, np.float32)
        r.ref_galileo_e1_code_gen_complex_sampled(vp(want), n, sig, cboc, prn, fs, shift, sec)
        got, m = gen(1, 2 * n, sig, prn, fs, shift, cboc, sec)
        assert m == n and np.array_equal(got, want), (sig, prn, fs, shift, cboc, sec)


@needs_ref
def test_gps_l5_codes_bit_exact():
    r = O.ref()
    for prn in range(0, 52):  # 0 and 51 are outside the table: the all-zero-register code, like the reference
        for what, f in ((2, r.ref_gps_l5i_code_gen_float), (3, r.ref_gps_l5q_code_gen_float)):
            want = np.zeros(10230, np.float32)
            f(vp(want), prn)
            got, _ = gen(what, 10230, prn=prn)
            assert np.array_equal(got, want), (what, prn)
    for fs in (10230000, 20000000, 12500000, 25000000):
        n = int(fs / 1000)
        for what, f in ((4, r.ref_gps_l5i_code_gen_complex_sampled), (5, r.ref_gps_l5q_code_gen_complex_sampled)):
            for prn in (1, 17, 32, 50):
                want = np.zeros(2 * n, np.float32)
                f(vp(want), n, prn, fs)
                got, m = gen(what, 2 * n, prn=prn, fs=fs)
                assert m == n and np.array_equal(got, want), (what, prn, fs)


@needs_ref
def test_galileo_e5a_codes_bit_exact():
    r = O.ref()
    for sig in (b"5I", b"5Q", b"5X"):
        for prn in range(1, 51):
            want = np.zeros(2 * 10230, np.float32)
            r.ref_galileo_e5_a_code_gen_complex_primary(vp(want), prn, sig)
            got, _ = gen(6, 2 * 10230, sig, prn)
            assert np.array_equal(got, want), (sig, prn)
    for sig, prn, fs, shift in [(b"5X", 3, 20000000, 0), (b"5Q", 24, 10230000, 17), (b"5I", 50, 25000000, 9000), (b"5X", 11, 12500000, 1)]:
        n = int(fs / 1000)
        want = np.zeros(2 * n, np.float32)
        r.ref_galileo_e5_a_code_gen_complex_sampled(vp(want), n, prn, sig, fs, shift)
        got, m = gen(7, 2 * n, sig, prn, fs, shift)
        assert m == n and np.array_equal(got, want), (sig, prn, fs, shift)
    # outside the table: the destination is left as it was (early return)
    got, _ = gen(6, 2 * 10230, b"5X", 51, fill=7.0)
    assert np.all(got == 7.0)
