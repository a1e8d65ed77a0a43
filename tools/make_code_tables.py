"""Writes gnss_sim_receiver_b200/data/gnss_memory_codes.bin: the Galileo memory codes the product's replica generators read.

Galileo E1-B / E1-C (4092 chips) and E5a-I / E5a-Q (10230 chips) primary codes are tabulated in the Galileo OS SIS ICD
(Annex C); the reference carries them as hex strings (src/core/system_parameters/Galileo_E1.h:44-760, Galileo_E5a.h) and
reads them in galileo_e1_code_gen_int / galileo_e5_a_code_gen_complex_primary.  This script takes the chips from the
OUTPUT of the reference's own generators compiled under oracle/_ref (dev container only) and packs them one bit per chip
(bit = 1 for a chip of -1), PRN 1..50 each.  The file is embedded into libgnss_b200_host.so (b200_gnss_codes.cc, .incbin).

Layout (little-endian): magic "GB2C", u32 version=1, u32 n_tables=4, then per table: u32 n_prn, u32 n_chips, followed
by the tables in the same order, each PRN padded to whole bytes, chips LSB-first.  Order: E1B, E1C, E5aI, E5aQ, then the
E5a-Q secondary codes CS100 (100 chips per PRN, bit = the ICD's binary digit; OS SIS ICD table 20), which no generator
function of the reference outputs: those strings (47: the reference's array leaves rows 48..50 empty) are read from Galileo_E5a.h:3585-3633 by this script.
"""
import ctypes as C
import os
import struct
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "tests"))
import _oracle as O  # noqa: E402


def main():
    r = O.ref()
    tables = []
    for sig in (b"1B", b"1C"):
        t = np.zeros((50, 4092), np.uint8)
        for prn in range(1, 51):
            buf = np.zeros(8184, np.float32)
            r.ref_galileo_e1_code_gen_sinboc11_float(buf.ctypes.data_as(C.c_void_p), 8184, sig, prn)
            assert np.array_equal(buf[1::2], -buf[0::2]) and set(np.unique(buf)) == {-1.0, 1.0}
            t[prn - 1] = buf[0::2] < 0
        tables.append(t)
    ti, tq = np.zeros((50, 10230), np.uint8), np.zeros((50, 10230), np.uint8)
    for prn in range(1, 51):
        buf = np.zeros(2 * 10230, np.float32)
        r.ref_galileo_e5_a_code_gen_complex_primary(buf.ctypes.data_as(C.c_void_p), prn, b"5X")
        assert set(np.unique(buf)) == {-1.0, 1.0}
        ti[prn - 1], tq[prn - 1] = buf[0::2] < 0, buf[1::2] < 0
    tables += [ti, tq]
    import re
    hdr = open("/root/reference/src/core/system_parameters/Galileo_E5a.h").read()
    body = hdr[hdr.index("GALILEO_E5A_Q_SECONDARY_CODE[GALILEO_E5A_NUMBER_OF_CODES]"):]
    strings = re.findall(r'"([01]{100})"', body)
    assert len(strings) == 47  # the reference's table initialises 47 of its 50 rows; the table here holds those 47
    tables.append(np.array([[int(ch) for ch in st] for st in strings], np.uint8))
    out = os.path.join(os.path.dirname(__file__), "..", "gnss_sim_receiver_b200", "data", "gnss_memory_codes.bin")
    with open(out, "wb") as f:
        f.write(b"GB2C" + struct.pack("<II", 1, len(tables)))
        for t in tables:
            f.write(struct.pack("<II", *t.shape))
        for t in tables:
            f.write(np.packbits(t, axis=1, bitorder="little").tobytes())
    print(out, os.path.getsize(out), "bytes")


if __name__ == "__main__":
    main()
