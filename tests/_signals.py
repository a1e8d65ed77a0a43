"""Test-side signal tables and synthetic IF for the four in-scope signals (GPS L1 C/A, Galileo E1 B/C, GPS L5 I/Q, Galileo E5a I/Q).

Code tables and sampled replicas come from the host library's generators (`gb200h_code_gen`, CPU-only code), which
tests/test_codes_cpu.py proves bit-exact against the reference's own generators compiled into oracle/_ref; the IF model is
SURVEY §8d's (sum of satellites + CN(0,1) noise), with per-code-period sign sequences for data symbols / secondary codes.
"""
import numpy as np

from gnss_sim_receiver_b200 import _hostapi as H

# what-codes of gb200h_code_gen (host/b200_host_capi.cc)
_E1_TABLE, _E1_SAMPLED, _L5I, _L5Q, _L5I_S, _L5Q_S, _E5A_PRIMARY, _E5A_SAMPLED = range(8)


def _gen(what, n_floats, sig=b"", prn=1, fs=0, chip_shift=0, cboc=0, secondary=0):
    buf = np.zeros(n_floats, np.float32)
    n = H.check(H.lib().gb200h_code_gen(what, buf.ctypes.data, sig, prn, int(fs), chip_shift, cboc, secondary))
    return buf, n


def gps_l1_table(prn):
    buf = np.zeros(1023, np.float32)
    H.check(H.lib().gb200h_gps_l1_ca_code_gen_float(buf.ctypes.data, prn, 0))
    return buf


def gps_l1_sampled(prn, fs):
    n = int(fs) // 1000
    buf = np.zeros(2 * n, np.float32)
    H.check(H.lib().gb200h_gps_l1_ca_code_gen_complex_sampled(buf.ctypes.data, n, prn, int(fs), 0))
    return buf.view(np.complex64).copy()


def e1_table(prn, sig="1B"):
    """8184 entries (2 per chip, sinBOC(1,1)): galileo_e1_code_gen_sinboc11_float"""
    return _gen(_E1_TABLE, 8184, sig.encode(), prn)[0]


def e1_sampled(prn, fs, sig="1B", cboc=False):
    n = int(int(fs) / (1.023e6 / 4092))
    buf, m = _gen(_E1_SAMPLED, 2 * n, sig.encode(), prn, fs, 0, int(cboc), 0)
    assert m == n
    return buf.view(np.complex64).copy()


def l5_table(prn, q=False):
    return _gen(_L5Q if q else _L5I, 10230, b"", prn)[0]


def l5i_sampled(prn, fs):
    n = int(fs) // 1000
    buf, m = _gen(_L5I_S, 2 * n, b"", prn, fs)
    assert m == n
    return buf.view(np.complex64).copy()


def e5a_primary(prn, sig="5X"):
    """10230 complex entries: "5I" -> (I, 0), "5Q" -> (Q, 0), "5X" -> (I, Q)"""
    return _gen(_E5A_PRIMARY, 2 * 10230, sig.encode(), prn)[0].view(np.complex64).copy()


def e5a_sampled(prn, fs, sig="5I"):
    n = int(fs) // 1000
    buf, m = _gen(_E5A_SAMPLED, 2 * n, sig.encode(), prn, fs)
    assert m == n
    return buf.view(np.complex64).copy()


def synth_if(n, fs, comps, seed=1, f_carrier=1575.42e6, noise=True, t0=0.0):
    """complex64[n]:  sum over components of  A * sym[period] * table[floor(t f_chip (1 + fd / f_carrier) - delay) mod L] * exp(j(2 pi fd t + phase)) + CN(0, 1).

    comps: dicts with table (real or complex, L entries), f_chip (table entries per second), doppler, delay (table entries), cn0 (dB-Hz,
    N0 = 1/fs), optional phase, optional `symbols` (+-1 per code period, repeated cyclically), optional `weight` (amplitude share)."""
    rng = np.random.default_rng(seed)
    t = t0 + np.arange(n, dtype=np.float64) / fs
    x = np.zeros(n, dtype=np.complex128)
    for c in comps:
        tab = np.asarray(c["table"])
        L = len(tab)
        amp = np.sqrt(10.0 ** (c["cn0"] / 10.0) / fs) * c.get("weight", 1.0)
        k = np.floor(t * c["f_chip"] * (1.0 + c["doppler"] / f_carrier) - c["delay"]).astype(np.int64)
        v = tab[k % L].astype(np.complex128)
        if "symbols" in c:
            sym = np.asarray(c["symbols"], dtype=np.float64)
            v = v * sym[(k // L) % len(sym)]
        x += amp * v * np.exp(1j * (2 * np.pi * c["doppler"] * t + c.get("phase", 0.0)))
    if noise:
        x += (rng.standard_normal(n) + 1j * rng.standard_normal(n)) / np.sqrt(2.0)
    return x.astype(np.complex64)
