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


# ---------------------------------------------------------------------------------------------------------------- frozen acquisition cases
# Seeded acquisition scenarios shared by tests/golden/make_acq_golden.py (which freezes the oracle's answers into
# tests/golden/acq_vectors.json), tests/test_acq_golden_cpu.py (the oracle must keep reproducing them, fp32- and float64-FFT alike)
# and tests/test_acq_golden_gpu.py (the engine must reproduce them with no oracle in the loop).
ACQ_CASES = {
    # name: (fs, fft_size, consumed, doppler_max, doppler_step, signal, PRNs searched, PRNs present, item, bit_transition, use_cfar, dwells, seeds)
    "c1_gps_4msps": (4e6, 4000, 4000, 10000, 250, "1C", list(range(1, 9)), [1, 4, 6, 7], "gr_complex", False, True, 1, [1, 2, 3, 4, 5]),
    "c2_e1c_bit_transition_cshort": (4e6, 32000, 32000, 5000, 125, "1Cp", [3, 8, 11, 19], [3, 11], "cshort", True, False, 1, [1, 2]),
    "c3_gps_25msps": (25e6, 25000, 25000, 10000, 250, "1C", [2, 5, 9, 13], [5, 13], "gr_complex", False, True, 1, [1, 2, 3]),
    "c3_e1b_25msps": (25e6, 100000, 100000, 5000, 125, "1B", [1, 7], [7], "gr_complex", False, True, 1, [1]),
    "c4_l5i_4_dwells": (20e6, 20000, 20000, 10000, 250, "L5", [1, 3, 5], [3], "gr_complex", False, True, 4, [1, 2]),
    "c5_gps_int8_100msps": (100e6, 100000, 100000, 10000, 250, "1C", [4, 10], [10], "cbyte", False, True, 1, [1]),
}


def acq_case(name, seed):
    """One frozen scenario: dict with the engine / oracle configuration, the local replicas, the per-dwell inputs as the engine takes
    them (`items`) and as complex64 (`sigs`)."""
    fs, fft_size, consumed, dmax, dstep, signal, prns, present, item, bt, cfar, nd, _ = ACQ_CASES[name]
    rng = np.random.default_rng(1000 * seed + sum(map(ord, name)))
    spc = consumed // 2 if bt else consumed
    comps = []
    for p in present:
        base = dict(doppler=float(rng.uniform(-0.9 * dmax, 0.9 * dmax)), delay=float(rng.uniform(0, 1000)), cn0=float(rng.uniform(47, 51) + (4 if nd > 1 else 0)), phase=float(rng.uniform(0, 6.28)))
        if signal == "1C":
            comps.append(dict(table=gps_l1_table(p), f_chip=1.023e6, **base))
        elif signal in ("1B", "1Cp"):
            comps.append(dict(table=e1_table(p, "1B"), f_chip=2.046e6, weight=np.sqrt(0.5), symbols=rng.choice([-1.0, 1.0], 8), **base))
            comps.append(dict(table=e1_table(p, "1C"), f_chip=2.046e6, weight=-np.sqrt(0.5), symbols=rng.choice([-1.0, 1.0], 25), **base))
        else:
            comps.append(dict(table=l5_table(p), f_chip=10.23e6, weight=np.sqrt(0.5), symbols=rng.choice([-1.0, 1.0], 10), **base))
            comps.append(dict(table=1j * l5_table(p, q=True), f_chip=10.23e6, weight=np.sqrt(0.5), **base))
    x = synth_if(nd * consumed, fs, comps, seed=seed, f_carrier=1176.45e6 if signal == "L5" else 1575.42e6)
    if item == "gr_complex":
        items_all, sig_all, iq = x, x, 0
    else:
        scale, lim, t = (400.0 * np.sqrt(2.0), 32767, np.int16) if item == "cshort" else (25.0 * np.sqrt(2.0), 127, np.int8)
        q = np.empty(2 * len(x), dtype=t)
        q[0::2] = np.clip(np.rint(x.real * scale), -lim, lim).astype(t)
        q[1::2] = np.clip(np.rint(x.imag * scale), -lim, lim).astype(t)
        items_all, iq = q, (1 if item == "cshort" else 2)
        sig_all = (q[0::2].astype(np.float32) + 1j * q[1::2].astype(np.float32)).astype(np.complex64)
    if signal == "1C":
        locals_ = {p: gps_l1_sampled(p, fs) for p in prns}
        chip_rate = 1.023e6
    elif signal in ("1B", "1Cp"):
        locals_ = {p: e1_sampled(p, fs, "1C" if signal == "1Cp" else "1B") for p in prns}
        chip_rate = 1.023e6
    else:
        locals_ = {p: l5i_sampled(p, fs) for p in prns}
        chip_rate = 10.23e6
    k = 1 if item == "gr_complex" else 2
    return dict(fs=fs, fft_size=fft_size, consumed=consumed, dmax=dmax, dstep=dstep, spc=float(spc), spchip=int(np.ceil(fs / chip_rate)), prns=prns, present=present,
                locals_=locals_, iq_dtype=iq, bit_transition=bt, use_cfar=cfar, max_dwells=nd,
                items=[items_all[k * d * consumed:k * (d + 1) * consumed] for d in range(nd)], sigs=[sig_all[d * consumed:(d + 1) * consumed] for d in range(nd)])
