"""Seeded synthetic IF generator: sum of GNSS signals + complex white noise at baseband.

s(t) = sum_k A_k * c_k(floor((t - tau_k) * f_chip * (1 + fd_k / f_carrier)) mod L) * exp(j(2 pi fd_k t + phi_k)) + CN(0, 1)
A_k from CN0_k:  A = sqrt(10^(CN0/10) / fs)  (noise variance 1 per complex sample, i.e. N0 = 1/fs).
Restates the signal model of signal_generator_c (src/algorithms/signal_generator/gnuradio_blocks/signal_generator_c.cc:87-326)
without data bits (acquisition windows are one code period) unless nav_bits is given.
Quantised outputs are interleaved I,Q exactly as File_Signal_Source + Ishort_To_Complex / Ibyte_To_Complex read them.
"""
import numpy as np


def gen_if(codes, fs, n, sats, seed=1, f_chip=1.023e6, f_carrier=1575.42e6, noise=True, t0=0.0):
    """codes: dict prn -> +-1 chip table (one entry per chip... or per half chip with f_chip doubled accordingly);
    sats: list of dicts {prn, doppler, delay_chips, cn0, phase}.  Returns complex64[n]."""
    rng = np.random.default_rng(seed)
    t = t0 + np.arange(n, dtype=np.float64) / fs
    x = np.zeros(n, dtype=np.complex128)
    for s in sats:
        c = np.asarray(codes[s["prn"]], dtype=np.float64)
        L = len(c)
        amp = np.sqrt(10.0 ** (s["cn0"] / 10.0) / fs)
        chip = np.floor(t * f_chip * (1.0 + s["doppler"] / f_carrier) - s["delay_chips"]).astype(np.int64) % L
        x += amp * c[chip] * np.exp(1j * (2 * np.pi * s["doppler"] * t + s.get("phase", 0.0)))
    if noise:
        x += (rng.standard_normal(n) + 1j * rng.standard_normal(n)) / np.sqrt(2.0)
    return x.astype(np.complex64)


def quantise(x, kind):
    """complex64 -> interleaved int16 ('ishort', sigma ~ 400 LSB) or int8 ('ibyte', sigma ~ 25 LSB)."""
    if kind == "ishort":
        scale, lim, t = 400.0 * np.sqrt(2.0), 32767, np.int16
    elif kind == "ibyte":
        scale, lim, t = 25.0 * np.sqrt(2.0), 127, np.int8
    else:
        raise ValueError(kind)
    out = np.empty(2 * len(x), dtype=t)
    out[0::2] = np.clip(np.rint(x.real * scale), -lim, lim).astype(t)
    out[1::2] = np.clip(np.rint(x.imag * scale), -lim, lim).astype(t)
    return out


def dequantise(q):
    return (q[0::2].astype(np.float32) + 1j * q[1::2].astype(np.float32)).astype(np.complex64)
