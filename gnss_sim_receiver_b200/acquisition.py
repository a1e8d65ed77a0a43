"""Python door onto the PCPS acquisition engine (C ABI gb200_acq_*), used by tests/ and bench.py.

Mirrors pcps_acquisition's configuration (Acq_Conf, src/algorithms/acquisition/libs/acq_conf.h:34-79) and the
AcquisitionInterface verbs (set_local_code, set_doppler_max/step/center) for ALL PRNs of one signal at once.
"""
import ctypes as C

import numpy as np

from . import _capi
from ._capi import IQ_CF32, IQ_CS16, IQ_CS8, AcqConfig, AcqResult, check

_IQ_NP = {IQ_CF32: np.complex64, IQ_CS16: np.int16, IQ_CS8: np.int8}


class AcquisitionEngine:
    def __init__(self, fs_in, fft_size, consumed_samples=None, samples_per_code=None, samples_per_chip=1, bit_transition=False, use_cfar=True,
                 max_dwells=1, max_prns=32, iq_dtype=IQ_CF32, device=0, keep_grid=False):
        self._lib = _capi.lib()
        consumed_samples = fft_size if consumed_samples is None else consumed_samples
        samples_per_code = float(consumed_samples) if samples_per_code is None else samples_per_code
        self.cfg = AcqConfig(device, iq_dtype, fft_size, consumed_samples, int(bit_transition), int(use_cfar), samples_per_chip, max_dwells, max_prns,
                             samples_per_code, float(fs_in))
        self._h = C.c_void_p()
        check(self._lib.gb200_acq_create(C.byref(self.cfg), C.byref(self._h)))
        if keep_grid:
            check(self._lib.gb200_acq_keep_grid(self._h, 1))
        self.iq_dtype = iq_dtype

    def close(self):
        if self._h:
            self._lib.gb200_acq_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_local_code(self, slot, code):
        code = np.ascontiguousarray(code, dtype=np.complex64)
        check(self._lib.gb200_acq_set_local_code(self._h, slot, code.ctypes.data, len(code)))

    def set_wipeoff_mode(self, exact):
        check(self._lib.gb200_acq_set_wipeoff_mode(self._h, int(exact)))

    def set_doppler_bins(self, freqs_hz):
        """explicit wipe-off frequencies, one per bin (a Doppler tile of a sweep, or the fine grid of a two-step search)"""
        f = np.ascontiguousarray(freqs_hz, dtype=np.float32)
        check(self._lib.gb200_acq_set_doppler_bins(self._h, f.ctypes.data_as(C.POINTER(C.c_float)), len(f)))

    def read_cells(self, n_slots):
        """(vmax, imax, sum) of every (slot, bin) of the last run, each [n_slots, n_bins]"""
        b = self.num_doppler_bins
        v, i, s = np.zeros((n_slots, b), np.float32), np.zeros((n_slots, b), np.int32), np.zeros((n_slots, b), np.float32)
        check(self._lib.gb200_acq_read_cells(self._h, n_slots, v.ctypes.data, i.ctypes.data, s.ctypes.data))
        return v, i, s

    def set_doppler_grid(self, doppler_max, doppler_step, doppler_center=0):
        check(self._lib.gb200_acq_set_doppler_grid(self._h, doppler_max, doppler_step, doppler_center))

    @property
    def num_doppler_bins(self):
        return self._lib.gb200_acq_num_doppler_bins(self._h)

    @property
    def neff(self):
        return self._lib.gb200_acq_neff(self._h)

    def run(self, iq, slots=None, dwell_index=0, n_slots=None):
        """One dwell of acquisition_core for the given code slots; returns a list of AcqResult."""
        iq = np.ascontiguousarray(iq)
        if self.iq_dtype == IQ_CF32:
            iq = iq.astype(np.complex64, copy=False)
            assert iq.size >= self.cfg.consumed_samples
        else:
            assert iq.dtype == _IQ_NP[self.iq_dtype] and iq.size >= 2 * self.cfg.consumed_samples
        if slots is None:
            n = n_slots or self.cfg.max_prns
            sl = None
        else:
            arr = np.ascontiguousarray(slots, dtype=np.int32)
            n, sl = len(arr), arr.ctypes.data
        out = (AcqResult * n)()
        check(self._lib.gb200_acq_run(self._h, iq.ctypes.data, dwell_index, sl, n, out))
        return out  # (a ctypes array of AcqResult: indexable and iterable like the list it used to be copied into)

    def run_device(self, dev_ptr, slots=None, dwell_index=0, n_slots=None):
        if slots is None:
            n = n_slots or self.cfg.max_prns
            sl = None
        else:
            arr = np.ascontiguousarray(slots, dtype=np.int32)
            n, sl = len(arr), arr.ctypes.data
        out = (AcqResult * n)()
        check(self._lib.gb200_acq_run_device(self._h, C.c_void_p(dev_ptr), dwell_index, sl, n, out))
        return out

    def replay(self, iters):
        ms = C.c_float()
        check(self._lib.gb200_acq_replay(self._h, iters, C.byref(ms)))
        return ms.value

    def threshold(self, pfa):
        return float(self._lib.gb200_acq_threshold(self._h, C.c_float(pfa)))

    def read_grid(self, slot):
        g = np.zeros((self.num_doppler_bins, self.neff), dtype=np.float32)
        check(self._lib.gb200_acq_read_grid(self._h, slot, g.ctypes.data))
        return g

    def fft(self, x, inverse=False):
        x = np.ascontiguousarray(x, dtype=np.complex64)
        assert len(x) == self.cfg.fft_size
        out = np.zeros_like(x)
        check(self._lib.gb200_acq_fft(self._h, x.ctypes.data, out.ctypes.data, int(inverse)))
        return out

    def set_input_folding(self, chunks, block_len):
        """QuickSync's folded input: every run reads block_len samples, wipes them off over the whole block and adds the first `chunks` pieces of
        fft_size samples into one transform input per Doppler bin (0 chunks: off)"""
        check(self._lib.gb200_acq_set_input_folding(self._h, int(chunks), int(block_len)))

    def folded_candidates(self, bin_, index, code, n_candidates):
        code = np.ascontiguousarray(code, dtype=np.complex64)
        out = np.zeros(n_candidates, np.complex64)
        check(self._lib.gb200_acq_folded_candidates(self._h, int(bin_), int(index), code.ctypes.data, len(code), int(n_candidates), out.ctypes.data))
        return out

    def grid_pair_max(self, slot_a, slot_b):
        """per Doppler bin: maximum / first index of kept grid row slot_a[bin] + row slot_b[bin] (slot_b < 0: row slot_a alone)"""
        a = np.ascontiguousarray(slot_a, dtype=np.int32)
        b = np.ascontiguousarray(slot_b, dtype=np.int32)
        assert len(a) == len(b) == self.num_doppler_bins
        v, i = np.zeros(len(a), np.float32), np.zeros(len(a), np.int32)
        check(self._lib.gb200_acq_grid_pair_max(self._h, a.ctypes.data, b.ctypes.data, v.ctypes.data, i.ctypes.data))
        return v, i

    def grid_probe(self, slot, bin_, index):
        s, b, i = (np.ascontiguousarray(x, dtype=np.int32) for x in (slot, bin_, index))
        out = np.zeros(len(s), np.float32)
        check(self._lib.gb200_acq_grid_probe(self._h, len(s), s.ctypes.data, b.ctypes.data, i.ctypes.data, out.ctypes.data))
        return out

    def spectrum_peak(self, x, rep, pad_factor=8):
        """estimate_Doppler's search (pcps_acquisition_fine_doppler_cc.cc:338-374): (index of the first maximum, value) of
        |FFT(x * rep repeated, zero-padded to pad_factor * len(x))|^2; len(x) a multiple of the engine's fft_size"""
        x = np.ascontiguousarray(x, dtype=np.complex64)
        rep = np.ascontiguousarray(rep, dtype=np.complex64)
        idx, val = C.c_int64(0), C.c_float(0.0)
        check(self._lib.gb200_acq_spectrum_peak(self._h, x.ctypes.data, len(x), rep.ctypes.data, len(rep), int(pad_factor), C.byref(idx), C.byref(val)))
        return int(idx.value), float(val.value)
