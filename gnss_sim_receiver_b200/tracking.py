"""Python door onto the tracking multicorrelator engine (C ABI gb200_trk_*), used by tests/ and bench.py.

The product host layer is the C++ under gnss_sim_receiver_b200/host/; this module only marshals numpy arrays
into the C structures.  Argument names follow Cpu_Multicorrelator_Real_Codes
(src/algorithms/tracking/libs/cpu_multicorrelator_real_codes.h:37-61 of the reference).
"""
import ctypes as C

import numpy as np

from . import _capi
from ._capi import GB200_MAX_TAPS, IQ_CF32, IQ_CS16, IQ_CS8, IngestFormat, TrkCmd, TrkConfig, check

SRC_CF32, SRC_CS16, SRC_CS8, SRC_S16, SRC_S8, SRC_2BIT, SRC_LABSAT2, SRC_LABSAT4 = range(8)  # GB200_SRC_*
TWOBIT_REAL, TWOBIT_IQ, TWOBIT_QI = range(3)  # GB200_2BIT_*

_CMD_DTYPE = np.dtype([("ch", "<i4"), ("n_samples", "<i4"), ("n_taps", "<i4"), ("flags", "<i4"), ("sample_offset", "<u8"),
                       ("rem_carrier_phase_rad", "<f4"), ("phase_step_rad", "<f4"), ("phase_rate_step_rad", "<f4"),
                       ("rem_code_phase_chips", "<f4"), ("code_phase_step_chips", "<f4"), ("code_phase_rate_step_chips", "<f4"),
                       ("shifts_chips", "<f4", (GB200_MAX_TAPS,))])
assert _CMD_DTYPE.itemsize == C.sizeof(TrkCmd)
_IQ_NP = {IQ_CF32: (np.complex64, 8), IQ_CS16: (np.int16, 4), IQ_CS8: (np.int8, 2)}


class TrackingEngine:
    """All channels of one GPU: shared device IQ ring + per-channel code tables."""

    def __init__(self, device=0, max_channels=64, max_code_len=10230, iq_dtype=IQ_CF32, ring_samples=1 << 20, max_block=1 << 17, max_cmds=None):
        self._lib = _capi.lib()
        self.cfg = TrkConfig(device, max_channels, max_code_len, iq_dtype, ring_samples, max_block, max_cmds or max_channels)
        self._h = C.c_void_p()
        self._host_bufs = []
        check(self._lib.gb200_trk_create(C.byref(self.cfg), C.byref(self._h)))
        self.iq_dtype = iq_dtype
        self._cmds = None
        self._n = 0

    def close(self):
        if self._h:
            self._lib.gb200_trk_destroy(self._h)
            self._h = C.c_void_p()
            # (arrays handed out by host_outputs() must not be used after close)
            for p in self._host_bufs:
                self._lib.gb200_host_free(p)
            self._host_bufs = []

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # set_local_code_and_taps (code part)
    def set_code(self, ch, code):
        code = np.ascontiguousarray(code)
        if np.iscomplexobj(code):
            code = code.astype(np.complex64)
            check(self._lib.gb200_trk_set_code_complex(self._h, ch, code.ctypes.data, len(code)))
        else:
            code = code.astype(np.float32)
            check(self._lib.gb200_trk_set_code(self._h, ch, code.ctypes.data, len(code)))

    # set_input_output_vectors (input part): append samples to the ring, returns the absolute index of the first one
    def push_iq(self, iq):
        np_t, eb = _IQ_NP[self.iq_dtype]
        iq = np.ascontiguousarray(iq)
        if self.iq_dtype == IQ_CF32:
            iq = iq.astype(np.complex64, copy=False)
            n = iq.size
        else:
            assert iq.dtype == np_t, f"ring holds interleaved {np_t.__name__} I,Q"
            n = iq.size // 2
        first = C.c_uint64()
        check(self._lib.gb200_trk_push_iq(self._h, iq.ctypes.data, n, C.byref(first)))
        return first.value

    # File_Signal_Source item types / Two_Bit_Packed_File_Signal_Source: raw bytes cross PCIe, converted on the device
    def push_raw(self, raw, kind, sample_type=0, item_size=1, big_endian_items=False, big_endian_bytes=False):
        """Append the samples held in `raw` (any contiguous numpy array, taken as bytes).  Returns (first_index, n_samples)."""
        raw = np.ascontiguousarray(raw)
        fmt = IngestFormat(int(kind), int(sample_type), int(item_size), int(bool(big_endian_items)), int(bool(big_endian_bytes)))
        first, n = C.c_uint64(), C.c_int64()
        check(self._lib.gb200_trk_push_raw(self._h, raw.ctypes.data, raw.nbytes, C.byref(fmt), C.byref(first), C.byref(n)))
        return first.value, n.value

    def read_iq(self, abs_index, n_samples):
        """Samples [abs_index, abs_index + n_samples) of the ring, in the ring's own item type (complex64 or (n, 2) integers)."""
        np_t, eb = _IQ_NP[self.iq_dtype]
        out = np.empty(n_samples, np.complex64) if self.iq_dtype == IQ_CF32 else np.empty((n_samples, 2), np_t)
        check(self._lib.gb200_trk_read_iq(self._h, abs_index, out.ctypes.data, n_samples))
        return out

    def push_iq_device(self, dev_ptr, n_samples):
        first = C.c_uint64()
        check(self._lib.gb200_trk_push_iq_device(self._h, C.c_void_p(dev_ptr), n_samples, C.byref(first)))
        return first.value

    @property
    def head(self):
        return self._lib.gb200_trk_head(self._h)

    @staticmethod
    def make_cmds(ch, sample_offset, n_samples, shifts, rem_carr, phase_step, rem_code, code_step, phase_rate=None, code_rate=None, flags=0):
        """Arrays (one entry per command) -> ctypes array of gb200_trk_cmd.  shifts: (n_cmds, n_taps) or (1, n_taps)."""
        shifts = np.atleast_2d(np.asarray(shifts, dtype=np.float32))
        n = len(ch)
        arr = (TrkCmd * n)()
        v = np.frombuffer(arr, dtype=_CMD_DTYPE)  # field-by-field view of the same memory
        v["ch"], v["n_samples"], v["n_taps"], v["flags"] = np.asarray(ch), np.asarray(n_samples), shifts.shape[1], flags
        v["sample_offset"] = np.asarray(sample_offset, dtype=np.uint64)
        v["rem_carrier_phase_rad"], v["phase_step_rad"] = np.asarray(rem_carr, dtype=np.float32), np.asarray(phase_step, dtype=np.float32)
        v["rem_code_phase_chips"], v["code_phase_step_chips"] = np.asarray(rem_code, dtype=np.float32), np.asarray(code_step, dtype=np.float32)
        if phase_rate is not None:
            v["phase_rate_step_rad"] = np.asarray(phase_rate, dtype=np.float32)
        if code_rate is not None:
            v["code_phase_rate_step_chips"] = np.asarray(code_rate, dtype=np.float32)
        v["shifts_chips"][:, : shifts.shape[1]] = shifts
        arr._n_taps = shifts.shape[1]
        return arr

    # Carrier_wipeoff_multicorrelator_resampler for a batch of commands: (n_cmds, n_taps) complex64
    def host_outputs(self, n_cmds):
        """(n_cmds, GB200_MAX_TAPS) complex64 array in page-locked memory (gb200_host_alloc): the caller-owned `corr_out` of
        Cpu_Multicorrelator_Real_Codes::set_input_output_vectors, kept across calls; results are DMA'd straight into it."""
        p = C.c_void_p()
        nbytes = n_cmds * GB200_MAX_TAPS * 8
        check(self._lib.gb200_host_alloc(C.byref(p), nbytes))
        buf = (C.c_char * nbytes).from_address(p.value)
        arr = np.frombuffer(buf, dtype=np.complex64).reshape(n_cmds, GB200_MAX_TAPS)
        self._host_bufs.append(p)
        return arr

    def correlate(self, cmds, out=None):
        n = len(cmds)
        if out is None:
            out = np.empty((n, GB200_MAX_TAPS), dtype=np.complex64)
        else:
            assert out.dtype == np.complex64 and out.shape == (n, GB200_MAX_TAPS) and out.flags.c_contiguous
        check(self._lib.gb200_trk_correlate(self._h, cmds, n, out.ctypes.data))
        nt = getattr(cmds, "_n_taps", None) or max(c.n_taps for c in cmds)
        return out[:, :nt]

    def enqueue(self, cmds):
        self._n = len(cmds)
        check(self._lib.gb200_trk_enqueue(self._h, cmds, self._n))

    def collect(self):
        out = np.zeros((self._n, GB200_MAX_TAPS), dtype=np.complex64)
        check(self._lib.gb200_trk_collect(self._h, out.ctypes.data))
        return out

    @property
    def last_kernel(self):
        return self._lib.gb200_trk_last_kernel(self._h).decode()

    def replay(self, iters):
        ms = C.c_float()
        check(self._lib.gb200_trk_replay(self._h, iters, C.byref(ms)))
        return ms.value
