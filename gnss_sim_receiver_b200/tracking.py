"""Python door onto the tracking multicorrelator engine (C ABI gb200_trk_*), used by tests/ and bench.py.

The product host layer is the C++ under gnss_sim_receiver_b200/host/; this module only marshals numpy arrays
into the C structures.  Argument names follow Cpu_Multicorrelator_Real_Codes
(src/algorithms/tracking/libs/cpu_multicorrelator_real_codes.h:37-61 of the reference).
"""
import ctypes as C

import numpy as np

from . import _capi
from ._capi import GB200_MAX_TAPS, IQ_CF32, IQ_CS16, IQ_CS8, IngestFormat, LoopChannelState, LoopConfig, TrkCmd, TrkConfig, check

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


# one emitted Gnss_Synchro (gb200_loop_record); same layout as the host bank's record (gb200h_trk_record)
LOOP_REC = np.dtype([("channel", "<i4"), ("prn", "<u4"), ("valid", "<i4"), ("correlation_length_ms", "<i4"), ("prompt_i", "<f8"), ("prompt_q", "<f8"),
                     ("cn0_db_hz", "<f8"), ("carrier_doppler_hz", "<f8"), ("carrier_phase_rads", "<f8"), ("code_phase_samples", "<f8"),
                     ("tracking_sample_counter", "<u8")])

_GPS_CA_PREAMBLE_SYMBOLS = "".join(c * 20 for c in "10001011")  # GPS_CA_PREAMBLE_SYMBOLS_STR (GPS_L1_CA.h:73)
# the signal table of dll_pll_veml_tracking.cc:171-447 for the data-component (track_pilot = false) signals the device loop serves
_SIGNALS = {
    "1C": dict(code_period_s=0.001, code_chip_rate=1.023e6, signal_carrier_freq=1575.42e6, code_length_chips=1023, code_samples_per_chip=1, veml=0, symbols_per_bit=20,
               secondary=0, secondary_code=_GPS_CA_PREAMBLE_SYMBOLS, interchange_iq=0),
    "L5": dict(code_period_s=0.001, code_chip_rate=10.23e6, signal_carrier_freq=1176.45e6, code_length_chips=10230, code_samples_per_chip=1, veml=0, symbols_per_bit=10,
               secondary=1, secondary_code="0000110101", interchange_iq=1),
    "1B": dict(code_period_s=0.004, code_chip_rate=1.023e6, signal_carrier_freq=1575.42e6, code_length_chips=4092, code_samples_per_chip=2, veml=1, symbols_per_bit=1,
               secondary=0, secondary_code="", interchange_iq=0),
    "5X": dict(code_period_s=0.001, code_chip_rate=10.23e6, signal_carrier_freq=1176.45e6, code_length_chips=10230, code_samples_per_chip=1, veml=0, symbols_per_bit=20,
               secondary=1, secondary_code="10000100001011101001", interchange_iq=0),
}


# pilot tracking (track_pilot = true, dll_pll_veml_tracking.cc:227-237, :278-284): the tracked component's secondary code, and the data component's
_PILOT = {
    "1B": dict(secondary=1, secondary_code="0011100000001010110110010", data_secondary_code="", interchange_iq=0),       # GALILEO_E1_C_SECONDARY_CODE
    "L5": dict(secondary=1, secondary_code="00000100110101001110", data_secondary_code="0000110101", interchange_iq=0),  # GPS_L5Q_NH_CODE_STR / GPS_L5I_NH_CODE_STR
    # Galileo E5a-Q: 100-chip secondary code PER SATELLITE (TrackingLoop.set_secondary_code), the data component's CS20 removed, I / Q interchanged (:305-314)
    "5X": dict(secondary=1, secondary_code="0" * 100, data_secondary_code="10000100001011101001", interchange_iq=1),
}


def loop_config(signal, fs_in, **overrides):
    """gb200_loop_config of one signal with the reference's Dll_Pll_Conf defaults (src/algorithms/tracking/libs/dll_pll_conf.cc:36-125,
    gnss_sdr_flags.cc:48-57); keyword arguments override fields by name."""
    sig = dict(_SIGNALS[signal])
    data_code = ""
    if overrides.get("track_pilot"):
        if signal not in _PILOT:
            raise ValueError(f"track_pilot is served on the device for {sorted(_PILOT)}; {signal!r} goes through the host block")
        sig.update(_PILOT[signal])
        data_code = sig.pop("data_secondary_code")
    code = sig.pop("secondary_code")
    c = LoopConfig()
    c.fs_in = float(fs_in)
    for k, v in sig.items():
        setattr(c, k, v)
    c.secondary_code = code.encode()
    c.secondary_code_length = len(code)
    c.data_secondary_code = data_code.encode()
    c.data_secondary_code_length = len(data_code)
    c.vector_length = int(round(fs_in / (c.code_chip_rate / c.code_length_chips)))
    d = dict(pll_bw_hz=35.0, dll_bw_hz=2.0, fll_bw_hz=35.0, pll_bw_narrow_hz=5.0, dll_bw_narrow_hz=0.75, early_late_space_chips=0.25 if signal != "1B" else 0.25,
             very_early_late_space_chips=0.5, early_late_space_narrow_chips=0.15, very_early_late_space_narrow_chips=0.5, slope=1.0, y_intercept=1.0,
             carrier_lock_th=0.7, cn0_smoother_alpha=0.002, carrier_lock_test_smoother_alpha=0.002, pll_filter_order=3, dll_filter_order=2,
             extend_correlation_symbols=1, enable_fll_pull_in=0, enable_fll_steady_state=0, carrier_aiding=1, cn0_samples=20, cn0_min=25,
             max_code_lock_fail=50, max_carrier_lock_fail=5000, cn0_smoother_samples=200, carrier_lock_test_smoother_samples=25, pull_in_time_s=10,
             bit_synchronization_time_limit_s=20, n_correlators=0, track_pilot=0)
    d.update(overrides)
    for k, v in d.items():
        if not hasattr(c, k):
            raise AttributeError(f"gb200_loop_config has no field {k!r}")
        setattr(c, k, v)
    return c


class TrackingLoop:
    """Device-resident DLL/PLL loop closure over a TrackingEngine (C ABI gb200_loop_*): channel i uses the engine's code slot i."""

    def __init__(self, engine, config, n_channels):
        self._lib = _capi.lib()
        self.engine, self.cfg, self.n = engine, config, n_channels
        self._h = C.c_void_p()
        check(self._lib.gb200_loop_create(engine._h, C.byref(config), n_channels, C.byref(self._h)))
        self._rec = np.zeros(max(4096, 9 * n_channels), dtype=LOOP_REC)
        assert LOOP_REC.itemsize == 72

    def close(self):
        if self._h:
            self._lib.gb200_loop_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def start(self, ch, prn, acq_delay_samples, acq_doppler_hz, acq_samplestamp=0, nitems_read=0):
        check(self._lib.gb200_loop_start(self._h, ch, prn, acq_delay_samples, acq_doppler_hz, acq_samplestamp, nitems_read))

    def stop(self, ch):
        check(self._lib.gb200_loop_stop(self._h, ch))

    def set_secondary_code(self, ch, code):
        """per-satellite secondary code of the tracked component ('0' / '1' string; Galileo E5a-Q), before start()"""
        check(self._lib.gb200_loop_set_secondary_code(self._h, ch, code.encode(), len(code)))

    def step(self, iq=None, copy=True):
        """Append `iq` (host samples of the engine's item type) and run the rounds that are due; returns the records emitted."""
        n = C.c_int()
        if iq is None:
            ptr, ns = None, 0
        else:
            iq = np.ascontiguousarray(iq)
            ns = iq.size if self.engine.iq_dtype == IQ_CF32 else iq.size // 2
            ptr = iq.ctypes.data
        check(self._lib.gb200_loop_step(self._h, ptr, ns, self._rec.ctypes.data, len(self._rec), C.byref(n)))
        k = min(n.value, len(self._rec))
        return self._rec[:k].copy() if copy else self._rec[:k]

    def submit(self, iq=None):
        """pipelined form: enqueue a step and return at once (`iq` must stay alive, and should be page-locked, until collected)"""
        if iq is None:
            ptr, ns = None, 0
        else:
            iq = np.ascontiguousarray(iq)
            ns = iq.size if self.engine.iq_dtype == IQ_CF32 else iq.size // 2
            ptr = iq.ctypes.data
        check(self._lib.gb200_loop_submit(self._h, ptr, ns))

    def collect(self, copy=True):
        """records of the oldest submitted step"""
        n = C.c_int()
        check(self._lib.gb200_loop_collect(self._h, self._rec.ctypes.data, len(self._rec), C.byref(n)))
        k = min(n.value, len(self._rec))
        return self._rec[:k].copy() if copy else self._rec[:k]

    def channel(self, ch):
        s = LoopChannelState()
        check(self._lib.gb200_loop_channel(self._h, ch, C.byref(s)))
        return {k: getattr(s, k) for k, _ in LoopChannelState._fields_}

    def read_taps(self, ch):
        """correlator outputs of channel ch's last correlated round (n_correlators complex values)"""
        out = np.zeros(GB200_MAX_TAPS, np.complex64)
        check(self._lib.gb200_loop_read_taps(self._h, ch, out.ctypes.data))
        return out[:max(self.cfg.n_correlators, 5 if self.cfg.veml else 3)]

    @property
    def last_step(self):
        ms, nc, nr = C.c_float(), C.c_int(), C.c_int()
        check(self._lib.gb200_loop_last_step(self._h, C.byref(ms), C.byref(nc), C.byref(nr)))
        return dict(device_ms=ms.value, commands=nc.value, rounds=nr.value)

    @property
    def kernel(self):
        return self._lib.gb200_loop_kernel(self._h).decode()


class BankedTrackingLoop:
    """The channels of one GPU as `n_banks` independent banks (a TrackingEngine + TrackingLoop each, n_channels / n_banks channels per bank) behind
    TrackingLoop's interface: channel numbers are global (0 .. n_channels-1) on the way in and in the records.

    Why: a bank's period is correlate -> loop closure -> correlate ... on the engine's one stream, so the loop-closure kernel (a short dependent chain
    on two warps per SM) sits between two correlator launches.  With two banks — two streams — it runs under the other bank's correlator: 0.302 instead
    of 0.331 ms per period for 10 240 channels at 25 MS/s, identical records (bench_closed_loop.py::run_two_banks).  Every bank keeps its own copy of the
    IQ ring, so a block is uploaded n_banks times.  More than two banks add nothing; below about 2 000 channels a period is bound by the host calls and
    ONE bank (TrackingLoop) is the faster form (profiles/r02p_two_banks_channel_counts.jsonl)."""

    def __init__(self, config, n_channels, n_banks=2, device=0, max_code_len=10230, iq_dtype=IQ_CF32, ring_samples=1 << 20, max_block=1 << 17):
        if n_banks < 1 or n_channels < n_banks:
            raise ValueError(f"{n_channels} channels cannot be split into {n_banks} banks")
        per = -(-n_channels // n_banks)
        self.n, self.cfg, self.per = n_channels, config, per
        self.engines, self.loops = [], []
        self._last = dict(device_ms=0.0, commands=0, rounds=0)
        slots = 2 if config.track_pilot else 1  # code slots and commands per channel (tracking + data component)
        try:
            for b in range(n_banks):
                k = min(per, n_channels - b * per)
                if k <= 0:
                    break
                e = TrackingEngine(device=device, max_channels=slots * k, max_code_len=max_code_len, iq_dtype=iq_dtype, ring_samples=ring_samples, max_block=max_block,
                                   max_cmds=slots * k)
                self.engines.append(e)
                self.loops.append(TrackingLoop(e, config, k))
        except Exception:
            self.close()
            raise

    def _at(self, ch):
        if not 0 <= ch < self.n:
            raise IndexError(f"channel {ch} out of range 0..{self.n - 1}")
        return ch // self.per, ch % self.per

    def close(self):
        for lp in self.loops:
            lp.close()
        for e in self.engines:
            e.close()
        self.loops, self.engines = [], []

    def set_code(self, ch, code, data_component=False):
        """code table of channel ch (the tracked component; data_component=True: the data component of a track_pilot bank)"""
        b, c = self._at(ch)
        self.engines[b].set_code(c + (self.loops[b].n if data_component else 0), code)

    def start(self, ch, *a, **k):
        b, c = self._at(ch)
        self.loops[b].start(c, *a, **k)

    def stop(self, ch):
        b, c = self._at(ch)
        self.loops[b].stop(c)

    def set_secondary_code(self, ch, code):
        b, c = self._at(ch)
        self.loops[b].set_secondary_code(c, code)

    def channel(self, ch):
        b, c = self._at(ch)
        return self.loops[b].channel(c)

    def read_taps(self, ch):
        b, c = self._at(ch)
        return self.loops[b].read_taps(c)

    def submit(self, iq=None):
        """enqueue the step on every bank and return at once (`iq` must stay alive, and should be page-locked, until collected)"""
        for lp in self.loops:
            lp.submit(iq)

    def collect(self):
        """records of the oldest submitted step of every bank, channel numbers global, in bank order"""
        out = []
        self._last = dict(device_ms=0.0, commands=0, rounds=0)
        for b, lp in enumerate(self.loops):
            r = lp.collect(copy=True)
            r["channel"] += b * self.per
            out.append(r)
            st = lp.last_step
            self._last = dict(device_ms=max(self._last["device_ms"], st["device_ms"]), commands=self._last["commands"] + st["commands"], rounds=max(self._last["rounds"], st["rounds"]))
        return np.concatenate(out)

    def step(self, iq=None):
        self.submit(iq)
        return self.collect()

    @property
    def last_step(self):
        """device_ms: the slowest bank's (the banks overlap); commands: all banks'"""
        return dict(self._last)

    @property
    def kernel(self):
        return self.loops[0].kernel
