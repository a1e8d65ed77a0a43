"""ctypes binding of the host C++ layer's C door (gnss_sim_receiver_b200/host/b200_host_capi.h, libgnss_b200_host.so).

Plumbing for tests/ and bench.py.  The library is built in-tree by gnss_sim_receiver_b200/host/Makefile; a missing
library raises ImportError with the build command.
"""
import ctypes as C
import os

import numpy as np

from . import _capi

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libgnss_b200_host.so")


class TrkRecord(C.Structure):
    _fields_ = [("channel", C.c_int32), ("prn", C.c_uint32), ("valid", C.c_int32), ("correlation_length_ms", C.c_int32),
                ("prompt_i", C.c_double), ("prompt_q", C.c_double), ("cn0_db_hz", C.c_double), ("carrier_doppler_hz", C.c_double),
                ("carrier_phase_rads", C.c_double), ("code_phase_samples", C.c_double), ("tracking_sample_counter", C.c_uint64)]


class AcqStatus(C.Structure):
    _fields_ = [("acq_delay_samples", C.c_double), ("acq_doppler_hz", C.c_double), ("acq_samplestamp_samples", C.c_uint64),
                ("test_statistics", C.c_float), ("threshold", C.c_float), ("mag", C.c_float), ("input_power", C.c_float),
                ("state", C.c_int32), ("active", C.c_int32), ("last_event", C.c_int32), ("indext", C.c_uint32), ("doppler_index", C.c_int32)]


class FineDopplerStatus(C.Structure):
    """gb200h_acqfd_status"""
    _fields_ = [("acq_delay_samples", C.c_double), ("acq_doppler_hz", C.c_double), ("acq_samplestamp_samples", C.c_uint64), ("sample_counter", C.c_uint64),
                ("fine_index", C.c_int64), ("test_statistics", C.c_float), ("threshold", C.c_float), ("state", C.c_int32), ("active", C.c_int32),
                ("last_event", C.c_int32), ("positive_acq", C.c_int32), ("num_doppler_points", C.c_int32), ("fft_size", C.c_int32),
                ("acq_doppler_step", C.c_uint32), ("pad_", C.c_int32)]


class CafStatus(C.Structure):
    """gb200h_acqcaf_status"""
    _fields_ = [("acq_delay_samples", C.c_double), ("acq_doppler_hz", C.c_double), ("acq_samplestamp_samples", C.c_uint64), ("sample_counter", C.c_uint64),
                ("test_statistics", C.c_float), ("threshold", C.c_float), ("mag", C.c_float), ("input_power", C.c_float), ("state", C.c_int32), ("active", C.c_int32),
                ("last_event", C.c_int32), ("num_doppler_bins", C.c_int32), ("fft_size", C.c_int32), ("well_count", C.c_int32), ("acq_doppler_step", C.c_uint32), ("pad_", C.c_int32)]


class QuickSyncStatus(C.Structure):
    """gb200h_acqqs_status"""
    _fields_ = [("acq_delay_samples", C.c_double), ("acq_doppler_hz", C.c_double), ("acq_samplestamp_samples", C.c_uint64), ("sample_counter", C.c_uint64),
                ("test_statistics", C.c_float), ("threshold", C.c_float), ("mag", C.c_float), ("input_power", C.c_float), ("state", C.c_int32), ("active", C.c_int32),
                ("last_event", C.c_int32), ("num_doppler_bins", C.c_int32), ("fft_size", C.c_int32), ("well_count", C.c_int32), ("item_length", C.c_int32),
                ("folding_factor", C.c_int32), ("sampled_ms", C.c_int32), ("acq_doppler_step", C.c_uint32)]


class Synchro(C.Structure):
    """gb200h_synchro: every field of the GnssSynchro message (docs/protobuf/gnss_synchro.proto of the reference)"""
    _fields_ = [("system", C.c_char), ("signal", C.c_char * 3), ("prn", C.c_uint32), ("channel_id", C.c_int32),
                ("acq_delay_samples", C.c_double), ("acq_doppler_hz", C.c_double), ("acq_samplestamp_samples", C.c_uint64),
                ("acq_doppler_step", C.c_uint32), ("flag_valid_acquisition", C.c_int32), ("fs", C.c_int64),
                ("prompt_i", C.c_double), ("prompt_q", C.c_double), ("cn0_db_hz", C.c_double), ("carrier_doppler_hz", C.c_double),
                ("carrier_phase_rads", C.c_double), ("code_phase_samples", C.c_double), ("tracking_sample_counter", C.c_uint64),
                ("flag_valid_symbol_output", C.c_int32), ("correlation_length_ms", C.c_int32), ("flag_valid_word", C.c_int32),
                ("tow_at_current_symbol_ms", C.c_uint32), ("pseudorange_m", C.c_double), ("rx_time", C.c_double),
                ("flag_valid_pseudorange", C.c_int32), ("flag_pll_180_deg_phase_locked", C.c_int32), ("interp_tow_ms", C.c_double)]


REC_DTYPE = np.dtype([("channel", "<i4"), ("prn", "<u4"), ("valid", "<i4"), ("correlation_length_ms", "<i4"), ("prompt_i", "<f8"), ("prompt_q", "<f8"),
                      ("cn0_db_hz", "<f8"), ("carrier_doppler_hz", "<f8"), ("carrier_phase_rads", "<f8"), ("code_phase_samples", "<f8"),
                      ("tracking_sample_counter", "<u8")])
assert REC_DTYPE.itemsize == C.sizeof(TrkRecord)

_vp, _f, _d, _i, _s = C.c_void_p, C.c_float, C.c_double, C.c_int, C.c_char_p
SYMBOLS = {
    "gb200h_last_error": (_s, []),
    "gb200h_conf_get": (_s, [_s, _s, _s]),
    "gb200h_dll_pll_conf": (_s, [_s, _s]),
    "gb200h_acq_conf": (_s, [_s, _s, _d, _d, _i]),
    "gb200h_factory": (_s, [_s, _s]),
    "gb200h_pll_cloop_two_quadrant_atan": (_d, [_f, _f]),
    "gb200h_pll_four_quadrant_atan": (_d, [_f, _f]),
    "gb200h_fll_four_quadrant_atan": (_d, [_f, _f, _f, _f, _d, _d]),
    "gb200h_fll_diff_atan": (_d, [_f, _f, _f, _f, _d, _d]),
    "gb200h_dll_nc_e_minus_l_normalized": (_d, [_f] * 7),
    "gb200h_dll_nc_vemlp_normalized": (_d, [_f] * 8),
    "gb200h_e5a_q_secondary_code": (C.c_int, [C.c_uint, C.c_char_p]),
    "gb200h_sinboc_slope_abs": (_d, [_d]),
    "gb200h_sinboc_yintercept_abs": (_d, [_d]),
    "gb200h_cn0_m2m4_estimator": (_f, [_vp, _i, _f]),
    "gb200h_cn0_svn_estimator": (_f, [_vp, _i, _f]),
    "gb200h_carrier_lock_detector": (_f, [_vp, _i]),
    "gb200h_tracking_loop_filter_run": (None, [_f, _f, _i, _i, _f, _vp, _vp, _i]),
    "gb200h_fll_pll_filter_run": (None, [_f, _f, _i, _f, _vp, _vp, _f, _vp, _i]),
    "gb200h_exponential_smoother_run": (None, [_f, _i, _f, _f, _vp, _vp, _i]),
    "gb200h_gps_l1_ca_code_gen_float": (_i, [_vp, C.c_uint, C.c_uint]),
    "gb200h_gps_l1_ca_code_gen_complex_sampled": (_i, [_vp, _i, C.c_uint, _i, C.c_uint]),
    "gb200h_code_gen": (_i, [_i, _vp, _s, C.c_uint, _i, C.c_uint, _i, _i]),
    "gb200h_bank_create": (_i, [_s, _s, _i, _i, C.POINTER(_vp)]),
    "gb200h_bank_destroy": (None, [_vp]),
    "gb200h_bank_set_codes": (_i, [_vp, _i, _vp, _vp, _i]),
    "gb200h_bank_start": (_i, [_vp, _i, C.c_uint, _d, _d, C.c_uint64, C.c_uint64]),
    "gb200h_bank_push": (_i, [_vp, _vp, C.c_int64]),
    "gb200h_bank_push_raw": (C.c_int64, [_vp, _vp, C.c_int64, _i, _i, _i, _i, _i]),
    "gb200h_bank_step": (_i, [_vp, _vp, _i]),
    "gb200h_bank_channel_state": (_i, [_vp, _i, _vp]),
    "gb200h_bank_launches": (C.c_uint64, [_vp]),
    "gb200h_bank_run_threaded": (_i, [_vp, _vp, C.c_int64, C.c_uint64, _i, _vp, _i]),
    "gb200h_engine_probe": (_s, [_s, _s, C.c_uint, _s, C.c_uint]),
    "gb200h_acq_create": (_i, [_s, _s, C.c_uint, C.POINTER(_vp)]),
    "gb200h_acq_work_gr": (C.c_int64, [_vp, _vp, C.c_int64, C.POINTER(C.c_int)]),
    "gb200h_gr_tracking_run": (_i, [_s, _s, C.c_uint, _d, _d, _vp, C.c_int64, C.c_int64, _vp, _i, _vp]),
    "gb200h_acq_create2": (_i, [_s, _s, C.c_uint, C.c_char, _s, C.POINTER(_vp)]),
    "gb200h_acq_local_code": (C.c_int64, [_vp, C.c_void_p, C.c_int64]),
    "gb200h_acq_destroy": (None, [_vp]),
    "gb200h_acq_arm": (_i, [_vp, _f]),
    "gb200h_acq_work": (C.c_int64, [_vp, _vp, C.c_int64, _i]),
    "gb200h_acq_get_status": (_i, [_vp, C.POINTER(AcqStatus)]),
    "gb200h_acqcaf_create": (_i, [C.c_char_p, C.c_char_p, C.c_uint, C.c_char_p, C.POINTER(_vp)]),
    "gb200h_acqcaf_destroy": (None, [_vp]),
    "gb200h_acqcaf_arm": (_i, [_vp, C.c_uint, C.c_uint, C.c_float]),
    "gb200h_acqcaf_work": (C.c_int64, [_vp, _vp, C.c_int64, _i]),
    "gb200h_acqcaf_get_status": (_i, [_vp, C.POINTER(CafStatus)]),
    "gb200h_acqcaf_caf_vector": (_i, [_vp, _vp, _i]),
    "gb200h_acqfd_work_gr": (C.c_int64, [_vp, _vp, C.c_int64, C.POINTER(C.c_int)]),
    "gb200h_acqcaf_work_gr": (C.c_int64, [_vp, _vp, C.c_int64, C.POINTER(C.c_int)]),
    "gb200h_acqqs_work_gr": (C.c_int64, [_vp, _vp, C.c_int64, C.POINTER(C.c_int)]),
    "gb200h_acqqs_create": (_i, [C.c_char_p, C.c_char_p, C.c_uint, C.POINTER(_vp)]),
    "gb200h_acqqs_create2": (_i, [C.c_char_p, C.c_char_p, C.c_uint, C.c_char, C.c_char_p, C.POINTER(_vp)]),
    "gb200h_acqqs_destroy": (None, [_vp]),
    "gb200h_acqqs_arm": (_i, [_vp, C.c_uint, C.c_uint, C.c_float]),
    "gb200h_acqqs_work": (C.c_int64, [_vp, _vp, C.c_int64, _i]),
    "gb200h_acqqs_get_status": (_i, [_vp, C.POINTER(QuickSyncStatus)]),
    "gb200h_acqfd_create": (_i, [C.c_char_p, C.c_char_p, C.c_uint, C.POINTER(_vp)]),
    "gb200h_acqfd_destroy": (None, [_vp]),
    "gb200h_acqfd_arm": (_i, [_vp, C.c_uint, C.c_uint, C.c_float]),
    "gb200h_acqfd_work": (C.c_int64, [_vp, _vp, C.c_int64, _i]),
    "gb200h_acqfd_get_status": (_i, [_vp, C.POINTER(FineDopplerStatus)]),
    "gb200h_mat_write_sample": (_i, [_s, _vp, _i, _i, C.c_int32, C.c_uint32, C.c_uint64, _f]),
    "gb200h_rx_create": (_i, [_s, _s, _s, _i, C.POINTER(C.c_uint), _i, C.POINTER(_vp)]),
    "gb200h_rx_destroy": (None, [_vp]),
    "gb200h_rx_push": (_i, [_vp, _vp, C.c_int64, _vp, _i]),
    "gb200h_rx_channel_info": (_i, [_vp, _i, _vp]),
    "gb200h_rx_sweeps": (C.c_uint64, [_vp]),
    "gb200h_rx_threshold": (_f, [_vp]),
    "gb200h_synchro_serialize": (_i, [_vp, _i, _vp, _i]),
    "gb200h_monitor_create": (_i, [_s, _s, _i, C.POINTER(_vp)]),
    "gb200h_monitor_destroy": (None, [_vp]),
    "gb200h_monitor_enabled": (_i, [_vp]),
    "gb200h_monitor_write": (C.c_int64, [_vp, _i, _vp, _i]),
    "gb200h_bank_monitor_datagrams": (C.c_int64, [_vp]),
}

_LIB = None


def lib():
    global _LIB
    if _LIB is None:
        _capi.lib()  # the CUDA C-ABI library first (the host library links against it)
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} is missing: build it with `make -C gnss_sim_receiver_b200/host`")
        _LIB = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            f = getattr(_LIB, name)
            f.restype, f.argtypes = res, args
    return _LIB


class HostError(RuntimeError):
    pass


def check(rc):
    if rc < 0:
        raise HostError(lib().gb200h_last_error().decode(errors="replace"))
    return rc


def kv(text):
    """'a=1;b=x;' -> dict"""
    if isinstance(text, bytes):
        text = text.decode()
    return dict(item.split("=", 1) for item in text.strip(";").split(";") if "=" in item)


class TrackingBank:
    """N tracking blocks (adapter selected by `<role>.implementation` in the .conf text) over one GPU engine."""

    def __init__(self, conf_text, role, n_channels, max_engine_channels=0):
        self._lib = lib()
        self._h = _vp()
        check(self._lib.gb200h_bank_create(conf_text.encode(), role.encode(), n_channels, max_engine_channels, C.byref(self._h)))
        self.n = n_channels
        self._rec = np.zeros(1 << 15, dtype=REC_DTYPE)

    def close(self):
        if self._h:
            self._lib.gb200h_bank_destroy(self._h)
            self._h = _vp()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_codes(self, ch, tracking_code, data_code=None):
        t = np.ascontiguousarray(tracking_code, dtype=np.float32)
        d = None if data_code is None else np.ascontiguousarray(data_code, dtype=np.float32)
        check(self._lib.gb200h_bank_set_codes(self._h, ch, t.ctypes.data, None if d is None else d.ctypes.data, len(t)))

    def start(self, ch, prn, acq_delay_samples, acq_doppler_hz, acq_samplestamp, nitems_read):
        check(self._lib.gb200h_bank_start(self._h, ch, prn, acq_delay_samples, acq_doppler_hz, acq_samplestamp, nitems_read))

    def push(self, iq):
        iq = np.ascontiguousarray(iq)
        n = iq.size if np.iscomplexobj(iq) else iq.size // 2
        check(self._lib.gb200h_bank_push(self._h, iq.ctypes.data, n))

    def push_raw(self, raw, kind, sample_type=0, item_size=1, big_endian_items=False, big_endian_bytes=False):
        raw = np.ascontiguousarray(raw)
        return check(self._lib.gb200h_bank_push_raw(self._h, raw.ctypes.data, raw.nbytes, kind, sample_type, item_size, int(big_endian_items), int(big_endian_bytes)))

    @property
    def monitor_datagrams(self):
        return self._lib.gb200h_bank_monitor_datagrams(self._h)

    def step(self):
        n = check(self._lib.gb200h_bank_step(self._h, self._rec.ctypes.data, len(self._rec)))
        return self._rec[:n].copy()

    def run_threaded(self, iq, first_index, n_threads):
        iq = np.ascontiguousarray(iq)
        n = iq.size if np.iscomplexobj(iq) else iq.size // 2
        rec = np.zeros(self.n * (n // 1000 + 8), dtype=REC_DTYPE)
        k = check(self._lib.gb200h_bank_run_threaded(self._h, iq.ctypes.data, n, first_index, n_threads, rec.ctypes.data, len(rec)))
        return rec[:k]

    def channel_state(self, ch):
        v = np.zeros(8)
        check(self._lib.gb200h_bank_channel_state(self._h, ch, v.ctypes.data))
        return dict(zip(["state", "carrier_doppler_hz", "cn0_db_hz", "carrier_lock_test", "code_freq_chips", "rem_code_phase_samples", "nitems_read", "prn_length"], v))

    @property
    def launches(self):
        return self._lib.gb200h_bank_launches(self._h)


class AcquisitionBlock:
    """One B200 acquisition adapter (GPS L1 C/A, Galileo E1, GPS L5I or Galileo E5a, chosen by `<role>.implementation`) + its
    block, driven like Channel/ChannelFsm drive it.  `system` / `signal` fill the channel's Gnss_Synchro."""

    def __init__(self, conf_text, role, prn, system="G", signal="1C"):
        self._lib = lib()
        self._h = _vp()
        check(self._lib.gb200h_acq_create2(conf_text.encode(), role.encode(), prn, system.encode(), signal.encode(), C.byref(self._h)))

    def local_code(self):
        """the replica the adapter handed to pcps_acquisition::set_local_code (after arm())"""
        n = self._lib.gb200h_acq_local_code(self._h, None, 0)
        out = np.zeros(n, np.complex64)
        self._lib.gb200h_acq_local_code(self._h, out.ctypes.data, n)
        return out

    def close(self):
        if self._h:
            self._lib.gb200h_acq_destroy(self._h)
            self._h = _vp()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def arm(self, threshold_if_no_pfa=0.0):
        check(self._lib.gb200h_acq_arm(self._h, threshold_if_no_pfa))

    def work(self, items, chunk=2048):
        items = np.ascontiguousarray(items)
        n = items.size if np.iscomplexobj(items) else items.size // 2
        return check(self._lib.gb200h_acq_work(self._h, items.ctypes.data, n, chunk))

    def work_gr(self, items):
        """the same items through the block's gr::block face (b200_acquisition_gr); returns (consumed, last event on the "events" port)"""
        items = np.ascontiguousarray(items)
        n = items.size if np.iscomplexobj(items) else items.size // 2
        ev = C.c_int()
        done = check(self._lib.gb200h_acq_work_gr(self._h, items.ctypes.data, n, C.byref(ev)))
        return done, ev.value

    def status(self):
        s = AcqStatus()
        check(self._lib.gb200h_acq_get_status(self._h, C.byref(s)))
        return s


class FineDopplerAcquisitionBlock:
    """GPS_L1_CA_PCPS_Acquisition_Fine_Doppler_B200 (host/b200_pcps_acquisition_fine_doppler.h) driven like a channel drives it"""

    def __init__(self, conf_text, role, prn):
        self._lib = lib()
        self._h = _vp()
        check(self._lib.gb200h_acqfd_create(conf_text.encode(), role.encode(), prn, C.byref(self._h)))

    def close(self):
        if self._h:
            self._lib.gb200h_acqfd_destroy(self._h)
            self._h = _vp()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def arm(self, doppler_max, doppler_step, threshold):
        check(self._lib.gb200h_acqfd_arm(self._h, int(doppler_max), int(doppler_step), float(threshold)))

    def work(self, items, noutput_items):
        items = np.ascontiguousarray(items, dtype=np.complex64)
        return check(self._lib.gb200h_acqfd_work(self._h, items.ctypes.data, items.size, int(noutput_items)))

    def work_gr(self, items):
        """the same stream through the block's gr::block face (host/b200_gr_blocks.h); returns (consumed, last event on the "events" port)"""
        items = np.ascontiguousarray(items, dtype=np.complex64)
        ev = C.c_int()
        done = check(self._lib.gb200h_acqfd_work_gr(self._h, items.ctypes.data, items.size, C.byref(ev)))
        return done, ev.value

    def status(self):
        s = FineDopplerStatus()
        check(self._lib.gb200h_acqfd_get_status(self._h, C.byref(s)))
        return s


class E5aCafAcquisitionBlock:
    """Galileo_E5a_Noncoherent_IQ_Acquisition_CAF_B200 (host/b200_e5a_noncoherent_iq_acquisition_caf.h) driven like a channel drives it"""

    def __init__(self, conf_text, role, prn, signal="5X"):
        self._lib = lib()
        self._h = _vp()
        check(self._lib.gb200h_acqcaf_create(conf_text.encode(), role.encode(), prn, signal.encode(), C.byref(self._h)))

    def close(self):
        if self._h:
            self._lib.gb200h_acqcaf_destroy(self._h)
            self._h = _vp()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def arm(self, doppler_max, doppler_step, threshold):
        check(self._lib.gb200h_acqcaf_arm(self._h, int(doppler_max), int(doppler_step), float(threshold)))

    def work(self, items, chunk):
        items = np.ascontiguousarray(items, dtype=np.complex64)
        return check(self._lib.gb200h_acqcaf_work(self._h, items.ctypes.data, items.size, int(chunk)))

    def work_gr(self, items):
        """the same stream through the block's gr::block face (host/b200_gr_blocks.h); returns (consumed, last event on the "events" port)"""
        items = np.ascontiguousarray(items, dtype=np.complex64)
        ev = C.c_int()
        done = check(self._lib.gb200h_acqcaf_work_gr(self._h, items.ctypes.data, items.size, C.byref(ev)))
        return done, ev.value

    def status(self):
        s = CafStatus()
        check(self._lib.gb200h_acqcaf_get_status(self._h, C.byref(s)))
        return s

    def caf_vector(self):
        n = self._lib.gb200h_acqcaf_caf_vector(self._h, None, 0)
        out = np.zeros(n, np.float32)
        if n:
            self._lib.gb200h_acqcaf_caf_vector(self._h, out.ctypes.data, n)
        return out


class QuickSyncAcquisitionBlock:
    """GPS_L1_CA_PCPS_QuickSync_Acquisition_B200 / Galileo_E1_PCPS_QuickSync_Ambiguous_Acquisition_B200 (host/b200_pcps_quicksync_acquisition.h) driven like a channel drives it"""

    def __init__(self, conf_text, role, prn, system="G", signal="1C"):
        self._lib = lib()
        self._h = _vp()
        check(self._lib.gb200h_acqqs_create2(conf_text.encode(), role.encode(), prn, system.encode(), signal.encode(), C.byref(self._h)))

    def close(self):
        if self._h:
            self._lib.gb200h_acqqs_destroy(self._h)
            self._h = _vp()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def arm(self, doppler_max, doppler_step, threshold):
        check(self._lib.gb200h_acqqs_arm(self._h, int(doppler_max), int(doppler_step), float(threshold)))

    def work(self, samples, max_items):
        """returns the vector ITEMS consumed"""
        samples = np.ascontiguousarray(samples, dtype=np.complex64)
        return check(self._lib.gb200h_acqqs_work(self._h, samples.ctypes.data, samples.size, int(max_items)))

    def work_gr(self, items):
        """the same stream through the block's gr::block face (host/b200_gr_blocks.h); returns (consumed, last event on the "events" port)"""
        items = np.ascontiguousarray(items, dtype=np.complex64)
        ev = C.c_int()
        done = check(self._lib.gb200h_acqqs_work_gr(self._h, items.ctypes.data, items.size, C.byref(ev)))
        return done, ev.value

    def status(self):
        s = QuickSyncStatus()
        check(self._lib.gb200h_acqqs_get_status(self._h, C.byref(s)))
        return s


class ChannelManager:
    """B200ChannelManager (host/b200_channel_manager.h): batched all-PRN acquisition sweep feeding a lock-step tracking bank."""

    def __init__(self, conf_text, acq_role, trk_role, n_channels, prns):
        self._lib = lib()
        self._h = _vp()
        self.n = n_channels
        p = (C.c_uint * len(prns))(*[int(x) for x in prns])
        check(self._lib.gb200h_rx_create(conf_text.encode(), acq_role.encode(), trk_role.encode(), n_channels, p, len(prns), C.byref(self._h)))
        self._rec = np.zeros(1 << 16, dtype=REC_DTYPE)

    def close(self):
        if self._h:
            self._lib.gb200h_rx_destroy(self._h)
            self._h = _vp()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def push(self, iq):
        iq = np.ascontiguousarray(iq)
        n = iq.size if np.iscomplexobj(iq) else iq.size // 2
        k = check(self._lib.gb200h_rx_push(self._h, iq.ctypes.data, n, self._rec.ctypes.data, len(self._rec)))
        return self._rec[:k].copy()

    def channel_info(self, ch):
        v = np.zeros(10)
        check(self._lib.gb200h_rx_channel_info(self._h, ch, v.ctypes.data))
        keys = ["state", "prn", "acquisitions", "losses", "acq_doppler_hz", "acq_delay_samples", "acq_samplestamp_samples", "test_statistics", "trk_state", "cn0_db_hz"]
        return dict(zip(keys, v))

    @property
    def sweeps(self):
        return self._lib.gb200h_rx_sweeps(self._h)

    @property
    def threshold(self):
        return self._lib.gb200h_rx_threshold(self._h)


def _synchro_array(items):
    arr = (Synchro * len(items))()
    for a, d in zip(arr, items):
        for k, v in d.items():
            setattr(a, k, v.encode() if isinstance(v, str) else v)
    return arr


def serialize_synchros(items):
    """Serdes_Gnss_Synchro::createProtobuffer on a list of dicts keyed by the proto field names -> bytes."""
    arr = _synchro_array(items)
    buf = C.create_string_buffer(512 * max(1, len(items)))
    n = check(lib().gb200h_synchro_serialize(arr, len(items), buf, len(buf)))
    return buf.raw[:n]


class SynchroMonitor:
    """gnss_synchro_monitor + Gnss_Synchro_Udp_Sink: decimated protobuf datagrams to the configured clients."""

    def __init__(self, conf_text, role, n_channels):
        self._lib = lib()
        self._h = _vp()
        check(self._lib.gb200h_monitor_create(conf_text.encode(), role.encode(), n_channels, C.byref(self._h)))

    @property
    def enabled(self):
        return bool(self._lib.gb200h_monitor_enabled(self._h))

    def write(self, channel, items):
        arr = _synchro_array(items)
        return check(self._lib.gb200h_monitor_write(self._h, channel, arr, len(items)))

    def close(self):
        if self._h:
            self._lib.gb200h_monitor_destroy(self._h)
            self._h = _vp()

    def __del__(self):
        self.close()


def gr_tracking_run(conf_text, role, prn, acq_delay_samples, acq_doppler_hz, iq, telemetry_fault_at=-1):
    """One tracking adapter behind its gr::block face (b200_tracking_gr over the GNU Radio stub): returns (records, events dict)."""
    iq = np.ascontiguousarray(iq)
    n = iq.size if np.iscomplexobj(iq) else iq.size // 2
    rec = np.zeros(n // 1000 + 64, dtype=REC_DTYPE)
    ev = (C.c_int64 * 4)()
    k = check(lib().gb200h_gr_tracking_run(conf_text.encode(), role.encode(), prn, acq_delay_samples, acq_doppler_hz, iq.ctypes.data, n, telemetry_fault_at, rec.ctypes.data, len(rec), ev))
    return rec[:k].copy(), dict(n_events=ev[0], last_event=ev[1], consumed=ev[2], state=ev[3])
