"""ctypes binding of the C ABI in include/gnss_b200.h (libgnss_b200.so, built in-tree by csrc/Makefile).

Plumbing only: the structures mirror the header field by field.  There is no CPU fallback; a missing library
raises ImportError with the build command, a missing GPU makes every *_create call fail with GB200_ENODEVICE.
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libgnss_b200.so")

GB200_MAX_TAPS = 16
IQ_CF32, IQ_CS16, IQ_CS8 = 0, 1, 2
TRK_HIGH_DYN = 1


class TrkConfig(C.Structure):
    _fields_ = [("device", C.c_int32), ("max_channels", C.c_int32), ("max_code_len", C.c_int32), ("iq_dtype", C.c_int32),
                ("ring_samples", C.c_int64), ("max_block", C.c_int32), ("max_cmds", C.c_int32)]


class IngestFormat(C.Structure):
    """gb200_ingest_format (include/gnss_b200.h)"""
    _fields_ = [("kind", C.c_int32), ("sample_type", C.c_int32), ("item_size", C.c_int32), ("big_endian_items", C.c_int32),
                ("big_endian_bytes", C.c_int32)]


class TrkCmd(C.Structure):
    _fields_ = [("ch", C.c_int32), ("n_samples", C.c_int32), ("n_taps", C.c_int32), ("flags", C.c_int32),
                ("sample_offset", C.c_uint64),
                ("rem_carrier_phase_rad", C.c_float), ("phase_step_rad", C.c_float), ("phase_rate_step_rad", C.c_float),
                ("rem_code_phase_chips", C.c_float), ("code_phase_step_chips", C.c_float), ("code_phase_rate_step_chips", C.c_float),
                ("shifts_chips", C.c_float * GB200_MAX_TAPS)]


class LoopConfig(C.Structure):
    """gb200_loop_config (include/gnss_b200.h): the signal table + Dll_Pll_Conf values the device-resident loop closure needs"""
    _fields_ = [("fs_in", C.c_double), ("code_period_s", C.c_double), ("code_chip_rate", C.c_double), ("signal_carrier_freq", C.c_double),
                ("code_length_chips", C.c_int32), ("code_samples_per_chip", C.c_int32), ("vector_length", C.c_int32), ("veml", C.c_int32),
                ("symbols_per_bit", C.c_int32), ("secondary", C.c_int32), ("secondary_code_length", C.c_int32), ("interchange_iq", C.c_int32),
                ("secondary_code", C.c_char * 192),
                ("pll_bw_hz", C.c_float), ("dll_bw_hz", C.c_float), ("fll_bw_hz", C.c_float), ("pll_bw_narrow_hz", C.c_float), ("dll_bw_narrow_hz", C.c_float),
                ("early_late_space_chips", C.c_float), ("very_early_late_space_chips", C.c_float), ("early_late_space_narrow_chips", C.c_float),
                ("very_early_late_space_narrow_chips", C.c_float), ("slope", C.c_float), ("y_intercept", C.c_float),
                ("carrier_lock_th", C.c_float), ("cn0_smoother_alpha", C.c_float), ("carrier_lock_test_smoother_alpha", C.c_float),
                ("pll_filter_order", C.c_int32), ("dll_filter_order", C.c_int32), ("extend_correlation_symbols", C.c_int32),
                ("enable_fll_pull_in", C.c_int32), ("enable_fll_steady_state", C.c_int32), ("carrier_aiding", C.c_int32),
                ("cn0_samples", C.c_int32), ("cn0_min", C.c_int32), ("max_code_lock_fail", C.c_int32), ("max_carrier_lock_fail", C.c_int32),
                ("cn0_smoother_samples", C.c_int32), ("carrier_lock_test_smoother_samples", C.c_int32),
                ("pull_in_time_s", C.c_uint32), ("bit_synchronization_time_limit_s", C.c_uint32), ("n_correlators", C.c_int32), ("track_pilot", C.c_int32),
                ("data_secondary_code_length", C.c_int32), ("data_secondary_code", C.c_char * 36)]


class LoopChannelState(C.Structure):
    _fields_ = [("state", C.c_int32), ("pull_in_transitory", C.c_int32), ("current_prn_length_samples", C.c_int32), ("narrow", C.c_int32),
                ("carrier_doppler_hz", C.c_double), ("cn0_db_hz", C.c_double), ("carrier_lock_test", C.c_double), ("code_freq_chips", C.c_double),
                ("rem_code_phase_samples", C.c_double), ("acc_carrier_phase_rad", C.c_double), ("nitems_read", C.c_uint64)]


class AcqConfig(C.Structure):
    _fields_ = [("device", C.c_int32), ("iq_dtype", C.c_int32), ("fft_size", C.c_int32), ("consumed_samples", C.c_int32),
                ("bit_transition", C.c_int32), ("use_cfar", C.c_int32), ("samples_per_chip", C.c_int32), ("max_dwells", C.c_int32),
                ("max_prns", C.c_int32), ("samples_per_code", C.c_float), ("fs_in", C.c_double)]


class AcqResult(C.Structure):
    _fields_ = [("indext", C.c_uint32), ("doppler_index", C.c_int32), ("doppler_hz", C.c_int32), ("peak", C.c_float),
                ("input_power", C.c_float), ("test_statistics", C.c_float), ("acq_delay_samples", C.c_float), ("acq_doppler_hz", C.c_float)]


# every symbol include/gnss_b200.h declares: name -> (restype, argtypes)
_vp = C.c_void_p
SYMBOLS = {
    "gb200_last_error": (C.c_char_p, []),
    "gb200_launch_count": (C.c_uint64, []),
    "gb200_device_count": (C.c_int, []),
    "gb200_version": (C.c_char_p, []),
    "gb200_host_alloc": (C.c_int, [C.POINTER(C.c_void_p), C.c_size_t]),
    "gb200_host_free": (C.c_int, [C.c_void_p]),
    "gb200_trk_create": (C.c_int, [C.POINTER(TrkConfig), C.POINTER(_vp)]),
    "gb200_trk_destroy": (C.c_int, [_vp]),
    "gb200_trk_set_code": (C.c_int, [_vp, C.c_int, _vp, C.c_int]),
    "gb200_trk_set_code_complex": (C.c_int, [_vp, C.c_int, _vp, C.c_int]),
    "gb200_trk_push_iq": (C.c_int, [_vp, _vp, C.c_int64, C.POINTER(C.c_uint64)]),
    "gb200_trk_push_iq_device": (C.c_int, [_vp, _vp, C.c_int64, C.POINTER(C.c_uint64)]),
    "gb200_trk_skip": (C.c_int, [_vp, C.c_int64]),
    "gb200_trk_write_iq": (C.c_int, [_vp, C.c_uint64, _vp, C.c_int64]),
    "gb200_trk_push_raw": (C.c_int, [_vp, _vp, C.c_int64, _vp, C.POINTER(C.c_uint64), C.POINTER(C.c_int64)]),
    "gb200_trk_read_iq": (C.c_int, [_vp, C.c_uint64, _vp, C.c_int64]),
    "gb200_trk_head": (C.c_uint64, [_vp]),
    "gb200_trk_correlate": (C.c_int, [_vp, _vp, C.c_int, _vp]),
    "gb200_trk_correlate_noalloc": (C.c_int, [C.c_int, _vp, C.c_int64, _vp, C.c_int, C.c_int, _vp, C.c_int, _vp, _vp]),
    "gb200_trk_enqueue": (C.c_int, [_vp, _vp, C.c_int]),
    "gb200_trk_collect": (C.c_int, [_vp, _vp]),
    "gb200_trk_replay": (C.c_int, [_vp, C.c_int, C.POINTER(C.c_float)]),
    "gb200_trk_last_kernel": (C.c_char_p, [_vp]),
    "gb200_loop_create": (C.c_int, [_vp, C.POINTER(LoopConfig), C.c_int, C.POINTER(_vp)]),
    "gb200_loop_destroy": (C.c_int, [_vp]),
    "gb200_loop_start": (C.c_int, [_vp, C.c_int, C.c_uint32, C.c_double, C.c_double, C.c_uint64, C.c_uint64]),
    "gb200_loop_stop": (C.c_int, [_vp, C.c_int]),
    "gb200_loop_set_secondary_code": (C.c_int, [_vp, C.c_int, C.c_char_p, C.c_int]),
    "gb200_loop_step": (C.c_int, [_vp, _vp, C.c_int64, _vp, C.c_int, C.POINTER(C.c_int)]),
    "gb200_loop_submit": (C.c_int, [_vp, _vp, C.c_int64]),
    "gb200_loop_collect": (C.c_int, [_vp, _vp, C.c_int, C.POINTER(C.c_int)]),
    "gb200_loop_channel": (C.c_int, [_vp, C.c_int, C.POINTER(LoopChannelState)]),
    "gb200_loop_read_taps": (C.c_int, [_vp, C.c_int, _vp]),
    "gb200_loop_last_step": (C.c_int, [_vp, C.POINTER(C.c_float), C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "gb200_loop_kernel": (C.c_char_p, [_vp]),
    "gb200_acq_create": (C.c_int, [C.POINTER(AcqConfig), C.POINTER(_vp)]),
    "gb200_acq_destroy": (C.c_int, [_vp]),
    "gb200_acq_set_local_code": (C.c_int, [_vp, C.c_int, _vp, C.c_int]),
    "gb200_acq_set_doppler_grid": (C.c_int, [_vp, C.c_int, C.c_int, C.c_int]),
    "gb200_acq_set_doppler_bins": (C.c_int, [_vp, C.POINTER(C.c_float), C.c_int]),
    "gb200_acq_set_wipeoff_mode": (C.c_int, [_vp, C.c_int]),
    "gb200_acq_num_doppler_bins": (C.c_int, [_vp]),
    "gb200_acq_neff": (C.c_int, [_vp]),
    "gb200_acq_run": (C.c_int, [_vp, _vp, C.c_int, _vp, C.c_int, _vp]),
    "gb200_acq_run_device": (C.c_int, [_vp, _vp, C.c_int, _vp, C.c_int, _vp]),
    "gb200_acq_replay": (C.c_int, [_vp, C.c_int, C.POINTER(C.c_float)]),
    "gb200_acq_threshold": (C.c_float, [_vp, C.c_float]),
    "gb200_acq_keep_grid": (C.c_int, [_vp, C.c_int]),
    "gb200_acq_read_grid": (C.c_int, [_vp, C.c_int, _vp]),
    "gb200_acq_read_cells": (C.c_int, [_vp, C.c_int, _vp, _vp, _vp]),
    "gb200_acq_fft": (C.c_int, [_vp, _vp, _vp, C.c_int]),
    "gb200_acq_set_input_folding": (C.c_int, [_vp, C.c_int, C.c_int]),
    "gb200_acq_folded_candidates": (C.c_int, [_vp, C.c_int, C.c_int, _vp, C.c_int, C.c_int, _vp]),
    "gb200_acq_grid_pair_max": (C.c_int, [_vp, _vp, _vp, _vp, _vp]),
    "gb200_acq_grid_probe": (C.c_int, [_vp, C.c_int, _vp, _vp, _vp, _vp]),
    "gb200_acq_spectrum_peak": (C.c_int, [_vp, _vp, C.c_int64, _vp, C.c_int, C.c_int, _vp, _vp]),
}

_LIB = None


def lib():
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} is missing: build it with `make -C gnss_sim_receiver_b200/csrc` (python -c 'import __graft_entry__ as g; g.build()')")
        _LIB = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            f = getattr(_LIB, name)  # AttributeError here = header/library mismatch, which must be loud
            f.restype, f.argtypes = res, args
    return _LIB


class GB200Error(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"gb200 error {code}: {msg}")
        self.code = code


def check(rc):
    if rc != 0:
        raise GB200Error(rc, lib().gb200_last_error().decode(errors="replace"))
