"""ctypes doors onto the TEST-ONLY oracles.

* ``oracle/liboracle.so``       — our plain-C restatement (always buildable; `make -C oracle liboracle.so`)
* ``oracle/_ref/libgnssref.so`` — the reference's own code compiled where it lies (built in the dev container only;
  the prebuilt .so travels to the GPU box, /root/reference does not)

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs import this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")

f32p = C.POINTER(C.c_float)
f64p = C.POINTER(C.c_double)
i32p = C.POINTER(C.c_int32)
i64p = C.POINTER(C.c_int64)


def _p(a, t=f32p):
    return a.ctypes.data_as(t)


def build_oracle():
    so = os.path.join(ORACLE_DIR, "liboracle.so")
    srcs = [os.path.join(ORACLE_DIR, f) for f in os.listdir(ORACLE_DIR) if f.endswith((".c", ".h"))]
    if not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
        subprocess.check_call(["make", "-C", ORACLE_DIR, "liboracle.so"], stdout=subprocess.DEVNULL)
    return so


class AcqConf(C.Structure):
    _fields_ = [
        ("fft_size", C.c_int32), ("consumed", C.c_int32), ("num_doppler_bins", C.c_int32), ("bit_transition", C.c_int32),
        ("use_cfar", C.c_int32), ("samples_per_chip", C.c_int32), ("doppler_max", C.c_int32), ("doppler_step", C.c_int32),
        ("doppler_center", C.c_int32), ("max_dwells", C.c_int32), ("samples_per_code", C.c_float), ("fs_in", C.c_double),
        ("fft_f64", C.c_int32), ("sincos_arch", C.c_int32),
    ]


class AcqResult(C.Structure):
    _fields_ = [
        ("indext", C.c_uint32), ("doppler_index", C.c_int32), ("doppler_hz", C.c_int32), ("peak", C.c_float),
        ("input_power", C.c_float), ("test_statistics", C.c_float), ("acq_delay_samples", C.c_float), ("acq_doppler_hz", C.c_float),
    ]


_ORC = None
_REF = None


def orc():
    global _ORC
    if _ORC is None:
        _ORC = C.CDLL(build_oracle())
        _ORC.orc_acq_threshold.restype = C.c_float
        _ORC.orc_acq_threshold.argtypes = [C.POINTER(AcqConf), C.c_float]
    return _ORC


def ref_available():
    return os.path.exists(os.path.join(ORACLE_DIR, "_ref", "libgnssref.so"))


def ref():
    """The reference's own compiled code (None when oracle/_ref was never built)."""
    global _REF
    if _REF is None:
        if not ref_available():
            return None
        _REF = C.CDLL(os.path.join(ORACLE_DIR, "_ref", "libgnssref.so"))
        _REF.ref_mc_real_codes_bench.restype = C.c_double
        for name in ("ref_pll_cloop_two_quadrant_atan", "ref_pll_four_quadrant_atan", "ref_fll_four_quadrant_atan", "ref_fll_diff_atan",
                     "ref_dll_nc_e_minus_l_normalized", "ref_dll_nc_vemlp_normalized", "ref_sinboc_slope_abs", "ref_sinboc_yintercept_abs"):
            getattr(_REF, name).restype = C.c_double
        for name in ("ref_cn0_m2m4_estimator", "ref_cn0_svn_estimator", "ref_carrier_lock_detector"):
            getattr(_REF, name).restype = C.c_float
    return _REF


# ------------------------------------------------------------------ tracking
def mc_real_codes(sig, code, shifts, rem_carr, phase_step, rem_code, code_step, n=None, high_dyn=False, phase_rate=0.0, code_rate=0.0, impl="orc", arch=0):
    """One Cpu_Multicorrelator_Real_Codes call. impl: 'orc' (restatement) | 'ref' (reference .so; arch 0 generic / 1 AVX)."""
    sig = np.ascontiguousarray(sig, dtype=np.complex64)
    code = np.ascontiguousarray(code, dtype=np.float32)
    shifts = np.ascontiguousarray(shifts, dtype=np.float32)
    n = len(sig) if n is None else n
    out = np.zeros(len(shifts), dtype=np.complex64)
    f = C.c_float
    if impl == "orc":
        rc = orc().orc_mc_real_codes(int(high_dyn), _p(sig), n, _p(code), len(code), _p(shifts), len(shifts), f(rem_carr), f(phase_step),
                                     f(phase_rate), f(rem_code), f(code_step), f(code_rate), _p(out))
    else:
        rc = ref().ref_mc_real_codes(arch, int(high_dyn), _p(sig), n, _p(code), len(code), _p(shifts), len(shifts), f(rem_carr), f(phase_step),
                                     f(phase_rate), f(rem_code), f(code_step), f(code_rate), _p(out))
    assert rc == 0
    return out


def mc_complex_codes(sig, code, shifts, rem_carr, phase_step, rem_code, code_step, impl="orc", arch=0):
    sig = np.ascontiguousarray(sig, dtype=np.complex64)
    code = np.ascontiguousarray(code, dtype=np.complex64)
    shifts = np.ascontiguousarray(shifts, dtype=np.float32)
    out = np.zeros(len(shifts), dtype=np.complex64)
    f = C.c_float
    if impl == "orc":
        rc = orc().orc_mc_complex_codes(_p(sig), len(sig), _p(code), len(code), _p(shifts), len(shifts), f(rem_carr), f(phase_step), f(rem_code), f(code_step), _p(out))
    else:
        rc = ref().ref_mc_complex_codes(arch, _p(sig), len(sig), _p(code), len(code), _p(shifts), len(shifts), f(rem_carr), f(phase_step), f(rem_code), f(code_step), _p(out))
    assert rc == 0
    return out


def mc_16sc(sig_i16, code_i16, shifts, rem_carr, phase_step, rem_code, code_step):
    """The reference's own int16 path (Cpu_Multicorrelator_16sc over the generic 16ic protokernels, oracle/_ref): interleaved int16
    (re, im) signal and code -> int16 complex outputs (saturating accumulation) as complex64."""
    sig = np.ascontiguousarray(sig_i16, dtype=np.int16)
    code = np.ascontiguousarray(code_i16, dtype=np.int16)
    shifts = np.ascontiguousarray(shifts, dtype=np.float32)
    out = np.zeros(2 * len(shifts), dtype=np.int16)
    f = C.c_float
    rc = ref().ref_mc_16sc(sig.ctypes.data_as(C.c_void_p), len(sig) // 2, code.ctypes.data_as(C.c_void_p), len(code) // 2, _p(shifts), len(shifts), f(rem_carr), f(phase_step),
                           f(rem_code), f(code_step), out.ctypes.data_as(C.c_void_p))
    assert rc == 0
    return (out[0::2].astype(np.float32) + 1j * out[1::2].astype(np.float32)).astype(np.complex64)


def mc_real_codes_f64(sig, code, shifts, rem_carr, phase_step, rem_code, code_step, mimic_inc=False):
    sig = np.ascontiguousarray(sig, dtype=np.complex64)
    code = np.ascontiguousarray(code, dtype=np.float32)
    shifts = np.ascontiguousarray(shifts, dtype=np.float32)
    out = np.zeros(len(shifts), dtype=np.complex128)
    f = C.c_float
    rc = orc().orc_mc_real_codes_f64(_p(sig), len(sig), _p(code), len(code), _p(shifts), len(shifts), f(rem_carr), f(phase_step), f(rem_code),
                                     f(code_step), int(mimic_inc), _p(out, f64p))
    assert rc == 0
    return out


def mc_batch(ring, offsets, lens, codes, shifts, rem_carr, phase_step, rem_code, code_step, nthreads=0):
    """Open-loop batch: channel c correlates ring[offsets[c] : offsets[c]+lens[c]] (scalar fp32 restatement, OpenMP over channels)."""
    ring = np.ascontiguousarray(ring, dtype=np.complex64)
    nch, L = codes.shape
    ntaps = shifts.shape[1]
    offsets = np.ascontiguousarray(offsets, dtype=np.int64)
    lens = np.ascontiguousarray(lens, dtype=np.int32)
    codes = np.ascontiguousarray(codes, dtype=np.float32)
    shifts = np.ascontiguousarray(shifts, dtype=np.float32)
    a = [np.ascontiguousarray(x, dtype=np.float32) for x in (rem_carr, phase_step, rem_code, code_step)]
    out = np.zeros((nch, ntaps), dtype=np.complex64)
    rc = orc().orc_mc_batch(_p(ring), C.c_int64(len(ring)), nch, ntaps, _p(offsets, i64p), _p(lens, i32p), _p(codes), L, _p(shifts),
                            _p(a[0]), _p(a[1]), _p(a[2]), _p(a[3]), _p(out), nthreads)
    assert rc == 0
    return out


def resampler_index(rem, step, shifts, L, n):
    shifts = np.ascontiguousarray(shifts, dtype=np.float32)
    idx = np.zeros((len(shifts), n), dtype=np.int32)
    orc().orc_resampler_index(_p(idx, i32p), C.c_float(rem), C.c_float(step), _p(shifts), L, len(shifts), n)
    return idx


# ------------------------------------------------------------------ codes
def gps_l1_ca_code(prn, chip_shift=0, impl="orc"):
    out = np.zeros(1023, dtype=np.float32)
    if impl == "orc":
        orc().orc_gps_l1_ca_code_gen_float(_p(out), prn, chip_shift)
    else:
        ref().ref_gps_l1_ca_code_gen_float(_p(out), prn, chip_shift)
    return out


def gps_l1_ca_code_sampled(prn, fs, n, chip_shift=0, impl="orc"):
    out = np.zeros(n, dtype=np.complex64)
    if impl == "orc":
        orc().orc_gps_l1_ca_code_gen_complex_sampled(_p(out), n, prn, int(fs), chip_shift)
    else:
        ref().ref_gps_l1_ca_code_gen_complex_sampled(_p(out), n, prn, int(fs), chip_shift)
    return out


# ------------------------------------------------------------------ FFT / acquisition
def fft_c32(x, inverse=False):
    x = np.ascontiguousarray(x, dtype=np.complex64)
    out = np.zeros_like(x)
    rc = orc().orc_fft_c32(_p(x), _p(out), len(x), int(inverse))
    assert rc == 0
    return out


def fft_c64(x, inverse=False):
    x = np.ascontiguousarray(x, dtype=np.complex128)
    out = np.zeros_like(x)
    rc = orc().orc_fft_c64(_p(x, f64p), _p(out, f64p), len(x), int(inverse))
    assert rc == 0
    return out


def acq_conf(fs, consumed, fft_size, doppler_max, doppler_step, samples_per_code, samples_per_chip, bit_transition=False, use_cfar=True,
             doppler_center=0, max_dwells=1, fft_f64=False):
    import math
    c = AcqConf()
    c.fft_size, c.consumed = fft_size, consumed
    c.num_doppler_bins = int(math.ceil(2.0 * doppler_max / doppler_step))
    c.bit_transition, c.use_cfar, c.samples_per_chip = int(bit_transition), int(use_cfar), samples_per_chip
    c.doppler_max, c.doppler_step, c.doppler_center, c.max_dwells = doppler_max, doppler_step, doppler_center, max_dwells
    c.samples_per_code, c.fs_in, c.fft_f64, c.sincos_arch = samples_per_code, float(fs), int(fft_f64), 0
    return c


def acq_neff(conf):
    return orc().orc_acq_neff(C.byref(conf))


def acq_set_local_code(conf, code):
    code = np.ascontiguousarray(code, dtype=np.complex64)
    out = np.zeros(conf.fft_size, dtype=np.complex64)
    rc = orc().orc_acq_set_local_code(C.byref(conf), _p(code), _p(out))
    assert rc == 0
    return out


def acq_dwell(conf, sig, fft_codes, grid=None, dwell_index=0, nthreads=0):
    sig = np.ascontiguousarray(sig, dtype=np.complex64)
    assert len(sig) >= conf.consumed
    if grid is None:
        grid = np.zeros((conf.num_doppler_bins, acq_neff(conf)), dtype=np.float32)
    rc = orc().orc_acq_dwell(C.byref(conf), _p(sig), _p(fft_codes), _p(grid), dwell_index, nthreads)
    assert rc == 0
    return grid


def acq_statistic(conf, grid, num_dwells=1):
    r = AcqResult()
    rc = orc().orc_acq_statistic(C.byref(conf), _p(grid), num_dwells, C.byref(r))
    assert rc == 0
    return r


def acq_threshold(conf, pfa):
    return float(orc().orc_acq_threshold(C.byref(conf), C.c_float(pfa)))


def acq_wipeoff(n, freq, fs):
    out = np.zeros(n, dtype=np.complex64)
    orc().orc_acq_wipeoff(_p(out), n, C.c_float(freq), C.c_double(fs))
    return out


# ------------------------------------------------------------------ closed loop (synthetic IF + the reference-object loop)
def gen_if(n, first_index, fs, codes, doppler_hz, delay_chips, phase0, cn0_dbhz, nav_bits=None, periods_per_bit=20,
           f_chip=1.023e6, f_carrier=1575.42e6, noise_sigma=1.0, seed=1):
    """complex64[n] of the stream starting at absolute sample `first_index` (oracle_loop.c: orc_gen_if); codes: (nsat, L)."""
    codes = np.ascontiguousarray(codes, dtype=np.float32)
    nsat, L = codes.shape
    d = [np.ascontiguousarray(x, dtype=np.float64) for x in (doppler_hz, delay_chips, phase0, cn0_dbhz)]
    out = np.empty(n, dtype=np.complex64)
    if nav_bits is not None:
        nav_bits = np.ascontiguousarray(nav_bits, dtype=np.int8)
        nb, nbp = nav_bits.shape[1], nav_bits.ctypes.data_as(C.c_void_p)
    else:
        nb, nbp = 0, None
    f = orc().orc_gen_if
    f.restype = None
    f.argtypes = [C.c_void_p, C.c_int64, C.c_int64, C.c_double, C.c_int, C.c_void_p, C.c_int, C.c_double, C.c_double,
                  C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_double, C.c_uint64]
    f(out.ctypes.data, n, first_index, fs, nsat, codes.ctypes.data, L, f_chip, f_carrier, d[0].ctypes.data, d[1].ctypes.data, d[2].ctypes.data,
      d[3].ctypes.data, nbp, nb, periods_per_bit, noise_sigma, seed)
    return out


class RefLoopParams(C.Structure):
    _fields_ = [(k, C.c_double) for k in ("fs_in", "code_period", "chip_rate", "carrier_freq", "carrier_lock_th")] + \
        [(k, C.c_float) for k in ("pll_bw", "dll_bw", "fll_bw", "pll_bw_narrow", "dll_bw_narrow", "el", "vel", "el_narrow", "vel_narrow",
                                  "cn0_smoother_alpha", "lock_smoother_alpha")] + \
        [(k, C.c_int32) for k in ("vector_length", "veml", "secondary", "track_pilot", "symbols_per_bit", "samples_per_chip", "code_length_chips",
                                  "pll_order", "dll_order", "extend", "cn0_samples", "cn0_min", "max_code_lock_fail", "max_carrier_lock_fail",
                                  "cn0_smoother_samples", "lock_smoother_samples", "carrier_aiding", "enable_fll_pull_in", "enable_fll_steady",
                                  "interchange_iq", "arch", "correlation_length_ms")] + \
        [(k, C.c_uint32) for k in ("pull_in_time_s", "bit_sync_time_limit_s", "secondary_code_length", "data_secondary_code_length")] + \
        [("secondary_code", C.c_char * 256), ("data_secondary_code", C.c_char * 256)]


REF_LOOP_REC = np.dtype([("channel", "<i4"), ("prn", "<u4"), ("valid", "<i4"), ("correlation_length_ms", "<i4"), ("prompt_i", "<f8"), ("prompt_q", "<f8"),
                         ("cn0_db_hz", "<f8"), ("carrier_doppler_hz", "<f8"), ("carrier_phase_rads", "<f8"), ("code_phase_samples", "<f8"),
                         ("tracking_sample_counter", "<u8")])


class RefLoop:
    """One tracking channel made of the reference's own objects (oracle/ref_loop.cc in oracle/_ref/libgnssref.so)."""

    def __init__(self, params, tracking_code, data_code, prn, acq_delay_samples, acq_doppler_hz, acq_samplestamp, nitems_read):
        r = ref()
        r.ref_loop_new.restype = C.c_void_p
        r.ref_loop_new.argtypes = [C.POINTER(RefLoopParams), C.c_void_p, C.c_void_p, C.c_int, C.c_uint, C.c_double, C.c_double, C.c_uint64, C.c_uint64]
        r.ref_loop_feed.restype = C.c_int
        r.ref_loop_feed.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int]
        r.ref_loop_state.argtypes = [C.c_void_p, C.c_void_p]
        r.ref_loop_free.argtypes = [C.c_void_p]
        self._r = r
        t = np.ascontiguousarray(tracking_code, dtype=np.float32)
        d = None if data_code is None else np.ascontiguousarray(data_code, dtype=np.float32)
        self._h = r.ref_loop_new(C.byref(params), t.ctypes.data, None if d is None else d.ctypes.data, len(t), prn, acq_delay_samples, acq_doppler_hz,
                                 acq_samplestamp, nitems_read)
        self._rec = np.zeros(4096, dtype=REF_LOOP_REC)

    def feed(self, sig):
        sig = np.ascontiguousarray(sig, dtype=np.complex64)
        n = self._r.ref_loop_feed(self._h, sig.ctypes.data, len(sig), self._rec.ctypes.data, len(self._rec))
        return self._rec[:n].copy()

    def state(self):
        v = np.zeros(8)
        self._r.ref_loop_state(self._h, v.ctypes.data)
        return dict(zip(["state", "carrier_doppler_hz", "cn0_db_hz", "carrier_lock_test", "code_freq_chips", "rem_code_phase_samples", "nitems_read", "prn_length"], v))

    def close(self):
        if self._h:
            self._r.ref_loop_free(self._h)
            self._h = None


def unpack_2bit(raw, item_size=1, big_endian_items=False, big_endian_bytes=False, reverse_interleaving=False):
    """unpack_2bit_samples::work on a byte array -> int8 array of 4 values per byte."""
    raw = np.ascontiguousarray(raw).view(np.int8).ravel()
    out = np.empty(4 * raw.size, np.int8)
    f = orc().orc_unpack_2bit
    f.restype = None
    f.argtypes = [C.c_void_p, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]
    f(raw.ctypes.data, raw.size, item_size, int(big_endian_items), int(big_endian_bytes), int(reverse_interleaving), out.ctypes.data)
    return out


def labsat_one_channel(raw, bits):
    """labsat23_source::decode_samples_one_channel over a payload of native int16 words -> complex64 (8 or 4 per word)."""
    w = np.ascontiguousarray(raw).view(np.int16).ravel()
    out = np.empty(w.size * (8 if bits == 2 else 4), np.complex64)
    f = orc().orc_labsat_one_channel
    f.restype = None
    f.argtypes = [C.c_void_p, C.c_int64, C.c_int, C.c_void_p]
    f(w.ctypes.data, w.size, bits, out.ctypes.data)
    return out


def items_to_complex(raw, kind):
    """File item types -> complex64 (GB200_SRC_* 0..4: gr_complex, ishort, ibyte, short, byte), no scaling."""
    raw = np.ascontiguousarray(raw)
    per = {0: 8, 1: 4, 2: 2, 3: 2, 4: 1}[kind]
    n = raw.nbytes // per
    out = np.empty(n, np.complex64)
    f = orc().orc_items_to_complex
    f.restype = None
    f.argtypes = [C.c_void_p, C.c_int64, C.c_int, C.c_void_p]
    f(raw.ctypes.data, n, kind, out.ctypes.data)
    return out


# ---------------------------------------------------------------------------------------------------------------- the reference's own acquisition block
class RefAcqStatus(C.Structure):
    _fields_ = [("state", C.c_int32), ("active", C.c_int32), ("positive_acq", C.c_int32), ("n_events", C.c_int32), ("events", C.c_int32 * 8),
                ("fsm_events", C.c_int32 * 8), ("n_fsm_events", C.c_int32),
                ("fft_size", C.c_uint32), ("consumed_samples", C.c_uint32), ("num_doppler_bins", C.c_uint32), ("dwells_counter", C.c_uint32), ("samples_per_chip", C.c_uint32),
                ("threshold", C.c_float), ("mag", C.c_float), ("input_power", C.c_float), ("test_statistics", C.c_float), ("samples_per_code", C.c_float),
                ("samples_per_ms", C.c_float), ("acq_delay_samples", C.c_double), ("acq_doppler_hz", C.c_double),
                ("acq_samplestamp_samples", C.c_uint64), ("sample_counter", C.c_uint64), ("consumed_total", C.c_int64),
                ("doppler_step", C.c_int32), ("doppler_max", C.c_int32), ("use_cfar", C.c_int32), ("bit_transition", C.c_int32), ("it_size", C.c_int32),
                ("step_two", C.c_int32), ("pad_", C.c_int32)]


class RefAcq:
    """pcps_acquisition of the reference (oracle/ref_acq.cc: the reference's own pcps_acquisition.cc + acq_conf.cc compiled unmodified into
    oracle/_ref against the stand-ins of oracle/shim/acq), driven like Channel and the adapters drive it."""

    def __init__(self, conf_text, role, ms_per_code, chip_rate, opt_freq, prn, system, signal, fft_f64=False):
        r = ref()
        r.ref_acq_new.restype = C.c_void_p
        r.ref_acq_new.argtypes = [C.c_char_p, C.c_char_p, C.c_uint, C.c_double, C.c_double, C.c_uint, C.c_char_p, C.c_char_p]
        r.ref_acq_free.argtypes = [C.c_void_p]
        r.ref_acq_init.argtypes = [C.c_void_p, C.c_uint, C.c_uint, C.c_float]
        r.ref_acq_set_threshold.argtypes = [C.c_void_p, C.c_float]
        r.ref_acq_set_doppler_center.argtypes = [C.c_void_p, C.c_int]
        r.ref_acq_set_local_code.argtypes = [C.c_void_p, C.c_void_p]
        r.ref_acq_arm.argtypes = [C.c_void_p]
        r.ref_acq_work.restype = C.c_long
        r.ref_acq_work.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
        r.ref_acq_get_status.argtypes = [C.c_void_p, C.POINTER(RefAcqStatus)]
        r.ref_acq_read_grid.restype = C.c_int
        r.ref_acq_read_grid.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int]
        r.ref_acq_fft_precision.argtypes = [C.c_int]
        self._r = r
        r.ref_acq_fft_precision(int(fft_f64))
        self._h = r.ref_acq_new(conf_text.encode(), role.encode(), ms_per_code, chip_rate, opt_freq, prn, system.encode(), signal.encode())

    def close(self):
        if self._h:
            self._r.ref_acq_free(self._h)
            self._h = None

    def init(self, doppler_max, doppler_step, threshold=0.0):
        self._r.ref_acq_init(self._h, int(doppler_max), int(doppler_step), float(threshold))

    def set_threshold(self, t):
        self._r.ref_acq_set_threshold(self._h, float(t))

    def set_local_code(self, code):
        """code: complex64, at least vector_length samples (what the adapter's code_ buffer holds)"""
        self._code = np.ascontiguousarray(code, dtype=np.complex64)
        self._r.ref_acq_set_local_code(self._h, self._code.ctypes.data)

    def arm(self):
        self._r.ref_acq_arm(self._h)

    def status(self):
        s = RefAcqStatus()
        self._r.ref_acq_get_status(self._h, C.byref(s))
        return s

    def work(self, items, scalars_per_item, chunk=None):
        """general_work over a stream of items until the block stops (positive / negative acquisition) or the stream ends; returns items consumed.
        chunk: at most that many items are offered per scheduler call (what gb200h_acq_work's `chunk` is on the product side); None = all that is left"""
        items = np.ascontiguousarray(items)
        n_items, isz = len(items) // scalars_per_item, items.itemsize * scalars_per_item
        pos, idle = 0, 0
        for _ in range(100000):
            st = self.status()
            if st.state == 0 and not st.active and not st.step_two:  # (between the two steps of make_two_steps the block is inactive for one call)
                break
            # (the block switches from buffering to searching on the call AFTER the one that filled its buffer, :985-989, so a call with
            # no new items is part of the sequence; two calls in a row without any effect end the loop)
            c = self._r.ref_acq_work(self._h, items.ctypes.data + pos * isz, n_items - pos if chunk is None else min(chunk, n_items - pos))
            pos += c
            st2 = self.status()
            idle = idle + 1 if (c == 0 and st2.state == st.state and st2.active == st.active) else 0
            if idle >= 2:
                break
        return pos

    def grid(self, bins, neff):
        g = np.zeros((bins, neff), dtype=np.float32)
        n = self._r.ref_acq_read_grid(self._h, g.ctypes.data, bins, neff)
        return g[:n]


class RefAcqFdStatus(C.Structure):
    """ref_acqfd_status (oracle/ref_acqfd.cc)"""
    _fields_ = [("acq_delay_samples", C.c_double), ("acq_doppler_hz", C.c_double), ("acq_samplestamp_samples", C.c_uint64), ("sample_counter", C.c_uint64),
                ("consumed_total", C.c_int64), ("test_statistics", C.c_float), ("threshold", C.c_float), ("state", C.c_int32), ("active", C.c_int32),
                ("positive_acq", C.c_int32), ("num_doppler_points", C.c_int32), ("fft_size", C.c_int32), ("well_count", C.c_int32), ("n_events", C.c_int32),
                ("events", C.c_int32 * 8), ("acq_doppler_step", C.c_uint32)]


class RefAcqFd:
    """pcps_acquisition_fine_doppler_cc of the reference (oracle/ref_acqfd.cc: the reference's own block compiled unmodified into oracle/_ref),
    filled and driven like adapters/gps_l1_ca_pcps_acquisition_fine_doppler.cc does"""

    def __init__(self, fs_in, doppler_max, max_dwells, prn, blocking_on_standby=False):
        r = ref()
        r.ref_acqfd_new.restype = C.c_void_p
        r.ref_acqfd_new.argtypes = [C.c_longlong, C.c_int, C.c_int, C.c_int, C.c_uint]
        r.ref_acqfd_free.argtypes = [C.c_void_p]
        r.ref_acqfd_arm.argtypes = [C.c_void_p, C.c_uint, C.c_uint, C.c_float]
        r.ref_acqfd_work.restype = C.c_longlong
        r.ref_acqfd_work.argtypes = [C.c_void_p, C.c_void_p, C.c_longlong, C.c_int]
        r.ref_acqfd_get_status.argtypes = [C.c_void_p, C.POINTER(RefAcqFdStatus)]
        self._r = r
        r.ref_acq_fft_precision.argtypes = [C.c_int]
        r.ref_acq_fft_precision(0)
        self._h = r.ref_acqfd_new(int(fs_in), int(doppler_max), int(max_dwells), int(blocking_on_standby), int(prn))

    def close(self):
        if self._h:
            self._r.ref_acqfd_free(self._h)
            self._h = None

    def arm(self, doppler_max, doppler_step, threshold):
        self._r.ref_acqfd_arm(self._h, int(doppler_max), int(doppler_step), float(threshold))

    def work(self, items, noutput_items):
        items = np.ascontiguousarray(items, dtype=np.complex64)
        return int(self._r.ref_acqfd_work(self._h, items.ctypes.data, items.size, int(noutput_items)))

    def status(self):
        s = RefAcqFdStatus()
        self._r.ref_acqfd_get_status(self._h, C.byref(s))
        return s


class RefAcqCafStatus(C.Structure):
    """ref_acqcaf_status (oracle/ref_acqcaf.cc)"""
    _fields_ = [("acq_delay_samples", C.c_double), ("acq_doppler_hz", C.c_double), ("acq_samplestamp_samples", C.c_uint64), ("sample_counter", C.c_uint64),
                ("consumed_total", C.c_int64), ("test_statistics", C.c_float), ("threshold", C.c_float), ("mag", C.c_float), ("input_power", C.c_float),
                ("state", C.c_int32), ("active", C.c_int32), ("num_doppler_bins", C.c_int32), ("fft_size", C.c_int32), ("well_count", C.c_int32), ("n_events", C.c_int32),
                ("events", C.c_int32 * 8), ("acq_doppler_step", C.c_uint32)]


class RefAcqCaf:
    """galileo_e5a_noncoherentIQ_acquisition_caf_cc of the reference (oracle/ref_acqcaf.cc: the reference's own block compiled unmodified into
    oracle/_ref), made and driven like adapters/galileo_e5a_noncoherent_iq_acquisition_caf.cc does"""

    def __init__(self, fs_in, sampled_ms, max_dwells, doppler_max, prn, signal="5X", channel_signal="5X", bit_transition=False, caf_window_hz=0, zero_padding=0):
        r = ref()
        r.ref_acqcaf_new.restype = C.c_void_p
        r.ref_acqcaf_new.argtypes = [C.c_longlong, C.c_uint, C.c_uint, C.c_uint, C.c_int, C.c_int, C.c_int, C.c_uint, C.c_char_p, C.c_char_p]
        r.ref_acqcaf_free.argtypes = [C.c_void_p]
        r.ref_acqcaf_arm.argtypes = [C.c_void_p, C.c_uint, C.c_uint, C.c_float]
        r.ref_acqcaf_work.restype = C.c_longlong
        r.ref_acqcaf_work.argtypes = [C.c_void_p, C.c_void_p, C.c_longlong, C.c_int]
        r.ref_acqcaf_get_status.argtypes = [C.c_void_p, C.POINTER(RefAcqCafStatus)]
        r.ref_acqcaf_caf_vector.restype = C.c_int
        r.ref_acqcaf_caf_vector.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
        self._r = r
        r.ref_acq_fft_precision.argtypes = [C.c_int]
        r.ref_acq_fft_precision(0)
        self._h = r.ref_acqcaf_new(int(fs_in), int(sampled_ms), int(max_dwells), int(doppler_max), int(bit_transition), int(caf_window_hz), int(zero_padding), int(prn),
                                   signal.encode(), channel_signal.encode())

    def close(self):
        if self._h:
            self._r.ref_acqcaf_free(self._h)
            self._h = None

    def arm(self, doppler_max, doppler_step, threshold):
        self._r.ref_acqcaf_arm(self._h, int(doppler_max), int(doppler_step), float(threshold))

    def work(self, items, chunk):
        items = np.ascontiguousarray(items, dtype=np.complex64)
        return int(self._r.ref_acqcaf_work(self._h, items.ctypes.data, items.size, int(chunk)))

    def status(self):
        s = RefAcqCafStatus()
        self._r.ref_acqcaf_get_status(self._h, C.byref(s))
        return s

    def caf_vector(self):
        n = self._r.ref_acqcaf_caf_vector(self._h, None, 0)
        out = np.zeros(n, np.float32)
        if n:
            self._r.ref_acqcaf_caf_vector(self._h, out.ctypes.data, n)
        return out


class RefAcqQsStatus(C.Structure):
    """ref_acqqs_status (oracle/ref_acqqs.cc)"""
    _fields_ = [("acq_delay_samples", C.c_double), ("acq_doppler_hz", C.c_double), ("acq_samplestamp_samples", C.c_uint64), ("sample_counter", C.c_uint64),
                ("consumed_total", C.c_int64), ("test_statistics", C.c_float), ("threshold", C.c_float), ("mag", C.c_float), ("input_power", C.c_float),
                ("state", C.c_int32), ("active", C.c_int32), ("num_doppler_bins", C.c_int32), ("fft_size", C.c_int32), ("well_count", C.c_int32), ("item_length", C.c_int32),
                ("folding_factor", C.c_int32), ("sampled_ms", C.c_int32), ("n_events", C.c_int32), ("events", C.c_int32 * 8), ("acq_doppler_step", C.c_uint32)]


class RefAcqQs:
    """pcps_quicksync_acquisition_cc of the reference (oracle/ref_acqqs.cc: the reference's own block compiled unmodified into oracle/_ref), made
    and driven like adapters/gps_l1_ca_pcps_quicksync_acquisition.cc does; folding_factor 0 = the adapter's default"""

    def __init__(self, fs_in, folding_factor, sampled_ms, max_dwells, doppler_max, prn, bit_transition=False, galileo_signal=None, cboc=False):
        r = ref()
        r.ref_acqqs_new_e1.restype = C.c_void_p
        r.ref_acqqs_new_e1.argtypes = [C.c_longlong, C.c_uint, C.c_uint, C.c_uint, C.c_uint, C.c_int, C.c_uint, C.c_char_p, C.c_int]
        r.ref_acqqs_new.restype = C.c_void_p
        r.ref_acqqs_new.argtypes = [C.c_longlong, C.c_uint, C.c_uint, C.c_uint, C.c_uint, C.c_int, C.c_uint]
        r.ref_acqqs_free.argtypes = [C.c_void_p]
        r.ref_acqqs_arm.argtypes = [C.c_void_p, C.c_uint, C.c_uint, C.c_float]
        r.ref_acqqs_work.restype = C.c_longlong
        r.ref_acqqs_work.argtypes = [C.c_void_p, C.c_void_p, C.c_longlong, C.c_int]
        r.ref_acqqs_get_status.argtypes = [C.c_void_p, C.POINTER(RefAcqQsStatus)]
        self._r = r
        r.ref_acq_fft_precision.argtypes = [C.c_int]
        r.ref_acq_fft_precision(0)
        if galileo_signal:  # the Galileo E1 adapter over the same block
            self._h = r.ref_acqqs_new_e1(int(fs_in), int(folding_factor), int(sampled_ms), int(max_dwells), int(doppler_max), int(bit_transition), int(prn), galileo_signal.encode(), int(cboc))
        else:
            self._h = r.ref_acqqs_new(int(fs_in), int(folding_factor), int(sampled_ms), int(max_dwells), int(doppler_max), int(bit_transition), int(prn))

    def close(self):
        if self._h:
            self._r.ref_acqqs_free(self._h)
            self._h = None

    def arm(self, doppler_max, doppler_step, threshold):
        self._r.ref_acqqs_arm(self._h, int(doppler_max), int(doppler_step), float(threshold))

    def work(self, samples, max_items):
        """returns the vector ITEMS consumed"""
        samples = np.ascontiguousarray(samples, dtype=np.complex64)
        return int(self._r.ref_acqqs_work(self._h, samples.ctypes.data, samples.size, int(max_items)))

    def status(self):
        s = RefAcqQsStatus()
        self._r.ref_acqqs_get_status(self._h, C.byref(s))
        return s


# ---------------------------------------------------------------------------------------------------------------- the reference's own tracking block
class RefTrkBlock:
    """dll_pll_veml_tracking of the reference (oracle/ref_trkblk.cc: the reference's own dll_pll_veml_tracking.cc + dll_pll_conf.cc compiled
    unmodified into oracle/_ref, the door playing the GNU Radio scheduler).  Same feed() / state() as RefLoop; the block generates its own
    code tables from (system, signal, PRN) in start_tracking."""
    SIGNALS = {"1C": ("G", "1C", 1.023e6, 1023.0), "1B": ("E", "1B", 1.023e6, 1023.0), "L5": ("G", "L5", 10.23e6, 10230.0), "5X": ("E", "5X", 10.23e6, 10230.0)}

    def __init__(self, conf_text, role, signal, prn, acq_delay_samples, acq_doppler_hz, acq_samplestamp=0):
        r = ref()
        r.ref_trkblk_new.restype = C.c_void_p
        r.ref_trkblk_new.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p, C.c_char_p, C.c_double, C.c_double, C.c_uint, C.c_double, C.c_double, C.c_uint64]
        r.ref_trkblk_feed.restype = C.c_int
        r.ref_trkblk_feed.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int]
        r.ref_trkblk_state.argtypes = [C.c_void_p, C.c_void_p]
        r.ref_trkblk_free.argtypes = [C.c_void_p]
        self._r = r
        system, sig, chip_rate, code_len = self.SIGNALS[signal]
        # vector_length = round(fs / (chip_rate / code_length)): 1 ms for GPS L1 / L5 / E5a; the E1 adapter uses the 4 ms code (4092 chips)
        if signal == "1B":
            code_len = 4092.0
        self._h = r.ref_trkblk_new(conf_text.encode(), role.encode(), system.encode(), sig.encode(), chip_rate, code_len, int(prn), float(acq_delay_samples),
                                   float(acq_doppler_hz), int(acq_samplestamp))
        self._rec = np.zeros(8192, dtype=REF_LOOP_REC)

    def feed(self, sig):
        sig = np.ascontiguousarray(sig, dtype=np.complex64)
        n = self._r.ref_trkblk_feed(self._h, sig.ctypes.data, len(sig), self._rec.ctypes.data, len(self._rec))
        return self._rec[:n].copy()

    def state(self):
        out = np.zeros(6, dtype=np.float64)
        self._r.ref_trkblk_state(self._h, out.ctypes.data)
        return dict(state=int(out[0]), cn0_db_hz=float(out[1]), carrier_doppler_hz=float(out[2]), code_freq_chips=float(out[3]), carrier_lock_test=float(out[4]),
                    last_event=int(out[5]))

    def close(self):
        if self._h:
            self._r.ref_trkblk_free(self._h)
            self._h = None
