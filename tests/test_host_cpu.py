"""Host C++ layer (gnss_sim_receiver_b200/host/), CPU-only checks:

* every scalar loop-closure piece (discriminators, loop filters, lock detectors, smoother, sinBOC slope) bit-for-bit
  against the reference's own compiled objects (oracle/_ref) wherever that library exists, and against the expected
  values of the reference's unit tests (tracking_loop_filter_test.cc:22-204, discriminator_test.cc:35-105) everywhere;
* the GPS L1 C/A generators bit-exact against the golden PRN tables produced by the reference's generators;
* .conf parsing, Dll_Pll_Conf / Acq_Conf defaults, clamps and derived parameters (dll_pll_conf.cc:36-125,
  acq_conf.cc:24-118) on the reference's own conf/gnss-sdr_GPS_L1_gr_complex.conf key set;
* the `.implementation` string chain (gnss_block_factory.cc pattern) and the library's exported symbols.
"""
import ctypes as C
import os
import re

import numpy as np
import pytest

import _oracle as O
from gnss_sim_receiver_b200 import _hostapi as H

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = np.load(os.path.join(ROOT, "tests", "golden", "ref_vectors.npz"))
needs_ref = pytest.mark.skipif(not O.ref_available(), reason="oracle/_ref not built (needs /root/reference)")

CONF_C1 = """
; shape of conf/gnss-sdr_GPS_L1_gr_complex.conf with the B200 implementation strings
[GNSS-SDR]
GNSS-SDR.internal_fs_sps=4000000
Channels_1C.count=8
Acquisition_1C.implementation=GPS_L1_CA_PCPS_Acquisition_B200
Acquisition_1C.item_type=gr_complex
Acquisition_1C.coherent_integration_time_ms=1
Acquisition_1C.pfa=0.01
Acquisition_1C.doppler_max=10000
Acquisition_1C.doppler_step=250
Acquisition_1C.max_dwells=1
Tracking_1C.implementation=GPS_L1_CA_DLL_PLL_Tracking_B200
Tracking_1C.item_type=gr_complex
Tracking_1C.pll_bw_hz=40.0;
Tracking_1C.dll_bw_hz=4.0 ; inline comment
Tracking_1C.order=3;
Tracking_1C.dump=false
"""


def test_every_declared_symbol_is_exported():
    lib = H.lib()
    hdr = open(os.path.join(ROOT, "gnss_sim_receiver_b200", "host", "b200_host_capi.h")).read()
    declared = set(re.findall(r"\b(gb200h_[a-z0-9_]+)\s*\(", hdr))
    assert declared == set(H.SYMBOLS), declared ^ set(H.SYMBOLS)
    for name in declared:
        getattr(lib, name)


def test_conf_reader():
    lib = H.lib()
    g = lambda k, d="": lib.gb200h_conf_get(CONF_C1.encode(), k.encode(), d.encode()).decode()
    assert g("GNSS-SDR.internal_fs_sps") == "4000000"
    assert g("Tracking_1C.pll_bw_hz") == "40.0;"[:-1] or g("Tracking_1C.pll_bw_hz") == "40.0;"  # trailing ';' without a space is part of the value in inih
    assert g("Tracking_1C.dll_bw_hz") == "4.0"
    assert g("Missing.key", "dflt") == "dflt"


def test_dll_pll_conf_defaults_and_keys():
    kv = H.kv(H.lib().gb200h_dll_pll_conf(CONF_C1.encode(), b"Tracking_1C"))
    assert float(kv["fs_in"]) == 4e6 and float(kv["pll_bw_hz"]) == 40.0 and float(kv["dll_bw_hz"]) == 4.0
    # defaults of dll_pll_conf.h:38-87
    assert float(kv["pll_bw_narrow_hz"]) == 5.0 and float(kv["dll_bw_narrow_hz"]) == 0.75 and float(kv["fll_bw_hz"]) == 35.0
    assert float(kv["early_late_space_chips"]) == 0.25 and float(kv["very_early_late_space_chips"]) == 0.5
    assert int(kv["pll_filter_order"]) == 3 and int(kv["dll_filter_order"]) == 2 and int(kv["fll_filter_order"]) == 2
    assert int(kv["pull_in_time_s"]) == 10 and int(kv["bit_synchronization_time_limit_s"]) == 20 and int(kv["smoother_length"]) == 10
    # clamps (dll_pll_conf.cc:64-95)
    kv = H.kv(H.lib().gb200h_dll_pll_conf(b"T.dll_filter_order=7\nT.pll_filter_order=1\nT.item_type=bogus\nT.smoother_length=0\n", b"T"))
    assert int(kv["dll_filter_order"]) == 3 and int(kv["pll_filter_order"]) == 2 and int(kv["fll_filter_order"]) == 1
    assert kv["item_type"] == "gr_complex" and int(kv["smoother_length"]) == 1


def test_acq_conf_derived_params():
    kv = H.kv(H.lib().gb200h_acq_conf(CONF_C1.encode(), b"Acquisition_1C", 1.023e6, 1e6, 1))
    assert int(kv["fs_in"]) == 4000000 and int(kv["doppler_max"]) == 10000 and float(kv["doppler_step"]) == 250.0
    # acq_conf.cc:113-117 computes float(fs) * 0.001F: 4000.00024 in fp32, not 4000 — kept, fmod(indext, samples_per_code) depends on it
    spm = float(np.float32(4e6) * np.float32(0.001))
    assert np.float32(kv["samples_per_ms"]) == np.float32(spm) and int(kv["samples_per_chip"]) == 4 and np.float32(kv["samples_per_code"]) == np.float32(spm)
    assert int(kv["use_CFAR_algorithm_flag"]) == 1 and int(kv["it_size"]) == 8
    # pfa <= 0 disables CFAR; pfa2 falls back to pfa (acq_conf.cc:73-90)
    kv = H.kv(H.lib().gb200h_acq_conf(b"GNSS-SDR.internal_fs_sps=25000000\nA.item_type=cshort\n", b"A", 1.023e6, 1e6, 1))
    assert int(kv["use_CFAR_algorithm_flag"]) == 0 and int(kv["it_size"]) == 4 and int(kv["samples_per_chip"]) == 25
    # the per-signal constants the four adapters hand to SetFromConfiguration (chip rate, optimum resampled rate, ms per code):
    # Galileo E1 at 4 MS/s (C2): 4 ms code -> sampled_ms snaps to a multiple of 4, samples_per_code = 16000, 4 samples per chip
    kv = H.kv(H.lib().gb200h_acq_conf(b"GNSS-SDR.internal_fs_sps=4000000\nA.coherent_integration_time_ms=4\nA.bit_transition_flag=true\n", b"A", 1.023e6, 2e6, 4))
    assert int(kv["sampled_ms"]) == 4 and abs(float(kv["samples_per_code"]) - 16000.0) < 0.01 and int(kv["samples_per_chip"]) == 4 and int(kv["bit_transition_flag"]) == 1
    kv = H.kv(H.lib().gb200h_acq_conf(b"GNSS-SDR.internal_fs_sps=4000000\nA.coherent_integration_time_ms=3\n", b"A", 1.023e6, 2e6, 4))
    assert int(kv["sampled_ms"]) == 4  # not a multiple of ms_per_code: reset to one code (acq_conf.cc:48-53)
    # GPS L5 / Galileo E5a at 20 MS/s (C4): 10.23 Mchip/s -> 2 samples per chip, 20000 samples per code, 20 dwells
    kv = H.kv(H.lib().gb200h_acq_conf(b"GNSS-SDR.internal_fs_sps=20000000\nA.max_dwells=20\nA.pfa=0.01\n", b"A", 10.23e6, 1e7, 1))
    assert int(kv["samples_per_chip"]) == 2 and abs(float(kv["samples_per_code"]) - 20000.0) < 0.01 and int(kv["max_dwells"]) == 20 and int(kv["use_CFAR_algorithm_flag"]) == 1
    # automatic resampler: fs 20 MS/s over the 10 MS/s optimum -> decimation 2 (acq_conf.cc:92-107)
    kv = H.kv(H.lib().gb200h_acq_conf(b"GNSS-SDR.internal_fs_sps=20000000\nGNSS-SDR.use_acquisition_resampler=true\n", b"A", 10.23e6, 1e7, 1))
    assert int(kv["resampled_fs"]) == 10000000 and float(kv["resampler_ratio"]) == 2.0
    # unknown item type throws std::invalid_argument like the reference (acq_conf.cc:28-31)
    kv = H.kv(H.lib().gb200h_acq_conf(b"A.item_type=bogus\n", b"A", 1.023e6, 1e6, 1))
    assert "invalid_argument" in kv


def test_factory_strings():
    f = lambda conf, role: H.lib().gb200h_factory(conf.encode(), role.encode()).decode()
    assert f(CONF_C1, "Tracking_1C").startswith("GPS_L1_CA_DLL_PLL_Tracking_B200;item_size=8;role=Tracking_1C")
    assert f(CONF_C1, "Acquisition_1C").startswith("GPS_L1_CA_PCPS_Acquisition_B200")
    assert f("T.implementation=Galileo_E1_DLL_PLL_VEML_Tracking_B200\nT.item_type=cshort\n", "T").startswith("Galileo_E1_DLL_PLL_VEML_Tracking_B200;item_size=4")
    # the four acquisition adapters (gnss_block_factory.cc GetAcqBlock chain): GPS L1 C/A, Galileo E1, GPS L5I, Galileo E5a
    for impl in ("GPS_L1_CA_PCPS_Acquisition_B200", "Galileo_E1_PCPS_Ambiguous_Acquisition_B200", "GPS_L5i_PCPS_Acquisition_B200", "Galileo_E5a_Pcps_Acquisition_B200"):
        assert f(f"A.implementation={impl}\n", "A").startswith(impl + ";acquisition")
    assert f("T.implementation=GPS_L1_CA_DLL_PLL_Tracking\n", "T") == ""  # the reference's own string is not ours
    # unknown item type -> item_size 0, which makes Channel::connect throw (channel.cc:118-122); Dll_Pll_Conf first resets it to gr_complex
    assert "item_size=8" in f("T.implementation=GPS_L5_DLL_PLL_Tracking_B200\nT.item_type=bogus\n", "T")


# ---- loop filter: expected values of the reference's own unit tests
@pytest.mark.parametrize("order,last,expected,tol", [
    (1, 0, [0.0, 0.0, 20.0, 0.0, 0.0, 0.0], 1e-6),
    (1, 1, [0.0, 0.0, 0.01, 0.02, 0.02, 0.02], 1e-4),
    (2, 0, [0.0, 0.0, 13.37778, 0.0889, 0.0889, 0.0889], 1e-4),
    (2, 1, [0.0, 0.0, 0.006689, 0.013422, 0.013511, 0.013600], 1e-4),
    (3, 0, [0.0, 0.0, 15.31877, 0.04494, 0.04520, 0.04546], 1e-4),
    (3, 1, [0.0, 0.0, 0.007659, 0.015341, 0.015386, 0.015432], 1e-4),
])
def test_tracking_loop_filter_reference_unit_test_values(order, last, expected, tol):
    x = np.array([0, 0, 1, 0, 0, 0], dtype=np.float32)
    y = np.zeros_like(x)
    H.lib().gb200h_tracking_loop_filter_run(0.001, 5.0, order, last, 0.0, x.ctypes.data, y.ctypes.data, len(x))
    assert np.allclose(y, expected, atol=tol), y


def test_bpsk_discriminator_reference_unit_test_property():
    # discriminator_test.cc:35-70: E-L normalised on a BPSK triangle returns the code error inside the linear region
    lib = H.lib()
    tri = lambda t: max(0.0, 1.0 - abs(t))
    for spacing in (0.5, 0.25, 0.1, 0.01):
        for A in (1 + 0j, -1 + 0j, 1j, 1 + 1j):
            for err in np.arange(-0.5, 0.5, 0.05):
                if abs(err) >= spacing:
                    continue
                E, L = A * np.float32(tri(err - spacing)), A * np.float32(tri(err + spacing))
                out = lib.gb200h_dll_nc_e_minus_l_normalized(E.real, E.imag, L.real, L.imag, spacing, 1.0, 1.0)
                assert abs(out - err) < 1e-4, (spacing, A, err, out)


@needs_ref
def test_scalar_pieces_bit_exact_vs_reference_objects():
    lib, ref = H.lib(), O.ref()
    f, d = C.c_float, C.c_double
    ref.ref_fll_four_quadrant_atan.argtypes = ref.ref_fll_diff_atan.argtypes = [f, f, f, f, d, d]
    ref.ref_pll_cloop_two_quadrant_atan.argtypes = ref.ref_pll_four_quadrant_atan.argtypes = [f, f]
    ref.ref_dll_nc_e_minus_l_normalized.argtypes = [f] * 7
    ref.ref_dll_nc_vemlp_normalized.argtypes = [f] * 8
    ref.ref_sinboc_slope_abs.argtypes = ref.ref_sinboc_yintercept_abs.argtypes = [d]
    rng = np.random.default_rng(5)
    for _ in range(300):
        v = rng.standard_normal(8).astype(np.float32) * rng.choice([1e-3, 1.0, 1e4])
        a = [float(x) for x in v]
        assert lib.gb200h_pll_cloop_two_quadrant_atan(a[0], a[1]) == ref.ref_pll_cloop_two_quadrant_atan(a[0], a[1])
        assert lib.gb200h_pll_four_quadrant_atan(a[0], a[1]) == ref.ref_pll_four_quadrant_atan(a[0], a[1])
        assert lib.gb200h_fll_diff_atan(a[0], a[1], a[2], a[3], 0.0, 0.001) == ref.ref_fll_diff_atan(a[0], a[1], a[2], a[3], 0.0, 0.001)
        assert lib.gb200h_fll_four_quadrant_atan(a[0], a[1], a[2], a[3], 0.0, 0.004) == ref.ref_fll_four_quadrant_atan(a[0], a[1], a[2], a[3], 0.0, 0.004)
        assert lib.gb200h_dll_nc_e_minus_l_normalized(a[0], a[1], a[2], a[3], 0.25, 1.0, 1.0) == ref.ref_dll_nc_e_minus_l_normalized(a[0], a[1], a[2], a[3], 0.25, 1.0, 1.0)
        assert lib.gb200h_dll_nc_vemlp_normalized(*a) == ref.ref_dll_nc_vemlp_normalized(*a)
    assert lib.gb200h_pll_cloop_two_quadrant_atan(0.0, 1.0) == 0.0
    for x in (0.05, 0.15, 0.25, 0.3, 0.5, 0.75):
        assert lib.gb200h_sinboc_slope_abs(x) == ref.ref_sinboc_slope_abs(x)
        assert lib.gb200h_sinboc_yintercept_abs(x) == ref.ref_sinboc_yintercept_abs(x)


@needs_ref
def test_estimators_and_filters_bit_exact_vs_reference_objects():
    lib, ref = H.lib(), O.ref()
    rng = np.random.default_rng(6)
    for n, amp in ((20, 50.0), (100, 5.0), (1000, 0.3), (7, 0.0)):
        p = ((amp + rng.standard_normal(n)) * rng.choice([-1, 1], n) + 1j * rng.standard_normal(n)).astype(np.complex64)
        for name in ("cn0_m2m4_estimator", "cn0_svn_estimator"):
            a = getattr(lib, "gb200h_" + name)(p.ctypes.data, n, 0.001)
            getattr(ref, "ref_" + name).argtypes = [C.c_void_p, C.c_int, C.c_float]
            b = getattr(ref, "ref_" + name)(p.ctypes.data, n, 0.001)
            assert a == b or (np.isnan(a) and np.isnan(b)), (name, n, a, b)
        ref.ref_carrier_lock_detector.argtypes = [C.c_void_p, C.c_int]
        assert lib.gb200h_carrier_lock_detector(p.ctypes.data, 1) == ref.ref_carrier_lock_detector(p.ctypes.data, 1)
        assert lib.gb200h_carrier_lock_detector(p.ctypes.data, n) == ref.ref_carrier_lock_detector(p.ctypes.data, n)
    x = rng.standard_normal(500).astype(np.float32) * 0.1
    x2 = rng.standard_normal(500).astype(np.float32) * 0.05
    vp, f, i = C.c_void_p, C.c_float, C.c_int
    ref.ref_tracking_loop_filter_run.argtypes = [f, f, i, i, f, vp, vp, i]
    ref.ref_fll_pll_filter_run.argtypes = [f, f, i, f, vp, vp, f, vp, i]
    ref.ref_exponential_smoother_run.argtypes = [f, i, f, f, vp, vp, i]
    for order in (1, 2, 3):
        for last in (0, 1):
            for T, bw in ((0.001, 2.0), (0.004, 0.75), (0.02, 15.0)):
                a, b = np.zeros_like(x), np.zeros_like(x)
                lib.gb200h_tracking_loop_filter_run(T, bw, order, last, 0.5, x.ctypes.data, a.ctypes.data, len(x))
                ref.ref_tracking_loop_filter_run(T, bw, order, last, 0.5, x.ctypes.data, b.ctypes.data, len(x))
                assert np.array_equal(a.view(np.uint32), b.view(np.uint32)), (order, last, T, bw)
    for order in (2, 3):
        for T in (0.001, 0.02):
            a, b = np.zeros_like(x), np.zeros_like(x)
            lib.gb200h_fll_pll_filter_run(10.0, 35.0, order, 1234.5, x.ctypes.data, x2.ctypes.data, T, a.ctypes.data, len(x))
            ref.ref_fll_pll_filter_run(10.0, 35.0, order, 1234.5, x.ctypes.data, x2.ctypes.data, T, b.ctypes.data, len(x))
            assert np.array_equal(a.view(np.uint32), b.view(np.uint32)), (order, T)
    cn0 = (40 + rng.standard_normal(600)).astype(np.float32)
    cn0[:50] = 20.0  # forces one restart of the warm-up (mean below min_value + offset)
    for alpha, n_init, mn, off in ((0.002, 200, 25.0, 12.0), (0.1, 25, -1.0, 0.0), (1.5, 0, 25.0, 12.0)):
        a, b = np.zeros_like(cn0), np.zeros_like(cn0)
        lib.gb200h_exponential_smoother_run(alpha, n_init, mn, off, cn0.ctypes.data, a.ctypes.data, len(cn0))
        ref.ref_exponential_smoother_run(alpha, n_init, mn, off, cn0.ctypes.data, b.ctypes.data, len(cn0))
        assert np.array_equal(a.view(np.uint32), b.view(np.uint32)), (alpha, n_init)


def test_gps_l1_ca_codes_bit_exact_vs_golden_reference_tables():
    lib = H.lib()
    for prn in (1, 7, 10, 22, 32, 37):
        code = np.zeros(1023, dtype=np.float32)
        assert lib.gb200h_gps_l1_ca_code_gen_float(code.ctypes.data, prn, 0) == 0
        assert np.array_equal(code, O.gps_l1_ca_code(prn)), prn  # the oracle's tables are pinned to the reference's (test_oracle_pin)
        if f"ca_{prn}" in GOLD.files:
            assert np.array_equal(code, GOLD[f"ca_{prn}"])
    shifted = np.zeros(1023, dtype=np.float32)
    lib.gb200h_gps_l1_ca_code_gen_float(shifted.ctypes.data, 5, 100)
    assert np.array_equal(shifted, O.gps_l1_ca_code(5, chip_shift=100))
    for fs, n in ((4000000, 4000), (25000000, 25000), (2048000, 2048)):
        s = np.zeros(n, dtype=np.complex64)
        assert lib.gb200h_gps_l1_ca_code_gen_complex_sampled(s.ctypes.data, n, 10, fs, 0) == 0
        assert np.array_equal(s, O.gps_l1_ca_code_sampled(10, fs, n)), fs
    assert lib.gb200h_gps_l1_ca_code_gen_float(code.ctypes.data, 99, 0) == -1 and b"PRN" in lib.gb200h_last_error()


@needs_ref
def test_gps_l1_ca_codes_bit_exact_vs_reference_generator_live():
    lib = H.lib()
    for prn in range(1, 38):
        a = np.zeros(1023, dtype=np.float32)
        lib.gb200h_gps_l1_ca_code_gen_float(a.ctypes.data, prn, 0)
        assert np.array_equal(a, O.gps_l1_ca_code(prn, impl="ref")), prn
    s = np.zeros(25000, dtype=np.complex64)
    lib.gb200h_gps_l1_ca_code_gen_complex_sampled(s.ctypes.data, 25000, 3, 25000000, 0)
    assert np.array_equal(s, O.gps_l1_ca_code_sampled(3, 25000000, 25000, impl="ref"))


def test_gpu_entry_points_fail_loudly_without_a_device():
    from gnss_sim_receiver_b200 import _capi
    if _capi.lib().gb200_device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(H.HostError, match="no CUDA device|CPU fallback"):
        H.TrackingBank(CONF_C1, "Tracking_1C", 2)
    with pytest.raises(H.HostError, match="no CUDA device|CPU fallback"):
        H.AcquisitionBlock(CONF_C1, "Acquisition_1C", 1)
    # the SURVEY 8f-4 strategy blocks likewise: no host path stands in for the engines
    base = "GNSS-SDR.internal_fs_sps=4000000\n"
    with pytest.raises(H.HostError, match="no CUDA device|CPU fallback"):
        H.FineDopplerAcquisitionBlock(base + "Acquisition_1C.implementation=GPS_L1_CA_PCPS_Acquisition_Fine_Doppler_B200\n", "Acquisition_1C", 1)
    with pytest.raises(H.HostError, match="no CUDA device|CPU fallback"):
        H.QuickSyncAcquisitionBlock(base + "Acquisition_1C.implementation=GPS_L1_CA_PCPS_QuickSync_Acquisition_B200\n", "Acquisition_1C", 1)
    with pytest.raises(H.HostError, match="no CUDA device|CPU fallback"):
        H.E5aCafAcquisitionBlock("GNSS-SDR.internal_fs_sps=12000000\nAcquisition_5X.implementation=Galileo_E5a_Noncoherent_IQ_Acquisition_CAF_B200\n", "Acquisition_5X", 1)


def test_mat_writer_is_readable_by_scipy(tmp_path):
    """The acquisition dump writer (MAT level 5): classes, shapes and column-major data survive a round trip through
    scipy.io.loadmat, the reader the reference's python utilities use."""
    import ctypes as C
    from scipy.io import loadmat
    from gnss_sim_receiver_b200 import _hostapi as H
    rng = np.random.default_rng(1)
    grid = rng.random((5, 37)).astype(np.float32)  # 5 Doppler bins of 37 code phases, one bin after the other
    f = str(tmp_path / "acq_G_1C_ch_0_1_sat_7.mat")
    assert H.lib().gb200h_mat_write_sample(f.encode(), grid.ctypes.data, 37, 5, -5000, 7, 2**40 + 3, np.float32(2.5)) == 0
    m = loadmat(f)
    assert m["acq_grid"].dtype == np.float32 and m["acq_grid"].shape == (37, 5)
    assert np.array_equal(m["acq_grid"], grid.T)
    assert m["doppler_max"].dtype == np.int32 and int(m["doppler_max"][0, 0]) == -5000
    assert m["PRN"].dtype == np.uint32 and int(m["PRN"][0, 0]) == 7
    assert m["sample_counter"].dtype == np.uint64 and int(m["sample_counter"][0, 0]) == 2**40 + 3
    assert m["threshold"].dtype == np.float32 and float(m["threshold"][0, 0]) == 2.5
