"""Pins the oracle (oracle/liboracle.so, our C restatement) before anything trusts it.

1. against tests/golden/ref_vectors.npz — outputs of the reference's OWN code (compiled from /root/reference into
   oracle/_ref by oracle/Makefile, vectors written by tests/golden/make_golden.py); runs everywhere;
2. against oracle/_ref/libgnssref.so live, on more random cases, wherever that library exists (dev container; it also
   travels to the GPU box as a prebuilt file);
3. the FFT stand-in against numpy's pocketfft in float64, and the acquisition restatement against the known answer of
   the reference's own synthetic acquisition test (GpsL1CaPcpsAcquisitionGSoC2013Test: PRN 10, 750 Hz, 600 chips —
   src/tests/unit-tests/signal-processing-blocks/acquisition/gps_l1_ca_pcps_acquisition_gsoc2013_test.cc:204-230,391-399).
"""
import os
import sys

import numpy as np
import pytest

import _oracle as O

GOLD = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_vectors.npz"))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))


def _case_inputs(i):
    n, L, nt, cstep, remc, pstep, remcode, seed = GOLD["mc_cases"][i]
    n, L, nt, seed = int(n), int(L), int(nt), int(seed)
    rng = np.random.default_rng(seed)
    sig = (rng.uniform(-0.5, 0.5, n) + 1j * rng.uniform(-0.5, 0.5, n)).astype(np.complex64)
    code = rng.choice([-1.0, 1.0], L).astype(np.float32)
    shifts = ((np.arange(nt) - (nt - 1) / 2.0) * 0.5).astype(np.float32)
    return sig, code, shifts, np.float32(remc), np.float32(pstep), np.float32(remcode), np.float32(cstep)


@pytest.mark.parametrize("i", range(8))
def test_multicorrelator_bit_exact_vs_reference_generic(i):
    sig, code, shifts, remc, pstep, remcode, cstep = _case_inputs(i)
    got = O.mc_real_codes(sig, code, shifts, remc, pstep, remcode, cstep)
    ref = GOLD[f"mc_{i}_arch0"]
    assert np.array_equal(got.view(np.uint32), ref.view(np.uint32)), (got, ref)
    # the AVX protokernels (what an x86 host really runs) stay within the reference's own QA tolerance of the generic ones
    avx = GOLD[f"mc_{i}_arch1"]
    assert np.max(np.abs(avx - ref)) / np.max(np.abs(ref)) < 1e-3  # kernel_tests.h:43-44,103-105
    # resampler law
    idx = O.resampler_index(remcode, cstep, shifts, len(code), len(sig))
    rep = code[idx]
    assert np.array_equal(rep[:, :64], GOLD[f"mc_{i}_rep_head"])
    assert np.array_equal(rep.astype(np.float64).sum(axis=1), GOLD[f"mc_{i}_rep_sum"])
    # float64 truth agrees with the fp32 reference to 1e-4 of the largest output
    f64 = O.mc_real_codes_f64(sig, code, shifts, remc, pstep, remcode, cstep, mimic_inc=True)
    assert np.max(np.abs(f64 - ref)) / np.max(np.abs(ref)) < 1e-4


def test_complex_code_and_high_dynamics_variants():
    rng = np.random.default_rng(21)
    n, L = 4096, 1023
    sig = (rng.uniform(0, 1, n) + 1j * rng.uniform(0, 1, n)).astype(np.complex64)
    codec = (rng.choice([-1.0, 1.0], L) + 1j * rng.choice([-1.0, 1.0], L)).astype(np.complex64)
    shifts = np.array([-0.5, 0.0, 0.5], np.float32)
    got = O.mc_complex_codes(sig, codec, shifts, 0.0, 0.1, 0.4, 0.3)
    assert np.array_equal(got.view(np.uint32), GOLD["mcc_ref"].view(np.uint32))
    rng = np.random.default_rng(22)
    n = 4000
    sig = (rng.uniform(-0.5, 0.5, n) + 1j * rng.uniform(-0.5, 0.5, n)).astype(np.complex64)
    code = rng.choice([-1.0, 1.0], 1023).astype(np.float32)
    got = O.mc_real_codes(sig, code, shifts, 0.3, 0.05, 0.4, 0.25575, high_dyn=True, phase_rate=1e-7, code_rate=1e-9)
    ref = GOLD["mchd_ref"]
    # the reference calls cpowf per sample; the restatement uses its documented closed form: equal to fp32 noise level
    assert np.max(np.abs(got - ref)) / np.max(np.abs(ref)) < 1e-4


def test_prn_tables_bit_exact():
    for prn in range(1, 33):
        assert np.array_equal(O.gps_l1_ca_code(prn).astype(np.int8), GOLD["gps_l1_ca"][prn - 1]), prn
    for prn, fs in ((1, 4000000), (10, 4000000), (7, 25000000)):
        s = O.gps_l1_ca_code_sampled(prn, fs, fs // 1000)
        assert np.all(s.real == 0)
        assert np.array_equal(s.imag.astype(np.int8), GOLD[f"gps_l1_ca_sampled_{prn}_{fs}"])


def test_wipeoff_and_index_max_kernels():
    w = O.acq_wipeoff(4000, 1750.0, 4e6)
    ref = GOLD["sincos_generic_1750Hz_4M"]
    assert np.array_equal(w.view(np.uint32), ref.view(np.uint32))
    v = GOLD["index_max_in"]
    assert int(GOLD["index_max_out"][0]) == 100 == int(np.argmax(v))  # first index wins ties


@pytest.mark.skipif(not O.ref_available(), reason="oracle/_ref/libgnssref.so not built (needs /root/reference)")
@pytest.mark.parametrize("seed", range(6))
def test_live_reference_random_cases(seed):
    rng = np.random.default_rng(100 + seed)
    n = int(rng.integers(500, 30000))
    L = int(rng.choice([1023, 2046, 8184, 10230]))
    nt = int(rng.choice([1, 3, 5, 11]))
    sig = (rng.standard_normal(n) + 1j * rng.standard_normal(n)).astype(np.complex64)
    code = rng.choice([-1.0, 1.0], L).astype(np.float32)
    shifts = np.sort(rng.uniform(-1.5, 1.5, nt)).astype(np.float32)
    args = (np.float32(rng.uniform(-3, 3)), np.float32(rng.uniform(-0.2, 0.2)), np.float32(rng.uniform(-1, 1)), np.float32(rng.uniform(0.03, 0.9)))
    got = O.mc_real_codes(sig, code, shifts, *args)
    ref = O.mc_real_codes(sig, code, shifts, *args, impl="ref", arch=0)
    assert np.array_equal(got.view(np.uint32), ref.view(np.uint32))
    for prn in rng.integers(1, 33, 3):
        assert np.array_equal(O.gps_l1_ca_code(int(prn)), O.gps_l1_ca_code(int(prn), impl="ref"))


@pytest.mark.parametrize("n", [8, 60, 1400, 4000, 16000, 25000, 32000])
def test_fft_stand_in_vs_pocketfft(n):
    rng = np.random.default_rng(n)
    x = (rng.standard_normal(n) + 1j * rng.standard_normal(n))
    ref = np.fft.fft(x)
    assert np.linalg.norm(O.fft_c64(x) - ref) / np.linalg.norm(ref) < 1e-13
    assert np.linalg.norm(O.fft_c32(x) - ref) / np.linalg.norm(ref) < 3e-6
    assert np.linalg.norm(O.fft_c64(ref, inverse=True) - n * x) / (n * np.linalg.norm(x)) < 1e-13


def test_acquisition_known_answer_gsoc2013_scenario():
    from gen_if import gen_if
    fs, n = 4e6, 4000
    codes = {p: O.gps_l1_ca_code(p) for p in (10, 15, 21, 22)}
    sats = [dict(prn=10, doppler=750.0, delay_chips=600.0, cn0=47.0), dict(prn=15, doppler=4000.0, delay_chips=300.0, cn0=44.0),
            dict(prn=21, doppler=-1250.0, delay_chips=720.0, cn0=44.0), dict(prn=22, doppler=9000.0, delay_chips=10.0, cn0=44.0)]
    sig = gen_if(codes, fs, n, sats, seed=3)
    for f64 in (False, True):
        conf = O.acq_conf(fs, n, n, 10000, 250, float(n), 4, fft_f64=f64)
        assert conf.num_doppler_bins == 80 and O.acq_neff(conf) == 4000
        fc = O.acq_set_local_code(conf, O.gps_l1_ca_code_sampled(10, fs, n))
        r = O.acq_statistic(conf, O.acq_dwell(conf, sig, fc))
        delay_chips = r.acq_delay_samples * 1.023e6 / fs
        assert abs(delay_chips - 600.0) < 0.5 and abs(r.acq_doppler_hz - 750.0) < 666.7
        assert r.test_statistics > O.acq_threshold(conf, 0.001)
    # threshold law: 2 * P^-1(2 * dwells, (1 - pfa)^(1/cells)) checked against scipy's regularised gamma inverse
    from scipy.special import gammaincinv
    conf = O.acq_conf(fs, n, n, 10000, 250, float(n), 4, max_dwells=3)
    expect = 2.0 * gammaincinv(2.0 * 3, (1.0 - float(np.float32(0.001))) ** (1.0 / 320000.0))
    assert abs(O.acq_threshold(conf, 0.001) - expect) < 1e-4 * expect
