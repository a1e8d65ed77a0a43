#!/usr/bin/env python
"""Generates tests/golden/ref_vectors.npz from the REFERENCE'S OWN CODE compiled into oracle/_ref/libgnssref.so
(oracle/Makefile; only possible in the dev container where /root/reference exists).  The vectors travel with the repo so
that the oracle restatement and the product host code stay pinned on machines without the reference.

    python tests/golden/make_golden.py

Everything here is seeded; re-running reproduces the committed file bit for bit.
"""
import ctypes as C
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import _oracle as O  # noqa: E402

f32p = O.f32p
F = C.c_float


def p(a):
    return a.ctypes.data_as(f32p)


def main():
    r = O.ref()
    assert r is not None, "oracle/_ref/libgnssref.so missing: run `make -C oracle` where /root/reference exists"
    out = {}
    # ---- multicorrelator cases: (n, L, ntaps, code_step, rem_carr, phase_step, rem_code, seed), reference generic (0) and AVX (1)
    cases = [(4000, 1023, 3, 0.25575, 0.25, 0.1, 0.4, 1), (8111, 1023, 3, 0.3, 0.25, 0.1, 0.4, 2), (2048, 2046, 3, 0.3, 0.0, 0.1, 0.4, 3),
             (16000, 8184, 5, 0.5115, -1.3, -0.02137, -0.7, 4), (16000, 8184, 11, 0.5115, 2.9, 0.0, 0.0, 5), (20000, 10230, 11, 0.5115, 1.0, 0.003, 0.2, 6),
             (25000, 1023, 11, 0.04092, 0.5, 0.0012, -0.3, 7), (777, 1023, 1, 0.26, 0.1, 0.2, 0.3, 8)]
    out["mc_cases"] = np.array(cases, dtype=np.float64)
    for i, (n, L, nt, cstep, remc, pstep, remcode, seed) in enumerate(cases):
        n, L, nt, seed = int(n), int(L), int(nt), int(seed)
        rng = np.random.default_rng(seed)
        sig = (rng.uniform(-0.5, 0.5, n) + 1j * rng.uniform(-0.5, 0.5, n)).astype(np.complex64)
        code = rng.choice([-1.0, 1.0], L).astype(np.float32)
        shifts = ((np.arange(nt) - (nt - 1) / 2.0) * 0.5).astype(np.float32)
        for arch in (0, 1):
            out[f"mc_{i}_arch{arch}"] = O.mc_real_codes(sig, code, shifts, remc, pstep, remcode, cstep, impl="ref", arch=arch)
        # resampler index table of tap 0 (generic law) as int16 delta-coded to keep the file small: store only a checksum and head
        rep = np.zeros((nt, n), dtype=np.float32)
        ptrs = (f32p * nt)(*[p(rep[t]) for t in range(nt)])
        sh = shifts.copy()
        r.ref_resampler_32f(0, ptrs, p(code), F(remcode), F(cstep), p(sh), L, nt, n)
        out[f"mc_{i}_rep_head"] = rep[:, :64].copy()
        out[f"mc_{i}_rep_sum"] = rep.astype(np.float64).sum(axis=1)
    # complex-code sibling
    rng = np.random.default_rng(21)
    n, L, nt = 4096, 1023, 3
    sig = (rng.uniform(0, 1, n) + 1j * rng.uniform(0, 1, n)).astype(np.complex64)
    codec = (rng.choice([-1.0, 1.0], L) + 1j * rng.choice([-1.0, 1.0], L)).astype(np.complex64)
    shifts = np.array([-0.5, 0.0, 0.5], np.float32)
    out["mcc_ref"] = O.mc_complex_codes(sig, codec, shifts, 0.0, 0.1, 0.4, 0.3, impl="ref", arch=0)
    # high-dynamics pair
    rng = np.random.default_rng(22)
    n = 4000
    sig = (rng.uniform(-0.5, 0.5, n) + 1j * rng.uniform(-0.5, 0.5, n)).astype(np.complex64)
    code = rng.choice([-1.0, 1.0], 1023).astype(np.float32)
    out["mchd_ref"] = O.mc_real_codes(sig, code, shifts, 0.3, 0.05, 0.4, 0.25575, high_dyn=True, phase_rate=1e-7, code_rate=1e-9, impl="ref", arch=0)
    # ---- PRN tables (a17), stored as int8
    gps = np.zeros((32, 1023), np.float32)
    for prn in range(1, 33):
        r.ref_gps_l1_ca_code_gen_float(p(gps[prn - 1]), prn, 0)
    out["gps_l1_ca"] = gps.astype(np.int8)
    for prn, fs in ((1, 4000000), (10, 4000000), (7, 25000000)):
        n = fs // 1000
        buf = np.zeros(n, np.complex64)
        r.ref_gps_l1_ca_code_gen_complex_sampled(p(buf), n, prn, fs, 0)
        assert np.all(buf.real == 0)
        out[f"gps_l1_ca_sampled_{prn}_{fs}"] = buf.imag.astype(np.int8)
    e1 = {}
    for sig_id in (b"1B", b"1C"):
        for prn in (1, 11, 36):
            buf = np.zeros(8184, np.float32)
            r.ref_galileo_e1_code_gen_sinboc11_float(p(buf), 8184, sig_id, prn)
            e1[f"gal_e1_{sig_id.decode()}_{prn}"] = buf.astype(np.int8)
    out.update(e1)
    buf = np.zeros(16000, np.complex64)
    r.ref_galileo_e1_code_gen_complex_sampled(p(buf), 16000, b"1B", 0, 11, 4000000, 0, 0)
    out["gal_e1_1B_11_sampled_4M_re"] = buf.real.astype(np.int8)
    out["gal_e1_1B_11_sampled_4M_im"] = buf.imag.astype(np.int8)
    for prn in (1, 17, 32):
        bi = np.zeros(10230, np.float32)
        bq = np.zeros(10230, np.float32)
        r.ref_gps_l5i_code_gen_float(p(bi), prn)
        r.ref_gps_l5q_code_gen_float(p(bq), prn)
        out[f"gps_l5i_{prn}"] = bi.astype(np.int8)
        out[f"gps_l5q_{prn}"] = bq.astype(np.int8)
        e5 = np.zeros(10230, np.complex64)
        r.ref_galileo_e5_a_code_gen_complex_primary(p(e5), prn, b"5X")
        out[f"gal_e5a_5X_{prn}_re"] = e5.real.astype(np.int8)
        out[f"gal_e5a_5X_{prn}_im"] = e5.imag.astype(np.int8)
    # ---- sincos / index_max kernels used by acquisition
    ph = np.zeros(1, np.float32)
    sc = np.zeros(4000, np.complex64)
    r.ref_sincos_32fc(0, p(sc), F(-2 * np.pi * 1750.0 / 4e6), p(ph), 4000)
    out["sincos_generic_1750Hz_4M"] = sc
    rng = np.random.default_rng(23)
    v = rng.uniform(0, 1, 5000).astype(np.float32)
    v[[100, 2000, 4999]] = 2.0  # ties: the first index wins
    idx = C.c_uint32()
    r.ref_index_max_32u(0, C.byref(idx), p(v), 5000)
    out["index_max_in"] = v
    out["index_max_out"] = np.array([idx.value], np.uint32)
    # ---- loop closure (a8): discriminators, loop filters, lock detectors, smoother
    rng = np.random.default_rng(24)
    z = (rng.standard_normal((64, 4)) + 1j * rng.standard_normal((64, 4))).astype(np.complex64)  # VE/E, E/P, L, VL style operands
    out["disc_in"] = z
    d = np.zeros((64, 8), np.float64)
    for i in range(64):
        a, b, c_, e = z[i]
        d[i, 0] = r.ref_pll_cloop_two_quadrant_atan(F(b.real), F(b.imag))
        d[i, 1] = r.ref_pll_four_quadrant_atan(F(b.real), F(b.imag))
        d[i, 2] = r.ref_fll_four_quadrant_atan(F(a.real), F(a.imag), F(b.real), F(b.imag), C.c_double(0.0), C.c_double(0.001))
        d[i, 3] = r.ref_fll_diff_atan(F(a.real), F(a.imag), F(b.real), F(b.imag), C.c_double(0.0), C.c_double(0.001))
        d[i, 4] = r.ref_dll_nc_e_minus_l_normalized(F(a.real), F(a.imag), F(c_.real), F(c_.imag), F(0.5), F(1.0), F(1.0))
        d[i, 5] = r.ref_dll_nc_e_minus_l_normalized(F(a.real), F(a.imag), F(c_.real), F(c_.imag), F(0.15), F(3.0), F(1.0))
        d[i, 6] = r.ref_dll_nc_vemlp_normalized(F(a.real), F(a.imag), F(b.real), F(b.imag), F(c_.real), F(c_.imag), F(e.real), F(e.imag))
    out["disc_out"] = d
    out["sinboc_slope_yint"] = np.array([[x, r.ref_sinboc_slope_abs(C.c_double(x)), r.ref_sinboc_yintercept_abs(C.c_double(x))] for x in (0.15, 0.25, 5.0 / 12.0, 0.6)])
    x = rng.standard_normal(200).astype(np.float32)
    out["filt_in"] = x
    k = 0
    for order in (1, 2, 3):
        for last in (0, 1):
            y = np.zeros(200, np.float32)
            r.ref_tracking_loop_filter_run(F(0.001), F(2.0), order, last, F(0.5), p(x), p(y), 200)
            out[f"loop_filter_o{order}_l{last}"] = y
            k += 1
    x2 = rng.standard_normal(200).astype(np.float32)
    out["filt_in2"] = x2
    for order in (2, 3):
        y = np.zeros(200, np.float32)
        r.ref_fll_pll_filter_run(F(10.0), F(35.0), order, F(1234.5), p(x), p(x2), F(0.001), p(y), 200)
        out[f"fll_pll_filter_o{order}"] = y
    pr = (rng.standard_normal(20) * 0.3 + 5.0 + 1j * rng.standard_normal(20) * 0.3).astype(np.complex64)
    out["lock_prompt"] = pr
    out["lock_out"] = np.array([r.ref_cn0_m2m4_estimator(p(pr), 20, F(0.001)), r.ref_cn0_svn_estimator(p(pr), 20, F(0.001)),
                                r.ref_carrier_lock_detector(p(pr), 1), r.ref_carrier_lock_detector(p(pr), 20)], np.float32)
    sm_in = (30 + rng.standard_normal(300)).astype(np.float32)
    sm_out = np.zeros(300, np.float32)
    r.ref_exponential_smoother_run(F(0.002), 200, F(0.0), F(0.0), p(sm_in), p(sm_out), 300)
    out["smoother_in"], out["smoother_out"] = sm_in, sm_out
    path = os.path.join(HERE, "ref_vectors.npz")
    np.savez_compressed(path, **out)
    print(f"wrote {path}: {len(out)} arrays, {os.path.getsize(path) / 1024:.0f} KiB")


if __name__ == "__main__":
    main()
