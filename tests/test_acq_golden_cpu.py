"""Pins the acquisition oracle (oracle/oracle_acq.c) on the frozen scenarios of tests/_signals.py: it must keep reproducing
tests/golden/acq_vectors.json — produced by tests/golden/make_acq_golden.py — with BOTH FFT precisions agreeing on every index.

What this pin is and is not: the reference's pcps_acquisition needs GNU Radio / FFTW / VOLK / Boost, none of which exist here, so the
vectors were frozen from the restatement; tests/test_acq_refblock_cpu.py then has the reference's OWN pcps_acquisition.cc (compiled
unmodified into oracle/_ref, with a stand-in DFT behind gr::fft) reproduce them.  The vectors freeze the restatement (which follows
pcps_acquisition.cc:175-208,496-697 and uses the reference's own sincos / index_max kernels where they compile) against regressions,
require the fp32-FFT and float64-FFT evaluations to name the same (code phase, Doppler bin) — so no frozen index hangs on FFT rounding —
and give the GPU tests an oracle-free target (tests/test_acq_golden_gpu.py).
"""
import json
import os

import pytest

import _oracle as O
import _signals as S

VEC = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "acq_vectors.json")))


def test_every_frozen_case_is_present_and_unambiguous():
    want = {f"{name}/seed{seed}" for name, spec in S.ACQ_CASES.items() for seed in spec[-1]}
    assert set(VEC) == want
    for key, v in VEC.items():
        name = key.split("/")[0]
        present = S.ACQ_CASES[name][7]
        for prn, r in v.items():
            if not isinstance(r, dict):
                continue
            assert not r["near_tie"], (key, prn)  # both FFT precisions name the same cell on every frozen vector
            if int(prn) in present and v["threshold_pfa_1e-3"] is not None:
                assert r["test_statistics"] > v["threshold_pfa_1e-3"], (key, prn)  # satellites in the air are detections


@pytest.mark.parametrize("key", sorted(k for k in VEC if not k.startswith(("c3_e1b", "c5_"))))  # (the two 100000-point cases are re-run by make_acq_golden.py only: 3 s each)
def test_oracle_reproduces_the_frozen_vectors(key):
    name, seed = key.split("/seed")
    case = S.acq_case(name, int(seed))
    for f64 in (False, True):
        conf = O.acq_conf(case["fs"], case["consumed"], case["fft_size"], case["dmax"], case["dstep"], case["spc"], case["spchip"], bit_transition=case["bit_transition"],
                          use_cfar=case["use_cfar"], max_dwells=case["max_dwells"], fft_f64=f64)
        for p in case["prns"][:4]:
            fc = O.acq_set_local_code(conf, case["locals_"][p])
            grid = None
            for d, s in enumerate(case["sigs"]):
                grid = O.acq_dwell(conf, s, fc, grid, d)
            r = O.acq_statistic(conf, grid, len(case["sigs"]))
            want = VEC[key][str(p)]
            assert (r.indext, r.doppler_index, r.doppler_hz) == (want["indext"], want["doppler_index"], want["doppler_hz"]), (key, p, f64)
            assert r.acq_delay_samples == want["acq_delay_samples"]
            ref_stat = want["f64_test_statistics"] if f64 else want["test_statistics"]
            assert abs(r.test_statistics - ref_stat) <= 1e-6 * ref_stat
