#!/usr/bin/env python
"""Freezes the acquisition oracle's answers on the seeded scenarios of tests/_signals.py (ACQ_CASES) into tests/golden/acq_vectors.json.

The reference's pcps_acquisition cannot be built here (GNU Radio, FFTW, VOLK, Boost absent: "parity unpinned"), so the
strongest pin this box allows is: the oracle restatement (oracle/oracle_acq.c, following pcps_acquisition.cc:175-208,496-697 line by
line, with the reference's own sincos and index_max kernels bit-pinned in tests/test_oracle_pin.py) evaluated with TWO independent FFT
precisions — complex64 and complex128 — which must agree on every (code phase, Doppler bin) before a vector is frozen.
A (PRN, seed) on which they disagree (an exact-arithmetic near-tie on a noise-only grid) is recorded as `near_tie` and not pinned.

    python tests/golden/make_acq_golden.py        # rewrites acq_vectors.json; seeded, reproducible bit for bit
"""
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import _oracle as O  # noqa: E402
import _signals as S  # noqa: E402


def oracle_answers(case):
    """{prn: dict(...)} after the last dwell, from the fp32-FFT oracle, plus whether the float64-FFT oracle agrees on the indices"""
    out = {}
    confs = {f64: O.acq_conf(case["fs"], case["consumed"], case["fft_size"], case["dmax"], case["dstep"], case["spc"], case["spchip"], bit_transition=case["bit_transition"],
                             use_cfar=case["use_cfar"], max_dwells=case["max_dwells"], fft_f64=f64) for f64 in (False, True)}
    for p in case["prns"]:
        res = {}
        for f64, conf in confs.items():
            fc = O.acq_set_local_code(conf, case["locals_"][p])
            grid = None
            for d, s in enumerate(case["sigs"]):
                grid = O.acq_dwell(conf, s, fc, grid, d)
            res[f64] = O.acq_statistic(conf, grid, len(case["sigs"]))
        a, b = res[False], res[True]
        out[str(p)] = dict(indext=int(a.indext), doppler_index=int(a.doppler_index), doppler_hz=int(a.doppler_hz), acq_delay_samples=float(a.acq_delay_samples),
                           test_statistics=float(a.test_statistics), peak=float(a.peak), near_tie=bool((a.indext, a.doppler_index) != (b.indext, b.doppler_index)),
                           f64_test_statistics=float(b.test_statistics))
    out["threshold_pfa_1e-3"] = O.acq_threshold(confs[False], 0.001) if case["use_cfar"] else None
    return out


def main():
    vec = {}
    for name, spec in S.ACQ_CASES.items():
        for seed in spec[-1]:
            vec[f"{name}/seed{seed}"] = oracle_answers(S.acq_case(name, seed))
            print(name, seed, {k: (v["indext"], v["doppler_index"], round(v["test_statistics"], 1), v["near_tie"]) for k, v in vec[f"{name}/seed{seed}"].items() if isinstance(v, dict)})
    with open(os.path.join(HERE, "acq_vectors.json"), "w") as f:
        json.dump(vec, f, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
