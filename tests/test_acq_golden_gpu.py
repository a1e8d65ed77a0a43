"""The acquisition engine against the FROZEN vectors of tests/golden/acq_vectors.json (every BASELINE config in small, seeds frozen per
config): code-phase index, Doppler bin, Doppler, Acq_delay_samples identical for every PRN searched — present or noise-only — and the
statistic within 1e-3.  No oracle runs here: the targets were frozen by tests/golden/make_acq_golden.py with the fp32- and float64-FFT
oracles in agreement on every index (tests/test_acq_golden_cpu.py keeps the oracle on them)."""
import json
import os

import pytest

import _signals as S

pytestmark = pytest.mark.gpu
VEC = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "acq_vectors.json")))


@pytest.mark.parametrize("key", sorted(VEC))
def test_engine_reproduces_the_frozen_vectors(key):
    from gnss_sim_receiver_b200.acquisition import AcquisitionEngine
    name, seed = key.split("/seed")
    c = S.acq_case(name, int(seed))
    eng = AcquisitionEngine(fs_in=c["fs"], fft_size=c["fft_size"], consumed_samples=c["consumed"], samples_per_code=c["spc"], samples_per_chip=c["spchip"],
                            bit_transition=c["bit_transition"], use_cfar=c["use_cfar"], max_dwells=c["max_dwells"], max_prns=len(c["prns"]), iq_dtype=c["iq_dtype"])
    eng.set_doppler_grid(c["dmax"], c["dstep"])
    for i, p in enumerate(c["prns"]):
        eng.set_local_code(i, c["locals_"][p])
    res = None
    for d, items in enumerate(c["items"]):
        res = eng.run(items, dwell_index=d)
    thr = eng.threshold(0.001) if c["use_cfar"] else None
    eng.close()
    if thr is not None:
        assert abs(thr - VEC[key]["threshold_pfa_1e-3"]) <= 1e-6 * thr
    for i, p in enumerate(c["prns"]):
        want, r = VEC[key][str(p)], res[i]
        assert (r.indext, r.doppler_index, r.doppler_hz) == (want["indext"], want["doppler_index"], want["doppler_hz"]), (key, p)
        assert r.acq_delay_samples == want["acq_delay_samples"]
        assert abs(r.test_statistics - want["test_statistics"]) <= 1e-3 * want["test_statistics"], (key, p)
        assert abs(r.peak - want["peak"]) <= 1e-3 * want["peak"]
