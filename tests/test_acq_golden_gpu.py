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


@pytest.mark.parametrize("name,seed", [("c1_gps_4msps", 3), ("c2_e1c_bit_transition_cshort", 1), ("c3_gps_25msps", 2), ("c4_l5i_4_dwells", 2)])
def test_engine_indices_equal_the_reference_blocks(name, seed):
    """north_star's gate, literally: acquisition code-phase and Doppler-bin results against the reference's pcps_acquisition — here the
    reference's OWN pcps_acquisition.cc, compiled unmodified into oracle/_ref (oracle/ref_acq.cc; the DFT library behind gr::fft is the one
    stand-in) and run live on the same block as the engine — for EVERY PRN searched, satellites in the air and noise-only alike."""
    import numpy as np

    import _oracle as O
    import test_acq_refblock_cpu as R
    if not O.ref_available():
        pytest.skip("oracle/_ref not built")
    from gnss_sim_receiver_b200.acquisition import AcquisitionEngine
    c = S.acq_case(name, seed)
    eng = AcquisitionEngine(fs_in=c["fs"], fft_size=c["fft_size"], consumed_samples=c["consumed"], samples_per_code=c["spc"], samples_per_chip=c["spchip"],
                            bit_transition=c["bit_transition"], use_cfar=c["use_cfar"], max_dwells=c["max_dwells"], max_prns=len(c["prns"]), iq_dtype=c["iq_dtype"])
    eng.set_doppler_grid(c["dmax"], c["dstep"])
    for i, p in enumerate(c["prns"]):
        eng.set_local_code(i, c["locals_"][p])
    res = None
    for d, items in enumerate(c["items"]):
        res = eng.run(items, dwell_index=d)
    eng.close()
    stream = np.concatenate(c["items"])
    k = 1 if S.ACQ_CASES[name][8] == "gr_complex" else 2
    for i, p in enumerate(c["prns"]):
        b = R._block(name, c, p, False, 0.001)
        b.set_threshold(3e38)  # take every dwell
        b.arm()
        b.work(stream, k)
        st = b.status()
        b.close()
        r = res[i]
        assert (r.acq_delay_samples, float(r.doppler_hz)) == (st.acq_delay_samples, st.acq_doppler_hz), (name, p, p in c["present"])
        assert abs(r.test_statistics - st.test_statistics) <= 1e-3 * st.test_statistics, (name, p)
