"""The acquisition oracle against the REFERENCE'S OWN acquisition block.

oracle/_ref/libgnssref.so carries src/algorithms/acquisition/gnuradio_blocks/pcps_acquisition.cc and src/algorithms/acquisition/libs/acq_conf.cc
compiled UNMODIFIED (oracle/ref_acq.cc, oracle/shim/acq/README.txt): the constructor's sizing, set_local_code, update_grid_doppler_wipeoffs,
general_work's buffering state machine, acquisition_core, max_to_input_power_statistic / first_vs_second_peak_statistic, the dwell logic,
calculate_threshold and the positive / negative acquisition events are the reference's code, running on the reference's own volk_gnsssdr
protokernels (sincos, index_max, 16ic convert).  Stand-ins: the DFT behind gr::fft (FFTW is not installed: oracle/oracle_fft.c, float32 or
float64 inside), five element-wise VOLK kernels, Boost's gamma_p_inv, and scheduler plumbing.

What is pinned here: the frozen vectors of tests/golden/acq_vectors.json (which the oracle restatement and, on the GPU, the engine reproduce)
are reproduced by the reference block itself — code phase, Doppler, peak, statistic, threshold, events — with both DFT precisions.
"""
import json
import os

import numpy as np
import pytest

import _oracle as O
import _signals as S

pytestmark = pytest.mark.skipif(not O.ref_available(), reason="oracle/_ref not built")
VEC = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "acq_vectors.json")))

# signal -> (role, system, Gnss_Synchro signal, ms_per_code, chip rate, optimum acquisition rate) as the adapters pass them:
# gps_l1_ca_pcps_acquisition.cc:49-50, galileo_e1_pcps_ambiguous_acquisition.cc:50-51, gps_l5i_pcps_acquisition.cc:50-51
SIG = {"1C": ("Acquisition_1C", "G", "1C", 1, 1.023e6, 2e6), "1B": ("Acquisition_1B", "E", "1B", 4, 1.023e6, 4e6),
       "1Cp": ("Acquisition_1B", "E", "1B", 4, 1.023e6, 4e6), "L5": ("Acquisition_L5", "G", "L5", 1, 10.23e6, 10e6)}


def _block(name, case, prn, fft_f64, pfa):
    fs, fft_size, consumed, dmax, dstep, signal, prns, present, item, bt, cfar, nd, _ = S.ACQ_CASES[name]
    role, system, sig2, ms_per_code, chip_rate, opt = SIG[signal]
    conf = f"GNSS-SDR.internal_fs_sps={int(fs)}\n{role}.item_type={item}\n{role}.doppler_max={dmax}\n{role}.doppler_step={dstep}\n" \
           f"{role}.coherent_integration_time_ms={ms_per_code}\n{role}.max_dwells={nd}\n{role}.bit_transition_flag={'true' if bt else 'false'}\n" \
           f"{role}.blocking=true\n" + (f"{role}.pfa={pfa}\n" if cfar else "")
    b = O.RefAcq(conf, role, ms_per_code, chip_rate, opt, prn, system, sig2, fft_f64=fft_f64)
    b.init(dmax, dstep, 0.0)
    st = b.status()
    # the block sized itself like the case: pcps_acquisition.cc:84-93, acq_conf.cc:113-118
    assert (st.fft_size, st.consumed_samples) == (fft_size, consumed), (st.fft_size, st.consumed_samples)
    assert st.num_doppler_bins == int(np.ceil(2 * dmax / dstep)) and st.use_cfar == int(cfar) and st.bit_transition == int(bt)
    # acq_conf.cc:115-117 in fp32: float(fs) * 0.001F * ms_per_code (4000.0002 at 4 MS/s) — the frozen cases use the exact value, which gives
    # the same fmod(index, samples_per_code) for every index below one code
    assert st.samples_per_code == np.float32(np.float32(fs) * np.float32(0.001)) * np.float32(ms_per_code) and st.samples_per_chip == case["spchip"]
    assert abs(st.samples_per_code - (case["spc"] if not bt else case["spc"])) <= 1e-3 * ms_per_code * 4
    code = np.zeros(fft_size, np.complex64)
    code[:len(case["locals_"][prn])] = case["locals_"][prn]
    b.set_local_code(code)
    return b


@pytest.mark.parametrize("key", sorted(k for k in VEC if not k.startswith(("c3_e1b", "c5_"))))  # cbyte is converted in front of the block by the adapter; 100000 points: below
def test_reference_block_reproduces_the_frozen_vectors(key):
    name, seed = key.split("/seed")
    case = S.acq_case(name, int(seed))
    item = S.ACQ_CASES[name][8]
    stream, k = (np.concatenate(case["items"]), 1) if item == "gr_complex" else (np.concatenate(case["items"]), 2)
    for f64 in (False, True):
        for p in case["prns"][:4]:
            want = VEC[key][str(p)]
            b = _block(name, case, p, f64, 0.001)
            thr = b.status().threshold
            if case["use_cfar"]:
                assert abs(thr - VEC[key]["threshold_pfa_1e-3"]) <= 1e-5 * thr  # calculate_threshold (:884-899; Boost's gamma_p_inv is a stand-in here)
            b.set_threshold(3e38)  # every dwell is taken and the search ends in a negative acquisition: the block's results stay readable
            b.arm()
            consumed = b.work(stream, k)
            st = b.status()
            assert consumed == case["max_dwells"] * case["consumed"]  # every dwell taken (the block's own dwell counter is reset when the search ends, :860-868)
            assert st.state == 0 and st.active == 0 and list(st.events[:st.n_events]) == [2]
            assert st.acq_delay_samples == want["acq_delay_samples"] and st.acq_doppler_hz == want["doppler_hz"], (key, p, f64, st.acq_delay_samples, st.acq_doppler_hz, want)
            ref_stat = want["f64_test_statistics"] if f64 else want["test_statistics"]
            assert abs(st.test_statistics - ref_stat) <= 2e-6 * ref_stat, (key, p, f64, st.test_statistics, ref_stat)
            b.close()


def test_reference_block_detects_with_its_own_threshold_and_stops_at_the_first_dwell():
    """the block's own decision path (pfa threshold, positive acquisition through the channel FSM, Acq_samplestamp_samples) on a strong satellite of the dwell case"""
    name, seed = "c4_l5i_4_dwells", 1
    case = S.acq_case(name, seed)
    p = case["present"][0]
    b = _block(name, case, p, False, 0.001)
    b.arm()
    consumed = b.work(np.concatenate(case["items"]), 1)
    st = b.status()
    # positive acquisition: raised on the channel FSM (Event_valid_acquisition, :353-357), nothing on the message port (the block's own
    # d_positive_acq flag is cleared again at the end of the search, :860-868)
    assert list(st.fsm_events[:st.n_fsm_events]) == [1] and st.n_events == 0 and st.state == 0 and st.active == 0
    dwells = consumed // case["consumed"]
    assert 1 <= dwells <= case["max_dwells"] and consumed == dwells * case["consumed"]
    assert st.test_statistics > st.threshold
    assert st.acq_samplestamp_samples == consumed  # pcps_acquisition.cc:695: the counter after the dwell's block
    # the oracle restatement on the same number of dwells
    conf = O.acq_conf(case["fs"], case["consumed"], case["fft_size"], case["dmax"], case["dstep"], case["spc"], case["spchip"], max_dwells=case["max_dwells"])
    fc = O.acq_set_local_code(conf, case["locals_"][p])
    grid = None
    for d in range(dwells):
        grid = O.acq_dwell(conf, case["sigs"][d], fc, grid, d)
    r = O.acq_statistic(conf, grid, dwells)
    assert (st.acq_delay_samples, st.acq_doppler_hz) == (r.acq_delay_samples, r.acq_doppler_hz)
    assert abs(st.test_statistics - r.test_statistics) <= 2e-6 * r.test_statistics and abs(st.input_power - r.input_power) <= 2e-6 * r.input_power
    b.close()


def test_reference_block_grid_equals_the_oracle_grid():
    """the whole search grid of one dwell, cell by cell (acquisition_core :600-672 against oracle_acq.c), and a 100000-point transform"""
    for name, seed, p in (("c1_gps_4msps", 2, 4), ("c3_e1b_25msps", 1, 7)):
        case = S.acq_case(name, seed)
        b = _block(name, case, p, False, 0.001)
        b.set_threshold(3e38)
        b.arm()
        b.work(case["items"][0], 1)
        st = b.status()
        g = b.grid(st.num_doppler_bins, case["fft_size"])
        conf = O.acq_conf(case["fs"], case["consumed"], case["fft_size"], case["dmax"], case["dstep"], case["spc"], case["spchip"])
        grid = O.acq_dwell(conf, case["sigs"][0], O.acq_set_local_code(conf, case["locals_"][p]), None, 0)
        grid = np.asarray(grid).reshape(g.shape)
        assert np.array_equal(g.view(np.uint32), grid.view(np.uint32)), (name, float(np.max(np.abs(g - grid)) / np.max(grid)))
        want = VEC[f"{name}/seed{seed}"][str(p)]
        assert st.acq_delay_samples == want["acq_delay_samples"] and st.acq_doppler_hz == want["doppler_hz"]
        b.close()
