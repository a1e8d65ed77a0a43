"""The closed-loop oracle against the REFERENCE'S OWN tracking block.

oracle/_ref/libgnssref.so carries src/algorithms/tracking/gnuradio_blocks/dll_pll_veml_tracking.cc and src/algorithms/tracking/libs/dll_pll_conf.cc
compiled UNMODIFIED (oracle/ref_trkblk.cc; stand-ins only for scheduler plumbing and containers: gr::block, pmt, boost::circular_buffer, gflags
variables, matio, glog — oracle/shim/acq/README.txt).  The door plays the GNU Radio scheduler for one block (forecast, general_work,
consume_each moving nitems_read at once like buffer_reader::update_read_pointer).

tests/test_closed_loop.py judges the product against oracle/ref_loop.cc, which drives the reference's objects (multicorrelator, filters,
discriminators, lock detectors, smoothers) through a RESTATEMENT of general_work's glue.  Here that restatement is pinned: on the same stream,
the reference block and the restated loop must emit the SAME records — every field, bit for bit — for every signal and mode the closed-loop
tests use.  With that, "the product agrees with ref_loop.cc" is "the product agrees with the reference's own block".
"""
import numpy as np
import pytest

import _oracle as O
import test_closed_loop as T

pytestmark = pytest.mark.skipif(not O.ref_available(), reason="oracle/_ref not built")


def conf_text(role, fs, **kw):
    """.conf text of a tracking role (keys of dll_pll_conf.cc:36-125)"""
    lines = [f"GNSS-SDR.internal_fs_sps={int(fs)}", f"{role}.item_type=gr_complex"]
    for k, v in kw.items():
        lines.append(f"{role}.{k}={'true' if v is True else 'false' if v is False else v}")
    return "\n".join(lines) + "\n"


def _same_records(a, b, label):
    assert len(a) == len(b) and len(a) > 50, (label, len(a), len(b))
    for f in a.dtype.names:
        if f == "channel":
            continue
        assert np.array_equal(a[f], b[f]), (label, f, int(np.argmax(a[f] != b[f])), a[f][a[f] != b[f]][:3], b[f][a[f] != b[f]][:3])


GPS_KW = dict(pll_bw_hz=35.0, dll_bw_hz=2.0, early_late_space_chips=0.5, pull_in_time_s=1, cn0_samples=20, cn0_min=25, max_lock_fail=50, max_carrier_lock_fail=5000,
              carrier_lock_th=0.5)


def test_gps_l1_block_and_restated_loop_emit_identical_records():
    nsat, seconds = 4, 12.0
    sc = T.scenario(nsat)
    a = [(int(sc["prn"][k]), float(sc["acq_delay"][k]), float(sc["doppler"][k] + sc["acq_err"][k])) for k in range(nsat)]
    blocks = [O.RefTrkBlock(conf_text("Tracking_1C", T.FS, **GPS_KW), "Tracking_1C", "1C", *a[k]) for k in range(nsat)]
    loops = [O.RefLoop(T.ref_params(0), sc["codes"][k], None, *a[k], 0, 0) for k in range(nsat)]
    A, B = [[] for _ in range(nsat)], [[] for _ in range(nsat)]
    for _, x in T.stream_chunks(sc, seconds):
        for k in range(nsat):
            A[k].append(blocks[k].feed(x))
            B[k].append(loops[k].feed(x))
    for k in range(nsat):
        assert blocks[k].state()["state"] == 4 and loops[k].state()["state"] == 4
        _same_records(np.concatenate(A[k]), np.concatenate(B[k]), f"GPS L1 ch {k}")
    for o in blocks + loops:
        o.close()


E1_KW = dict(pll_bw_hz=15.0, dll_bw_hz=0.75, pll_bw_narrow_hz=5.0, dll_bw_narrow_hz=0.2, early_late_space_chips=0.25, very_early_late_space_chips=0.5,
             early_late_space_narrow_chips=0.15, very_early_late_space_narrow_chips=0.5, extend_correlation_symbols=1, pull_in_time_s=1, cn0_samples=20, cn0_min=25,
             max_lock_fail=50, max_carrier_lock_fail=5000, carrier_lock_th=0.5)


@pytest.mark.parametrize("pilot", [True, False])
def test_galileo_e1_block_and_restated_loop_emit_identical_records(pilot):
    """E1: pilot tracking (E1-C with its 25-chip secondary code, data prompt from the E1-B correlator) and data tracking (E1-B alone), VEML, 4 ms codes;
    the block generates its own code tables, the loop gets the generator's through the test"""
    nsat = 3
    sc = T.e1_scenario(nsat, seed=29)
    a = [(int(sc["prn"][k]), float(sc["acq_delay"][k]), float(sc["doppler"][k] + sc["acq_err"][k])) for k in range(nsat)]
    blocks = [O.RefTrkBlock(conf_text("Tracking_1B", T.FS, track_pilot=pilot, **E1_KW), "Tracking_1B", "1B", *a[k]) for k in range(nsat)]
    p = T.e1_ref_params()
    if not pilot:
        p.track_pilot, p.secondary, p.secondary_code_length = 0, 0, 0
        p.secondary_code = b""
    loops = [O.RefLoop(p, sc["pilot_code"][k] if pilot else sc["data_code"][k], sc["data_code"][k] if pilot else None, *a[k], 0, 0) for k in range(nsat)]
    A, B = [[] for _ in range(nsat)], [[] for _ in range(nsat)]
    for _, x in T.e1_chunks(sc, 8.0):
        for k in range(nsat):
            A[k].append(blocks[k].feed(x))
            B[k].append(loops[k].feed(x))
    for k in range(nsat):
        assert blocks[k].state()["state"] == 4 and loops[k].state()["state"] == 4
        _same_records(np.concatenate(A[k]), np.concatenate(B[k]), f"E1 pilot={pilot} ch {k}")
    for o in blocks + loops:
        o.close()


L5_KW = dict(pll_bw_hz=20.0, dll_bw_hz=1.5, pll_bw_narrow_hz=5.0, dll_bw_narrow_hz=0.5, early_late_space_chips=0.5, early_late_space_narrow_chips=0.25,
             pull_in_time_s=1, cn0_samples=20, cn0_min=25, max_lock_fail=50, max_carrier_lock_fail=5000, carrier_lock_th=0.5, track_pilot=False)


@pytest.mark.parametrize("signal", ["L5", "5X"])
def test_l5_band_block_and_restated_loop_emit_identical_records(signal):
    """GPS L5I / Galileo E5aI data tracking: secondary code synchronised and removed, extended integration, Prompt_I / Prompt_Q interchange for L5I"""
    sec_str, spb, interchange = T.L5_BAND[signal]
    nsat, fs, seconds = 2, T.L5_FS, 5.0
    chunk = int(fs * 0.1)
    rng = np.random.default_rng(41 + spb)
    prns = rng.choice(np.arange(1, 33), nsat, replace=False)
    doppler, delay, phase = rng.uniform(-3000, 3000, nsat), rng.uniform(0, 10230, nsat), rng.uniform(0, 2 * np.pi, nsat)
    cn0, acq_err = rng.uniform(44, 49, nsat), rng.uniform(-10, 10, nsat)
    codes = np.stack([T._l5_band_code(signal, int(q)) for q in prns])
    sec = np.array([1 if c == "0" else -1 for c in sec_str], dtype=np.int8)
    n_bits = 8000 // spb
    bits = rng.choice(np.array([-1, 1], dtype=np.int8), (nsat, n_bits))
    sym = (np.repeat(bits, spb, axis=1) * np.tile(sec, n_bits)[None, :]).astype(np.int8)
    acq_delay = np.mod(delay / (10.23e6 * (1 + doppler / 1176.45e6)) * fs, fs / 1000.0)
    p = T.ref_params(0)
    p.fs_in, p.code_period, p.chip_rate, p.carrier_freq = fs, 0.001, 10.23e6, 1176.45e6
    p.pll_bw, p.dll_bw, p.pll_bw_narrow, p.dll_bw_narrow, p.el, p.el_narrow = 20.0, 1.5, 5.0, 0.5, 0.5, 0.25
    p.vector_length, p.veml, p.secondary, p.track_pilot, p.symbols_per_bit, p.samples_per_chip, p.code_length_chips = int(fs / 1000), 0, 1, 0, spb, 1, 10230
    p.interchange_iq = interchange
    p.secondary_code_length, p.data_secondary_code_length = len(sec_str), 0
    p.secondary_code = sec_str.encode()
    role = "Tracking_L5" if signal == "L5" else "Tracking_5X"
    a = [(int(prns[k]), float(acq_delay[k]), float(doppler[k] + acq_err[k])) for k in range(nsat)]
    blocks = [O.RefTrkBlock(conf_text(role, fs, **L5_KW), role, signal, *a[k]) for k in range(nsat)]
    loops = [O.RefLoop(p, codes[k], None, *a[k], 0, 0) for k in range(nsat)]
    A, B = [[] for _ in range(nsat)], [[] for _ in range(nsat)]
    n_total = int(seconds * fs)
    for first in range(0, n_total, chunk):
        x = O.gen_if(min(chunk, n_total - first), first, fs, codes, doppler, delay, phase, cn0, nav_bits=sym, periods_per_bit=1, f_chip=10.23e6, f_carrier=1176.45e6, seed=83)
        for k in range(nsat):
            A[k].append(blocks[k].feed(x))
            B[k].append(loops[k].feed(x))
    for k in range(nsat):
        assert blocks[k].state()["state"] == 4 and loops[k].state()["state"] == 4
        _same_records(np.concatenate(A[k]), np.concatenate(B[k]), f"{signal} ch {k}")
    for o in blocks + loops:
        o.close()
