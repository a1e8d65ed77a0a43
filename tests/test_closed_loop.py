"""Closed tracking loop on identical synthetic IF (north_star: "tracking-loop lock and PRN/CN0 within 0.1 dB over 60 s").

Product: B200TrackingBank (host C++ loop closure, gnss_sim_receiver_b200/host/) over the GPU multicorrelator, selected through
the `.conf` implementation string.  Oracle: one loop per channel made of the reference's OWN objects (oracle/ref_loop.cc in
oracle/_ref: Cpu_Multicorrelator_Real_Codes over the volk_gnsssdr protokernels, Tracking_loop_filter, Tracking_FLL_PLL_filter,
discriminators, lock detectors, Exponential_Smoother), glue restated from dll_pll_veml_tracking.cc — and that restatement is itself pinned:
tests/test_trk_refblock_cpu.py has the reference's OWN dll_pll_veml_tracking.cc (compiled unmodified into oracle/_ref) emit bit-identical
records on the same streams, for every signal and mode used here.

Gates: every channel the oracle keeps locked stays locked in the product, same PRN set, same number of
telemetry symbols with identical signs, CN0 (M2M4) within 0.1 dB, Doppler within 0.5 Hz RMS.
Duration: GB200_LOOP_SECONDS (default 60).
"""
import os

import numpy as np
import pytest

import _oracle as O

needs_ref = pytest.mark.skipif(not O.ref_available(), reason="oracle/_ref not built")

FS = 4e6
SECONDS = float(os.environ.get("GB200_LOOP_SECONDS", "60"))
CHUNK = 400000  # 100 ms of stream per push

CONF = """
[GNSS-SDR]
GNSS-SDR.internal_fs_sps=4000000
Tracking_1C.implementation=GPS_L1_CA_DLL_PLL_Tracking_B200
Tracking_1C.item_type=gr_complex
Tracking_1C.pll_bw_hz=35.0
Tracking_1C.dll_bw_hz=2.0
Tracking_1C.early_late_space_chips=0.5
Tracking_1C.pull_in_time_s=1
Tracking_1C.cn0_samples=20
Tracking_1C.cn0_min=25
Tracking_1C.max_lock_fail=50
Tracking_1C.max_carrier_lock_fail=5000
Tracking_1C.carrier_lock_th=0.5
"""


def scenario(nsat, seed=11):
    rng = np.random.default_rng(seed)
    prns = rng.choice(np.arange(1, 33), nsat, replace=False)
    sc = dict(prn=prns, doppler=rng.uniform(-4000, 4000, nsat), delay=rng.uniform(0, 1023, nsat), phase=rng.uniform(0, 2 * np.pi, nsat),
              cn0=rng.uniform(40, 50, nsat), acq_err=rng.uniform(-15, 15, nsat))
    bits = rng.choice(np.array([-1, 1], dtype=np.int8), (nsat, 3000))
    # telemetry preamble 10001011: the block synchronises bits on it (every 300 bits in a real message; every 50 here so that
    # short runs reach narrow tracking too)
    pre = np.array([1, -1, -1, -1, 1, -1, 1, 1], dtype=np.int8)
    for b in range(0, 3000, 50):
        bits[:, b:b + 8] = pre
    sc["bits"] = bits
    sc["codes"] = np.stack([O.gps_l1_ca_code(int(p)) for p in prns])
    # acquisition hand-over: code start of chip 0 in samples, modulo one code period
    f_chip = 1.023e6 * (1 + sc["doppler"] / 1575.42e6)
    sc["acq_delay"] = np.mod(sc["delay"] / f_chip * FS, FS / 1000.0)
    return sc


def ref_params(arch=0):
    p = O.RefLoopParams()
    p.fs_in, p.code_period, p.chip_rate, p.carrier_freq, p.carrier_lock_th = FS, 0.001, 1.023e6, 1575.42e6, 0.5
    p.pll_bw, p.dll_bw, p.fll_bw, p.pll_bw_narrow, p.dll_bw_narrow = 35.0, 2.0, 35.0, 5.0, 0.75
    p.el, p.vel, p.el_narrow, p.vel_narrow = 0.5, 0.5, 0.15, 0.5
    p.cn0_smoother_alpha, p.lock_smoother_alpha = 0.002, 0.002
    p.vector_length, p.veml, p.secondary, p.track_pilot, p.symbols_per_bit, p.samples_per_chip, p.code_length_chips = 4000, 0, 0, 0, 20, 1, 1023
    p.pll_order, p.dll_order, p.extend = 3, 2, 1
    p.cn0_samples, p.cn0_min, p.max_code_lock_fail, p.max_carrier_lock_fail = 20, 25, 50, 5000
    p.cn0_smoother_samples, p.lock_smoother_samples = 200, 25
    p.carrier_aiding, p.enable_fll_pull_in, p.enable_fll_steady, p.interchange_iq, p.arch, p.correlation_length_ms = 1, 0, 0, 0, arch, 1
    p.pull_in_time_s, p.bit_sync_time_limit_s = 1, 20
    pre = "".join(c * 20 for c in "10001011")
    p.secondary_code_length, p.data_secondary_code_length = 160, 0
    p.secondary_code = pre.encode()
    return p


def stream_chunks(sc, seconds):
    n_total = int(seconds * FS)
    for first in range(0, n_total, CHUNK):
        n = min(CHUNK, n_total - first)
        yield first, O.gen_if(n, first, FS, sc["codes"], sc["doppler"], sc["delay"], sc["phase"], sc["cn0"], nav_bits=sc["bits"], seed=77)


def run_ref_loops(sc, seconds, arch=0):
    nsat = len(sc["prn"])
    loops = [O.RefLoop(ref_params(arch), sc["codes"][k], None, int(sc["prn"][k]), float(sc["acq_delay"][k]), float(sc["doppler"][k] + sc["acq_err"][k]), 0, 0)
             for k in range(nsat)]
    recs = [[] for _ in range(nsat)]
    for _, x in stream_chunks(sc, seconds):
        for k, lp in enumerate(loops):
            recs[k].append(lp.feed(x))
    states = [lp.state() for lp in loops]
    for lp in loops:
        lp.close()
    return [np.concatenate(r) for r in recs], states


@needs_ref
def test_reference_object_loop_locks_on_the_synthetic_signal():
    """Sanity of the oracle itself (CPU): 4 s, 2 channels — reaches narrow tracking, Doppler and CN0 near the truth."""
    sc = scenario(2, seed=3)
    recs, states = run_ref_loops(sc, 4.0)
    for k in range(2):
        assert states[k]["state"] == 4, states[k]
        assert abs(states[k]["carrier_doppler_hz"] - sc["doppler"][k]) < 5.0
        assert abs(states[k]["cn0_db_hz"] - sc["cn0"][k]) < 2.0
        r = recs[k]
        assert len(r) > 50 and np.all(r["valid"] == 1)
        # telemetry symbols: 20 ms sums of the prompt reproduce the generated nav bits up to the PLL's half-cycle sign
        sym = np.sign(r["prompt_i"])
        f_chip = 1.023e6 * (1 + sc["doppler"][k] / 1575.42e6)
        t_end = (r["tracking_sample_counter"].astype(np.float64) - 2000) / FS  # middle of the last code period of the bit
        bit_idx = np.floor((t_end * f_chip - sc["delay"][k]) / (1023 * 20)).astype(int) % sc["bits"].shape[1]
        truth = sc["bits"][k][bit_idx]
        agree = np.mean(sym == truth)
        assert agree > 0.99 or agree < 0.01, agree


@pytest.mark.gpu
@needs_ref
def test_closed_loop_parity_vs_reference_objects():
    from gnss_sim_receiver_b200 import _hostapi as H
    nsat = 8
    sc = scenario(nsat)
    bank = H.TrackingBank(CONF, "Tracking_1C", nsat)
    for k in range(nsat):
        bank.start(k, int(sc["prn"][k]), float(sc["acq_delay"][k]), float(sc["doppler"][k] + sc["acq_err"][k]), 0, 0)
    loops = [O.RefLoop(ref_params(0), sc["codes"][k], None, int(sc["prn"][k]), float(sc["acq_delay"][k]), float(sc["doppler"][k] + sc["acq_err"][k]), 0, 0)
             for k in range(nsat)]
    got, ref = [[] for _ in range(nsat)], [[] for _ in range(nsat)]
    for _, x in stream_chunks(sc, SECONDS):
        bank.push(x)
        r = bank.step()
        for k in range(nsat):
            got[k].append(r[r["channel"] == k])
            ref[k].append(loops[k].feed(x))
    launches = bank.launches
    for k in range(nsat):
        g, r = np.concatenate(got[k]), np.concatenate(ref[k])
        gs, rs = bank.channel_state(k), loops[k].state()
        assert rs["state"] == 4, (k, rs)  # the oracle keeps every channel of this scenario locked ...
        assert gs["state"] == 4, (k, gs)  # ... and so does the product
        assert np.all(g["prn"] == sc["prn"][k]) and np.all(g["valid"] == 1) and np.all(r["valid"] == 1)
        assert len(g) == len(r), (k, len(g), len(r))
        # block lengths are floor(T_prn + remnant): a last-bit difference in the code NCO moves a boundary by one sample for a
        # period, never more, and never accumulates (the remnant carries it)
        d_cnt = np.abs(g["tracking_sample_counter"].astype(np.int64) - r["tracking_sample_counter"].astype(np.int64))
        assert d_cnt.max() <= 1, (k, d_cnt.max())
        assert np.array_equal(np.sign(g["prompt_i"]), np.sign(r["prompt_i"])), k
        half = len(g) // 2
        d_cn0 = np.max(np.abs(g["cn0_db_hz"][half:] - r["cn0_db_hz"][half:]))
        d_dop = np.sqrt(np.mean((g["carrier_doppler_hz"][half:] - r["carrier_doppler_hz"][half:]) ** 2))
        assert d_cn0 < 0.1, (k, d_cn0)   # dB
        assert d_dop < 0.5, (k, d_dop)   # Hz RMS
        # and both sit on the truth: CN0 within 2 dB, mean Doppler of the second half within 2 Hz
        assert abs(gs["cn0_db_hz"] - sc["cn0"][k]) < 2.0 and abs(np.mean(g["carrier_doppler_hz"][half:]) - sc["doppler"][k]) < 2.0
    # one launch per code period for all channels (plus the odd extra round when read pointers straddle a push)
    assert launches < 1.2 * SECONDS * 1000, launches
    for lp in loops:
        lp.close()
    bank.close()


@pytest.mark.gpu
@needs_ref
def test_thread_per_block_form_matches_the_lock_step_bank():
    """The GNU Radio shape (one host thread per tracking block calling general_work through the per-channel facade, requests
    batched by the engine) gives the same records as the lock-step bank."""
    from gnss_sim_receiver_b200 import _hostapi as H
    nsat, seconds = 4, 4.0  # pull-in ends 2 s after the acquisition stamp, bit sync follows within 1 s
    sc = scenario(nsat, seed=5)
    x = np.concatenate([c for _, c in stream_chunks(sc, seconds)])
    a, b = H.TrackingBank(CONF, "Tracking_1C", nsat), H.TrackingBank(CONF, "Tracking_1C", nsat)
    for bank in (a, b):
        for k in range(nsat):
            bank.start(k, int(sc["prn"][k]), float(sc["acq_delay"][k]), float(sc["doppler"][k] + sc["acq_err"][k]), 0, 0)
    ra = []
    for first in range(0, len(x), CHUNK):
        a.push(x[first:first + CHUNK])
        ra.append(a.step())
    ra = np.concatenate(ra)
    rb = b.run_threaded(x, 0, nsat)
    for k in range(nsat):
        ga, gb = ra[ra["channel"] == k], rb[rb["channel"] == k]
        assert len(ga) == len(gb) and len(ga) > 10
        assert np.array_equal(ga["tracking_sample_counter"], gb["tracking_sample_counter"])
        assert np.allclose(ga["prompt_i"], gb["prompt_i"], rtol=0, atol=1e-3 * np.max(np.abs(ga["prompt_i"])))
        assert np.allclose(ga["carrier_doppler_hz"], gb["carrier_doppler_hz"], atol=1e-6)
    assert b.launches < a.launches * 4  # requests of concurrent channels share launches
    a.close()
    b.close()


# ------------------------------------------------------------------------------------------------ Galileo E1, pilot + data
E1_SECONDS = float(os.environ.get("GB200_LOOP_SECONDS_E1", "12"))
E1_CS25 = "0011100000001010110110010"  # E1-C secondary code

CONF_E1 = """
GNSS-SDR.internal_fs_sps=4000000
Tracking_1B.implementation=Galileo_E1_DLL_PLL_VEML_Tracking_B200
Tracking_1B.item_type=gr_complex
Tracking_1B.track_pilot=true
Tracking_1B.pll_bw_hz=15.0
Tracking_1B.dll_bw_hz=0.75
Tracking_1B.pll_bw_narrow_hz=5.0
Tracking_1B.dll_bw_narrow_hz=0.2
Tracking_1B.early_late_space_chips=0.25
Tracking_1B.very_early_late_space_chips=0.5
Tracking_1B.early_late_space_narrow_chips=0.15
Tracking_1B.very_early_late_space_narrow_chips=0.5
Tracking_1B.extend_correlation_symbols=1
Tracking_1B.pull_in_time_s=1
Tracking_1B.cn0_samples=20
Tracking_1B.cn0_min=25
Tracking_1B.max_lock_fail=50
Tracking_1B.max_carrier_lock_fail=5000
Tracking_1B.carrier_lock_th=0.5
"""


def e1_code(sig, prn):
    buf = np.zeros(8184, dtype=np.float32)
    O.ref().ref_galileo_e1_code_gen_sinboc11_float(buf.ctypes.data_as(O.C.c_void_p), 8184, sig, int(prn))
    return buf


def e1_scenario(nsat, seed=21):
    """E1 open service: data component E1-B (250 symbols/s) minus pilot E1-C with its 25-chip secondary code, equal power."""
    rng = np.random.default_rng(seed)
    prns = rng.choice(np.arange(1, 31), nsat, replace=False)
    sc = dict(prn=prns, doppler=rng.uniform(-3000, 3000, nsat), delay=rng.uniform(0, 8184, nsat), phase=rng.uniform(0, 2 * np.pi, nsat),
              cn0=rng.uniform(42, 48, nsat), acq_err=rng.uniform(-10, 10, nsat))
    sc["data_code"] = np.stack([e1_code(b"1B", p) for p in prns])
    sc["pilot_code"] = np.stack([e1_code(b"1C", p) for p in prns])
    sec = np.array([1 if c == "0" else -1 for c in E1_CS25], dtype=np.int8)
    bits = rng.choice(np.array([-1, 1], dtype=np.int8), (nsat, 3000))
    # the generator sums independent components: rows 0..nsat-1 = data, nsat..2nsat-1 = pilot (sign folded into the table)
    sc["gen_codes"] = np.concatenate([sc["data_code"], -sc["pilot_code"]])
    sc["gen_bits"] = np.concatenate([bits, np.tile(sec, (nsat, 120))])
    f_entry = 2.046e6 * (1 + sc["doppler"] / 1575.42e6)  # table entries per second (2 per chip)
    sc["acq_delay"] = np.mod(sc["delay"] / f_entry * FS, FS * 0.004)
    return sc


def e1_chunks(sc, seconds):
    two = lambda a: np.concatenate([a, a])
    n_total = int(seconds * FS)
    for first in range(0, n_total, CHUNK):
        n = min(CHUNK, n_total - first)
        yield first, O.gen_if(n, first, FS, sc["gen_codes"], two(sc["doppler"]), two(sc["delay"]), two(sc["phase"]), two(sc["cn0"]) - 3.0103,
                              nav_bits=sc["gen_bits"], periods_per_bit=1, f_chip=2.046e6, seed=78)


def e1_ref_params():
    p = ref_params(0)
    p.code_period, p.chip_rate = 0.004, 1.023e6
    p.pll_bw, p.dll_bw, p.fll_bw, p.pll_bw_narrow, p.dll_bw_narrow = 15.0, 0.75, 35.0, 5.0, 0.2
    p.el, p.vel, p.el_narrow, p.vel_narrow = 0.25, 0.5, 0.15, 0.5
    p.vector_length, p.veml, p.secondary, p.track_pilot, p.symbols_per_bit, p.samples_per_chip, p.code_length_chips = 16000, 1, 1, 1, 1, 2, 4092
    p.correlation_length_ms = 4
    p.secondary_code_length, p.data_secondary_code_length = 25, 0
    p.secondary_code = E1_CS25.encode()
    return p


@pytest.mark.gpu
@needs_ref
def test_closed_loop_galileo_e1_pilot_vs_reference_objects():
    """C2 shape: Galileo E1, VEML (5 taps) on the pilot with secondary-code lock, one-tap data correlator for the symbols —
    the product's bank against loops made of the reference's own objects, on the same stream."""
    from gnss_sim_receiver_b200 import _hostapi as H
    nsat = 4
    sc = e1_scenario(nsat)
    bank = H.TrackingBank(CONF_E1, "Tracking_1B", nsat)
    loops = []
    for k in range(nsat):
        # no set_codes: start_tracking builds the E1-C / E1-B tables itself (dll_pll_veml_tracking.cc:684-699); the reference loop
        # gets the reference generator's tables
        bank.start(k, int(sc["prn"][k]), float(sc["acq_delay"][k]), float(sc["doppler"][k] + sc["acq_err"][k]), 0, 0)
        loops.append(O.RefLoop(e1_ref_params(), sc["pilot_code"][k], sc["data_code"][k], int(sc["prn"][k]), float(sc["acq_delay"][k]),
                               float(sc["doppler"][k] + sc["acq_err"][k]), 0, 0))
    got, ref = [[] for _ in range(nsat)], [[] for _ in range(nsat)]
    for _, x in e1_chunks(sc, E1_SECONDS):
        bank.push(x)
        r = bank.step()
        for k in range(nsat):
            got[k].append(r[r["channel"] == k])
            ref[k].append(loops[k].feed(x))
    for k in range(nsat):
        g, r = np.concatenate(got[k]), np.concatenate(ref[k])
        gs, rs = bank.channel_state(k), loops[k].state()
        assert rs["state"] == 4, (k, rs)
        assert gs["state"] == 4, (k, gs)
        assert len(g) == len(r) and len(g) > 500, (k, len(g), len(r))
        d_cnt = np.abs(g["tracking_sample_counter"].astype(np.int64) - r["tracking_sample_counter"].astype(np.int64))
        assert d_cnt.max() <= 1, (k, d_cnt.max())
        # telemetry symbols come from the DATA component's prompt: identical signs
        assert np.array_equal(np.sign(g["prompt_i"]), np.sign(r["prompt_i"])), k
        half = len(g) // 2
        assert np.max(np.abs(g["cn0_db_hz"][half:] - r["cn0_db_hz"][half:])) < 0.1, k
        assert np.sqrt(np.mean((g["carrier_doppler_hz"][half:] - r["carrier_doppler_hz"][half:]) ** 2)) < 0.5, k
        assert abs(np.mean(g["carrier_doppler_hz"][half:]) - sc["doppler"][k]) < 2.0
    for lp in loops:
        lp.close()
    bank.close()


# ------------------------------------------------------------------------------------------------ device-resident loop closure
def _device_loop(nsat, sc, iq_dtype=0, **cfg_over):
    from gnss_sim_receiver_b200.tracking import TrackingEngine, TrackingLoop, loop_config
    eng = TrackingEngine(max_channels=nsat, max_code_len=1023, iq_dtype=iq_dtype, ring_samples=4 * CHUNK, max_block=2 * 4000, max_cmds=nsat)
    kw = dict(pll_bw_hz=35.0, dll_bw_hz=2.0, early_late_space_chips=0.5, pull_in_time_s=1, cn0_samples=20, cn0_min=25, max_code_lock_fail=50,
              max_carrier_lock_fail=5000, carrier_lock_th=0.5)
    kw.update(cfg_over)
    cfg = loop_config("1C", FS, **kw)
    for k in range(nsat):
        eng.set_code(k, sc["codes"][k])
    loop = TrackingLoop(eng, cfg, nsat)
    for k in range(nsat):
        loop.start(k, int(sc["prn"][k]), float(sc["acq_delay"][k]), float(sc["doppler"][k] + sc["acq_err"][k]), 0, 0)
    return eng, loop


@pytest.mark.gpu
@needs_ref
def test_device_loop_parity_vs_reference_objects():
    """gb200_loop_*: the loop closure on the DEVICE (csrc/trk_loop.cuh: state in HBM, one thread per channel, commands formed on the
    device) against the loops made of the reference's own objects, same gates as the host bank above: every channel locked, same
    number of telemetry symbols with identical signs, sample counters within one sample, CN0 within 0.1 dB, Doppler within 0.5 Hz RMS."""
    nsat = 8
    seconds = min(SECONDS, 30.0)
    sc = scenario(nsat)
    eng, loop = _device_loop(nsat, sc)
    loops = [O.RefLoop(ref_params(0), sc["codes"][k], None, int(sc["prn"][k]), float(sc["acq_delay"][k]), float(sc["doppler"][k] + sc["acq_err"][k]), 0, 0)
             for k in range(nsat)]
    got, ref = [[] for _ in range(nsat)], [[] for _ in range(nsat)]
    for _, x in stream_chunks(sc, seconds):
        r = loop.step(x)
        for k in range(nsat):
            got[k].append(r[r["channel"] == k])
            ref[k].append(loops[k].feed(x))
    tail = loop.step(None)  # the commands formed by the last step are still pending: nothing new may appear without samples
    assert len(tail) <= nsat
    for k in range(nsat):
        g, r = np.concatenate(got[k]), np.concatenate(ref[k])
        gs, rs = loop.channel(k), loops[k].state()
        assert rs["state"] == 4 and gs["state"] == 4, (k, gs, rs)
        assert np.all(g["prn"] == sc["prn"][k]) and np.all(g["valid"] == 1)
        # the device loop runs one round behind the host loops (its last command is formed but not yet correlated): compare the common part
        m = min(len(g), len(r))
        assert m > 0.9 * len(r) and len(r) - len(g) <= 6, (k, len(g), len(r))
        g, r = g[:m], r[:m]
        d_cnt = np.abs(g["tracking_sample_counter"].astype(np.int64) - r["tracking_sample_counter"].astype(np.int64))
        assert d_cnt.max() <= 1, (k, d_cnt.max())
        assert np.array_equal(np.sign(g["prompt_i"]), np.sign(r["prompt_i"])), k
        half = m // 2
        d_cn0 = np.max(np.abs(g["cn0_db_hz"][half:] - r["cn0_db_hz"][half:]))
        d_dop = np.sqrt(np.mean((g["carrier_doppler_hz"][half:] - r["carrier_doppler_hz"][half:]) ** 2))
        print(f"[device loop] ch {k}: records {m}, CN0 diff {d_cn0:.4f} dB, Doppler diff {d_dop:.4f} Hz rms")
        assert d_cn0 < 0.1, (k, d_cn0)
        assert d_dop < 0.5, (k, d_dop)
        assert abs(gs["cn0_db_hz"] - sc["cn0"][k]) < 2.0 and abs(np.mean(g["carrier_doppler_hz"][half:]) - sc["doppler"][k]) < 2.0
    for lp in loops:
        lp.close()
    loop.close()
    eng.close()


@pytest.mark.gpu
def test_device_loop_matches_host_bank_and_loses_lock_like_it():
    """Device loop against the product's host loop (B200TrackingBank) on a signal that disappears: both lock, both declare loss of lock
    (carrier / CN0 fail counters, dll_pll_veml_tracking.cc:991-1027) and fall back to standby, records of the locked stretch agree."""
    from gnss_sim_receiver_b200 import _hostapi as H
    nsat = 4
    sc = scenario(nsat, seed=7)
    seconds_on, seconds = 3.0, 5.0
    # (on noise the M2M4 estimate sits near 29 dB-Hz, above cn0_min: it is the carrier lock test that fails, once per code period)
    eng, loop = _device_loop(nsat, sc, max_carrier_lock_fail=400)
    bank = H.TrackingBank(CONF.replace("max_carrier_lock_fail=5000", "max_carrier_lock_fail=400"), "Tracking_1C", nsat)
    for k in range(nsat):
        bank.start(k, int(sc["prn"][k]), float(sc["acq_delay"][k]), float(sc["doppler"][k] + sc["acq_err"][k]), 0, 0)
    got, ref = [], []
    for first, x in stream_chunks(sc, seconds):
        if first >= seconds_on * FS:  # the satellites vanish: noise only
            x = O.gen_if(len(x), first, FS, sc["codes"], sc["doppler"], sc["delay"], sc["phase"], np.full(nsat, -30.0), nav_bits=sc["bits"], seed=77)
        got.append(loop.step(x))
        bank.push(x)
        ref.append(bank.step())
    got, ref = np.concatenate(got), np.concatenate(ref)
    for k in range(nsat):
        g, r = got[got["channel"] == k], ref[ref["channel"] == k]
        assert loop.channel(k)["state"] == 0 and bank.channel_state(k)["state"] == 0, (k, loop.channel(k), bank.channel_state(k))
        gl, rl = g[g["valid"] == 0], r[r["valid"] == 0]
        assert len(gl) == 1 and len(rl) == 1  # exactly one loss-of-lock record each
        gv, rv = g[g["valid"] == 1], r[r["valid"] == 1]
        m = min(len(gv), len(rv))
        assert m > 20 and abs(len(gv) - len(rv)) <= 15  # on noise the two loops wander apart: the fail counters cross their limit at different periods
        on = gv["tracking_sample_counter"][:m] < seconds_on * FS
        assert np.array_equal(np.sign(gv["prompt_i"][:m][on]), np.sign(rv["prompt_i"][:m][on]))
        assert np.max(np.abs(gv["carrier_doppler_hz"][:m][on] - rv["carrier_doppler_hz"][:m][on])) < 1.0
        # the loss is declared within a few hundred code periods of the host loop's (noise-driven counters)
        assert abs(int(gl["tracking_sample_counter"][0]) - int(rl["tracking_sample_counter"][0])) <= 320 * 4000
    bank.close()
    loop.close()
    eng.close()


@pytest.mark.gpu
def test_device_loop_pipelined_form_gives_the_same_records():
    """gb200_loop_submit / gb200_loop_collect (two steps in flight) against gb200_loop_step: identical records, in the same order."""
    nsat = 4
    sc = scenario(nsat, seed=13)
    x = np.concatenate([c for _, c in stream_chunks(sc, 3.0)])
    e1, l1 = _device_loop(nsat, sc)
    e2, l2 = _device_loop(nsat, sc)
    per = 4000
    a, b = [], []
    blocks = [np.ascontiguousarray(x[i:i + per]) for i in range(0, len(x), per)]
    for blk in blocks:
        a.append(l1.step(blk))
    l2.submit(blocks[0])
    for blk in blocks[1:]:
        l2.submit(blk)
        b.append(l2.collect())
    b.append(l2.collect())
    a, b = np.concatenate(a), np.concatenate(b)
    assert len(a) > 40 and len(a) == len(b)
    for k in range(nsat):
        assert a[a["channel"] == k].tobytes() == b[b["channel"] == k].tobytes()
    from gnss_sim_receiver_b200._capi import GB200Error
    l2.submit(blocks[0])
    l2.submit(blocks[1])
    with pytest.raises(GB200Error):
        l2.submit(blocks[2])  # a third step in flight is refused
    l2.collect()
    l2.collect()
    for l, e in ((l1, e1), (l2, e2)):
        l.close()
        e.close()


@pytest.mark.gpu
def test_banked_loop_gives_the_records_of_one_bank():
    """BankedTrackingLoop (the GPU's channels as two banks, one bank's loop kernel under the other's correlator) against ONE TrackingLoop over the same
    channels and stream: identical records per channel, channel numbers global; uneven split (5 channels = 3 + 2)."""
    from gnss_sim_receiver_b200.tracking import BankedTrackingLoop, loop_config
    nsat = 5
    sc = scenario(nsat, seed=17)
    x = np.concatenate([c for _, c in stream_chunks(sc, 3.0)])
    e1, l1 = _device_loop(nsat, sc)
    cfg = loop_config("1C", FS, pll_bw_hz=35.0, dll_bw_hz=2.0, early_late_space_chips=0.5, pull_in_time_s=1, cn0_samples=20, cn0_min=25, max_code_lock_fail=50,
                      max_carrier_lock_fail=5000, carrier_lock_th=0.5)
    bl = BankedTrackingLoop(cfg, nsat, n_banks=2, max_code_len=1023, ring_samples=4 * CHUNK, max_block=2 * 4000)
    assert [lp.n for lp in bl.loops] == [3, 2]
    for k in range(nsat):
        bl.set_code(k, sc["codes"][k])
        bl.start(k, int(sc["prn"][k]), float(sc["acq_delay"][k]), float(sc["doppler"][k] + sc["acq_err"][k]), 0, 0)
    per = 4000
    blocks = [np.ascontiguousarray(x[i:i + per]) for i in range(0, len(x), per)]
    a, b = [], []
    for blk in blocks:
        a.append(l1.step(blk))
    bl.submit(blocks[0])
    for blk in blocks[1:]:
        bl.submit(blk)
        b.append(bl.collect())
    b.append(bl.collect())
    assert bl.last_step["commands"] == l1.last_step["commands"] == nsat
    a, b = np.concatenate(a), np.concatenate(b)
    assert len(a) > 50 and len(a) == len(b) and set(b["channel"]) == set(range(nsat))
    for k in range(nsat):
        assert a[a["channel"] == k].tobytes() == b[b["channel"] == k].tobytes()
        assert bl.channel(k)["state"] == l1.channel(k)["state"] >= 2
    with pytest.raises(IndexError):
        bl.start(nsat, 1, 0.0, 0.0)
    with pytest.raises(ValueError):
        BankedTrackingLoop(cfg, 1, n_banks=2)
    bl.close()
    l1.close()
    e1.close()


@pytest.mark.gpu
def test_device_loop_many_channels_run_kernel():
    """2048 channels at 25 MS/s (the run-based correlator kernel drives the loop): every channel reaches narrow tracking on its own
    satellite of a 32-satellite stream and reports Doppler / CN0 near the truth; one correlate launch + one loop kernel per code period."""
    from gnss_sim_receiver_b200 import _capi
    from gnss_sim_receiver_b200.tracking import TrackingEngine, TrackingLoop, loop_config
    fs, n, nsat, nch, periods = 25e6, 25000, 32, 2048, 2600
    rng = np.random.default_rng(5)
    codes = np.stack([O.gps_l1_ca_code(p) for p in range(1, 33)])
    dopp, delay, ph0, cn0 = rng.uniform(-4000, 4000, nsat), rng.uniform(0, 1023, nsat), rng.uniform(0, 6.28, nsat), rng.uniform(42, 48, nsat)
    bits = rng.choice(np.array([-1, 1], dtype=np.int8), (nsat, 400))
    pre = np.array([1, -1, -1, -1, 1, -1, 1, 1], dtype=np.int8)
    for b in range(0, 400, 10):
        bits[:, b:b + 8] = pre
    eng = TrackingEngine(max_channels=nch, max_code_len=1023, ring_samples=16 * n, max_block=n + 8, max_cmds=nch)
    cfg = loop_config("1C", fs, pull_in_time_s=1, cn0_samples=20, cn0_min=25, max_code_lock_fail=50, max_carrier_lock_fail=5000, carrier_lock_th=0.5, early_late_space_chips=0.5)
    sat = np.arange(nch) % nsat
    for c in range(nch):
        eng.set_code(c, codes[sat[c]])
    loop = TrackingLoop(eng, cfg, nch)
    assert loop.kernel.startswith("trk_run_kernel"), loop.kernel
    f_chip = 1.023e6 * (1 + dopp / 1575.42e6)
    acq_delay = np.mod(delay / f_chip * fs, fs / 1000.0)
    for c in range(nch):
        loop.start(c, int(sat[c]) + 1, float(acq_delay[sat[c]] + rng.uniform(-2, 2)), float(dopp[sat[c]] + rng.uniform(-30, 30)), 0, 0)
    lib = _capi.lib()
    l0 = lib.gb200_launch_count()
    nrec = 0
    for p in range(periods):
        x = O.gen_if(n, p * n, fs, codes, dopp, delay, ph0, cn0, nav_bits=bits, seed=9)
        nrec += len(loop.step(x, copy=False))
    launches = lib.gb200_launch_count() - l0
    states = [loop.channel(c) for c in range(0, nch, 37)]
    assert all(s["state"] == 4 for s in states), [s["state"] for s in states]
    for c, s in zip(range(0, nch, 37), states):
        assert abs(s["carrier_doppler_hz"] - dopp[sat[c]]) < 15.0 and abs(s["cn0_db_hz"] - cn0[sat[c]]) < 3.5, (c, s, dopp[sat[c]], cn0[sat[c]])  # 31 other satellites in the stream: their cross-correlation reads as noise
    assert nrec > nch * (periods - 2200) // 20 // 2
    assert launches <= 2 * periods + 2 * 40, launches  # correlate + loop kernel per period, a few catch-up rounds
    loop.close()
    eng.close()


# ------------------------------------------------------------------------------------------------ device loop on the other served signals
def _compare_device_with_ref(loop, loops, got, ref, nsat, truth_doppler, label, min_records, data_field="prompt_i"):
    for k in range(nsat):
        g, r = np.concatenate(got[k]), np.concatenate(ref[k])
        gs, rs = loop.channel(k), loops[k].state()
        assert rs["state"] == 4 and gs["state"] == 4, (label, k, gs, rs)
        m = min(len(g), len(r))
        assert m > min_records and len(r) - len(g) <= 6, (label, k, len(g), len(r))
        g, r = g[:m], r[:m]
        d_cnt = np.abs(g["tracking_sample_counter"].astype(np.int64) - r["tracking_sample_counter"].astype(np.int64))
        assert d_cnt.max() <= 1, (label, k, d_cnt.max())
        assert np.array_equal(g["correlation_length_ms"], r["correlation_length_ms"]), (label, k)
        # telemetry symbols: identical signs on the data-bearing arm (with d_interchange_iq the record's Prompt_Q; the other arm is noise
        # around zero), and both arms close in value
        assert np.array_equal(np.sign(g[data_field]), np.sign(r[data_field])), (label, k)
        amp = np.median(np.abs(r[data_field]))
        for f in ("prompt_i", "prompt_q"):
            assert np.max(np.abs(g[f][m // 4:] - r[f][m // 4:])) < 0.05 * amp, (label, k, f)
        half = m // 2
        d_cn0 = np.max(np.abs(g["cn0_db_hz"][half:] - r["cn0_db_hz"][half:]))
        d_dop = np.sqrt(np.mean((g["carrier_doppler_hz"][half:] - r["carrier_doppler_hz"][half:]) ** 2))
        print(f"[device loop {label}] ch {k}: records {m}, CN0 diff {d_cn0:.4f} dB, Doppler diff {d_dop:.4f} Hz rms")
        assert d_cn0 < 0.1, (label, k, d_cn0)
        assert d_dop < 0.5, (label, k, d_dop)
        assert abs(np.mean(g["carrier_doppler_hz"][half:]) - truth_doppler[k]) < 2.0, (label, k)


@pytest.mark.gpu
@needs_ref
def test_device_loop_galileo_e1b_vs_reference_objects():
    """The device loop on Galileo E1B (data component, track_pilot=false: dll_pll_veml_tracking.cc:264-292): 4 ms codes of 8184 table
    entries, the five-tap VEML correlator and discriminator, no secondary code, one symbol per code — against loops made of the
    reference's own objects on the E1 open-service stream (E1-B minus E1-C), same gates as for GPS L1."""
    from gnss_sim_receiver_b200.tracking import TrackingEngine, TrackingLoop, loop_config
    nsat = 4
    sc = e1_scenario(nsat, seed=23)
    eng = TrackingEngine(max_channels=nsat, max_code_len=8184, ring_samples=4 * CHUNK, max_block=2 * 16000, max_cmds=nsat)
    cfg = loop_config("1B", FS, pll_bw_hz=15.0, dll_bw_hz=0.75, pll_bw_narrow_hz=5.0, dll_bw_narrow_hz=0.2, early_late_space_chips=0.25,
                      very_early_late_space_chips=0.5, early_late_space_narrow_chips=0.15, very_early_late_space_narrow_chips=0.5,
                      pull_in_time_s=1, cn0_samples=20, cn0_min=25, max_code_lock_fail=50, max_carrier_lock_fail=5000, carrier_lock_th=0.5)
    p = e1_ref_params()
    p.track_pilot, p.secondary, p.secondary_code_length = 0, 0, 0
    p.secondary_code = b""
    cfg.slope, cfg.y_intercept = _e1_slope_intercept(0.25)
    loops = []
    for k in range(nsat):
        eng.set_code(k, sc["data_code"][k])
    loop = TrackingLoop(eng, cfg, nsat)
    for k in range(nsat):
        a = (int(sc["prn"][k]), float(sc["acq_delay"][k]), float(sc["doppler"][k] + sc["acq_err"][k]), 0, 0)
        loop.start(k, *a)
        loops.append(O.RefLoop(p, sc["data_code"][k], None, *a))
    got, ref = [[] for _ in range(nsat)], [[] for _ in range(nsat)]
    for _, x in e1_chunks(sc, E1_SECONDS):
        r = loop.step(x)
        for k in range(nsat):
            got[k].append(r[r["channel"] == k])
            ref[k].append(loops[k].feed(x))
    _compare_device_with_ref(loop, loops, got, ref, nsat, sc["doppler"], "E1B", 500)
    for lp in loops:
        lp.close()
    loop.close()
    eng.close()


def _e1_slope_intercept(spc):
    """slope / y_intercept the reference derives from SinBocCorrelationFunction<1,1> (dll_pll_veml_tracking.cc:275-277, tracking_discriminators.h):
    taken from the host library's configuration reader, which is bit-checked against the reference's in tests/test_host_cpu.py"""
    from gnss_sim_receiver_b200 import _hostapi as H
    lib = H.lib()
    return float(np.float32(-lib.gb200h_sinboc_slope_abs(float(spc)))), float(np.float32(lib.gb200h_sinboc_yintercept_abs(float(spc))))


L5_FS = 12.5e6
L5_SECONDS = float(os.environ.get("GB200_LOOP_SECONDS_L5", "8"))
# signal, secondary code (dll_pll_veml_tracking.cc:213-246 GPS_L5I_NH_CODE_STR; :296-330 GALILEO_E5A_I_SECONDARY_CODE), symbols per bit, d_interchange_iq
L5_BAND = {"L5": ("0000110101", 10, 1), "5X": ("10000100001011101001", 20, 0)}


def _l5_band_code(signal, prn):
    import _signals as S
    if signal == "L5":
        return S.l5_table(prn)
    return np.ascontiguousarray(S.e5a_primary(prn, "5I").real, dtype=np.float32)  # E5a-I primary code (the data component's table)


@pytest.mark.gpu
@needs_ref
@pytest.mark.parametrize("signal", ["L5", "5X"])
def test_device_loop_l5_band_vs_reference_objects(signal):
    """The device loop on GPS L5I and Galileo E5aI (data components, dll_pll_veml_tracking.cc:213-246, :296-330): 10 230-chip codes at
    10.23 Mcps (code step 0.82 entries per sample at 12.5 MS/s: the general correlator kernel), the secondary code (10-bit Neuman-Hofman /
    20-bit CS20) synchronised and removed, extended integration over the symbols of a bit, Prompt_I / Prompt_Q interchanged for L5I —
    against loops made of the reference's own objects."""
    from gnss_sim_receiver_b200.tracking import TrackingEngine, TrackingLoop, loop_config
    sec_str, spb, interchange = L5_BAND[signal]
    nsat, fs = 3, L5_FS
    chunk = int(fs * 0.1)
    rng = np.random.default_rng(31 + spb)
    prns = rng.choice(np.arange(1, 33), nsat, replace=False)
    doppler, delay, phase = rng.uniform(-3000, 3000, nsat), rng.uniform(0, 10230, nsat), rng.uniform(0, 2 * np.pi, nsat)
    cn0, acq_err = rng.uniform(44, 49, nsat), rng.uniform(-10, 10, nsat)
    codes = np.stack([_l5_band_code(signal, int(p)) for p in prns])
    sec = np.array([1 if c == "0" else -1 for c in sec_str], dtype=np.int8)
    n_bits = 12000 // spb
    bits = rng.choice(np.array([-1, 1], dtype=np.int8), (nsat, n_bits))
    sym = (np.repeat(bits, spb, axis=1) * np.tile(sec, n_bits)[None, :]).astype(np.int8)  # one entry per 1 ms code period
    f_chip = 10.23e6 * (1 + doppler / 1176.45e6)
    acq_delay = np.mod(delay / f_chip * fs, fs / 1000.0)
    kw = dict(pll_bw_hz=20.0, dll_bw_hz=1.5, pll_bw_narrow_hz=5.0, dll_bw_narrow_hz=0.5, early_late_space_chips=0.5, early_late_space_narrow_chips=0.25,
              pull_in_time_s=1, cn0_samples=20, cn0_min=25, max_code_lock_fail=50, max_carrier_lock_fail=5000, carrier_lock_th=0.5)
    cfg = loop_config(signal, fs, **kw)
    assert (cfg.secondary_code.decode(), cfg.symbols_per_bit, cfg.interchange_iq) == (sec_str, spb, interchange)
    p = ref_params(0)
    p.fs_in, p.code_period, p.chip_rate, p.carrier_freq = fs, 0.001, 10.23e6, 1176.45e6
    p.pll_bw, p.dll_bw, p.pll_bw_narrow, p.dll_bw_narrow, p.el, p.el_narrow = 20.0, 1.5, 5.0, 0.5, 0.5, 0.25
    p.vector_length, p.veml, p.secondary, p.track_pilot, p.symbols_per_bit, p.samples_per_chip, p.code_length_chips = int(fs / 1000), 0, 1, 0, spb, 1, 10230
    p.interchange_iq = interchange
    p.secondary_code_length, p.data_secondary_code_length = len(sec_str), 0
    p.secondary_code = sec_str.encode()
    eng = TrackingEngine(max_channels=nsat, max_code_len=10230, ring_samples=4 * chunk, max_block=2 * int(fs / 1000), max_cmds=nsat)
    for k in range(nsat):
        eng.set_code(k, codes[k])
    loop = TrackingLoop(eng, cfg, nsat)
    loops = []
    for k in range(nsat):
        a = (int(prns[k]), float(acq_delay[k]), float(doppler[k] + acq_err[k]), 0, 0)
        loop.start(k, *a)
        loops.append(O.RefLoop(p, codes[k], None, *a))
    got, ref = [[] for _ in range(nsat)], [[] for _ in range(nsat)]
    n_total = int(L5_SECONDS * fs)
    for first in range(0, n_total, chunk):
        n = min(chunk, n_total - first)
        x = O.gen_if(n, first, fs, codes, doppler, delay, phase, cn0, nav_bits=sym, periods_per_bit=1, f_chip=10.23e6, f_carrier=1176.45e6, seed=79)
        r = loop.step(x)
        for k in range(nsat):
            got[k].append(r[r["channel"] == k])
            ref[k].append(loops[k].feed(x))
    assert loop.kernel.startswith("trk_correlate_kernel"), loop.kernel
    _compare_device_with_ref(loop, loops, got, ref, nsat, doppler, signal, 250, data_field="prompt_q" if interchange else "prompt_i")
    for lp in loops:
        lp.close()
    loop.close()
    eng.close()


# ------------------------------------------------------------------------------------------------ tracking dump through the reference's reader
@pytest.mark.gpu
@needs_ref
def test_tracking_dump_is_read_by_the_reference_reader(tmp_path):
    """Tracking_XX.dump=true (dll_pll_veml_tracking.cc:1376-1470 log_data, :1675-1705 file naming): the per-channel .dat files the block
    writes are read back with the reference's OWN reader (src/tests/unit-tests/signal-processing-blocks/libs/tracking_dump_reader.cc,
    compiled into oracle/_ref) — whole epochs, and the fields agree with the Gnss_Synchro records of the same run."""
    from gnss_sim_receiver_b200 import _hostapi as H
    nsat, seconds = 3, 4.0
    sc = scenario(nsat, seed=19)
    base = str(tmp_path / "trk_dump_ch_")
    bank = H.TrackingBank(CONF + f"Tracking_1C.dump=true\nTracking_1C.dump_filename={base}\n", "Tracking_1C", nsat)
    for k in range(nsat):
        bank.start(k, int(sc["prn"][k]), float(sc["acq_delay"][k]), float(sc["doppler"][k] + sc["acq_err"][k]), 0, 0)
    got = [[] for _ in range(nsat)]
    for _, x in stream_chunks(sc, seconds):
        bank.push(x)
        r = bank.step()
        for k in range(nsat):
            got[k].append(r[r["channel"] == k])
    bank.close()  # closes the dump files
    rd = O.ref().ref_trk_dump_read
    rd.restype = O.C.c_long
    rd.argtypes = [O.C.c_char_p, O.C.c_void_p, O.C.c_long]
    for k in range(nsat):
        path = base + f"{k}.dat"
        assert os.path.exists(path), os.listdir(tmp_path)
        assert os.path.getsize(path) % (8 + 8 + 19 * 4 + 4) == 0  # whole epochs of the reference's record (tracking_dump_reader.cc:74-79)
        buf = np.zeros((8192, 10), dtype=np.float64)
        n = rd(path.encode(), buf.ctypes.data, len(buf))
        assert n > 1500, (k, n)
        d = buf[:n]
        g = np.concatenate(got[k])
        assert np.all(d[:, 9] == sc["prn"][k])
        # one dump epoch per code period once the loop runs (the first periods of pull-in included); Gnss_Synchro records are emitted for
        # the same periods from tracking on: match them on the PRN start sample stamp
        stamp = d[:, 5].astype(np.uint64)
        assert np.all(np.diff(stamp.astype(np.int64)) > 0)
        idx = np.searchsorted(stamp, g["tracking_sample_counter"])
        ok = (idx < n) & (stamp[np.minimum(idx, n - 1)] == g["tracking_sample_counter"])
        assert np.count_nonzero(ok) > 0.95 * len(g), (k, np.count_nonzero(ok), len(g))
        e, gg = d[idx[ok]], g[ok]
        # (the dump's prompt is that of the last code period, d_Prompt, :1397-1400; the record's is accumulated over the symbols of a bit,
        # :1952-1958 — not comparable; Doppler and CN0 are the same variables in both)
        assert np.allclose(e[:, 6], gg["carrier_doppler_hz"].astype(np.float32), rtol=1e-6, atol=1e-3)
        assert np.allclose(e[:, 7], gg["cn0_db_hz"].astype(np.float32), rtol=1e-6, atol=1e-4)
        assert np.all(d[-500:, 1] > d[-500:, 0]) and np.all(d[-500:, 1] > d[-500:, 2])  # locked: |P| above |E| and |L|
        assert np.all(np.hypot(d[-500:, 3], d[-500:, 4]) > 0.0)


# ------------------------------------------------------------------------------------------------ device loop, pilot tracking
def _conf_text(role, fs, **kw):
    lines = [f"GNSS-SDR.internal_fs_sps={int(fs)}", f"{role}.item_type=gr_complex"]
    for k, v in kw.items():
        lines.append(f"{role}.{k}={'true' if v is True else 'false' if v is False else v}")
    return "\n".join(lines) + "\n"


def _device_vs_block_pilot(label, loop, blocks, chunks, nsat, truth_doppler, min_records):
    got, ref = [[] for _ in range(nsat)], [[] for _ in range(nsat)]
    for x in chunks:
        r = loop.step(x)
        for k in range(nsat):
            got[k].append(r[r["channel"] == k])
            ref[k].append(blocks[k].feed(x))
    for k in range(nsat):
        g, r = np.concatenate(got[k]), np.concatenate(ref[k])
        gs, rs = loop.channel(k), blocks[k].state()
        assert rs["state"] == 4 and gs["state"] == 4, (label, k, gs, rs)
        m = min(len(g), len(r))
        assert m > min_records and len(r) - len(g) <= 6, (label, k, len(g), len(r))
        g, r = g[:m], r[:m]
        d_cnt = np.abs(g["tracking_sample_counter"].astype(np.int64) - r["tracking_sample_counter"].astype(np.int64))
        assert d_cnt.max() <= 1, (label, k, d_cnt.max())
        assert np.array_equal(g["correlation_length_ms"], r["correlation_length_ms"]), (label, k)
        # the symbols sit on the arm that carries the data component's energy (in phase with the pilot for E1, in quadrature for L5)
        f = "prompt_i" if np.median(np.abs(r["prompt_i"])) >= np.median(np.abs(r["prompt_q"])) else "prompt_q"
        assert np.array_equal(np.sign(g[f]), np.sign(r[f])), (label, k, f)
        amp = np.median(np.abs(r[f]))
        for ff in ("prompt_i", "prompt_q"):  # (auxiliary: both arms close in value; the gates are the signs above and CN0 / Doppler below)
            assert np.max(np.abs(g[ff][m // 4:] - r[ff][m // 4:])) < 0.1 * amp, (label, k, ff)
        half = m // 2
        d_cn0 = np.max(np.abs(g["cn0_db_hz"][half:] - r["cn0_db_hz"][half:]))
        d_dop = np.sqrt(np.mean((g["carrier_doppler_hz"][half:] - r["carrier_doppler_hz"][half:]) ** 2))
        print(f"[device loop {label}] ch {k}: records {m}, symbols on {f}, CN0 diff {d_cn0:.4f} dB, Doppler diff {d_dop:.4f} Hz rms")
        assert d_cn0 < 0.1 and d_dop < 0.5, (label, k, d_cn0, d_dop)
        assert abs(np.mean(g["carrier_doppler_hz"][half:]) - truth_doppler[k]) < 2.0, (label, k)


@pytest.mark.gpu
@needs_ref
def test_device_loop_galileo_e1_pilot_vs_the_reference_block():
    """track_pilot=true on the device: E1-C tracked by the five-tap VEML correlator (25-chip secondary code synchronised and wiped), the
    E1-B symbols from a second one-tap correlator in the same launch, no Costas loop — against the reference's OWN dll_pll_veml_tracking
    block (oracle/ref_trkblk.cc) on the E1 open-service stream."""
    from gnss_sim_receiver_b200.tracking import TrackingEngine, TrackingLoop, loop_config
    nsat = 4
    sc = e1_scenario(nsat, seed=37)
    kw = dict(pll_bw_hz=15.0, dll_bw_hz=0.75, pll_bw_narrow_hz=5.0, dll_bw_narrow_hz=0.2, early_late_space_chips=0.25, very_early_late_space_chips=0.5,
              early_late_space_narrow_chips=0.15, very_early_late_space_narrow_chips=0.5, extend_correlation_symbols=1, pull_in_time_s=1, cn0_samples=20, cn0_min=25,
              carrier_lock_th=0.5)
    eng = TrackingEngine(max_channels=2 * nsat, max_code_len=8184, ring_samples=4 * CHUNK, max_block=2 * 16000, max_cmds=2 * nsat)
    cfg = loop_config("1B", FS, track_pilot=True, max_code_lock_fail=50, max_carrier_lock_fail=5000, **kw)
    cfg.slope, cfg.y_intercept = _e1_slope_intercept(0.25)
    for k in range(nsat):
        eng.set_code(k, sc["pilot_code"][k])         # tracking code: slot i
        eng.set_code(nsat + k, sc["data_code"][k])   # data code: slot n_channels + i
    loop = TrackingLoop(eng, cfg, nsat)
    blocks = []
    for k in range(nsat):
        a = (int(sc["prn"][k]), float(sc["acq_delay"][k]), float(sc["doppler"][k] + sc["acq_err"][k]))
        loop.start(k, *a, 0, 0)
        blocks.append(O.RefTrkBlock(_conf_text("Tracking_1B", FS, track_pilot=True, max_lock_fail=50, max_carrier_lock_fail=5000, **kw), "Tracking_1B", "1B", *a))
    _device_vs_block_pilot("E1 pilot", loop, blocks, (x for _, x in e1_chunks(sc, E1_SECONDS)), nsat, sc["doppler"], 500)
    for b in blocks:
        b.close()
    loop.close()
    eng.close()


@pytest.mark.gpu
@needs_ref
@pytest.mark.parametrize("signal", ["L5", "5X"])
def test_device_loop_l5_band_pilot_vs_the_reference_block(signal):
    """track_pilot=true in the L5 band, against the reference's own block:
    GPS L5 — L5Q tracked (20-bit Neuman-Hofman code synchronised and wiped), the L5I symbols from the data correlator with their own 10-bit
    code removed (d_data_secondary_code_string, :227-237, :1305-1335);
    Galileo E5a — E5a-Q tracked with the 100-chip secondary code OF THE SATELLITE (gb200_loop_set_secondary_code; :733-735), the E5a-I
    symbols with CS20 removed, Prompt_I / Prompt_Q interchanged (:305-314)."""
    import _signals as S
    from gnss_sim_receiver_b200 import _hostapi as H
    from gnss_sim_receiver_b200.tracking import TrackingEngine, TrackingLoop, loop_config
    nsat, fs = 3, L5_FS
    chunk = int(fs * 0.1)
    rng = np.random.default_rng(53 if signal == "L5" else 59)
    prns = rng.choice(np.arange(1, 33), nsat, replace=False)
    doppler, delay, phase = rng.uniform(-3000, 3000, nsat), rng.uniform(0, 10230, nsat), rng.uniform(0, 2 * np.pi, nsat)
    cn0, acq_err = rng.uniform(45, 49, nsat), rng.uniform(-10, 10, nsat)
    signs = lambda code: np.array([1 if c == "0" else -1 for c in code], dtype=np.int8)
    if signal == "L5":
        ci = np.stack([S.l5_table(int(p)) for p in prns])
        cq = np.stack([S.l5_table(int(p), q=True) for p in prns])
        data_sec, spb, role = "0000110101", 10, "Tracking_L5"
        pilot_sec = ["00000100110101001110"] * nsat
    else:
        ci = np.stack([np.ascontiguousarray(S.e5a_primary(int(p), "5I").real, dtype=np.float32) for p in prns])
        cq = np.stack([np.ascontiguousarray(S.e5a_primary(int(p), "5Q").real, dtype=np.float32) for p in prns])
        data_sec, spb, role = "10000100001011101001", 20, "Tracking_5X"
        pilot_sec = []
        for p in prns:
            buf = C_create_string_buffer(101)
            assert H.lib().gb200h_e5a_q_secondary_code(int(p), buf) == 100
            pilot_sec.append(buf.value.decode())
    n_bits = 12000 // spb
    bits = rng.choice(np.array([-1, 1], dtype=np.int8), (nsat, n_bits))
    sym_i = (np.repeat(bits, spb, axis=1) * np.tile(signs(data_sec), n_bits)[None, :]).astype(np.int8)
    sym_q = np.stack([np.tile(signs(c), 12000 // len(c)) for c in pilot_sec]).astype(np.int8)
    two = lambda a: np.concatenate([a, a])
    gen_codes, gen_sym = np.concatenate([ci, cq]), np.concatenate([sym_i, sym_q])
    gen_phase = np.concatenate([phase, phase + np.pi / 2])  # the pilot in quadrature with the data component
    acq_delay = np.mod(delay / (10.23e6 * (1 + doppler / 1176.45e6)) * fs, fs / 1000.0)
    kw = dict(pll_bw_hz=20.0, dll_bw_hz=1.5, pll_bw_narrow_hz=5.0, dll_bw_narrow_hz=0.5, early_late_space_chips=0.5, early_late_space_narrow_chips=0.25,
              pull_in_time_s=1, cn0_samples=20, cn0_min=25, carrier_lock_th=0.5)
    eng = TrackingEngine(max_channels=2 * nsat, max_code_len=10230, ring_samples=4 * chunk, max_block=2 * int(fs / 1000), max_cmds=2 * nsat)
    cfg = loop_config(signal, fs, track_pilot=True, max_code_lock_fail=50, max_carrier_lock_fail=5000, **kw)
    for k in range(nsat):
        eng.set_code(k, cq[k])
        eng.set_code(nsat + k, ci[k])
    loop = TrackingLoop(eng, cfg, nsat)
    blocks = []
    for k in range(nsat):
        a = (int(prns[k]), float(acq_delay[k]), float(doppler[k] + acq_err[k]))
        if signal == "5X":
            loop.set_secondary_code(k, pilot_sec[k])
        loop.start(k, *a, 0, 0)
        blocks.append(O.RefTrkBlock(_conf_text(role, fs, track_pilot=True, max_lock_fail=50, max_carrier_lock_fail=5000, **kw), role, signal, *a))
    n_total = int(L5_SECONDS * fs)
    chunks = (O.gen_if(min(chunk, n_total - first), first, fs, gen_codes, two(doppler), two(delay), gen_phase, two(cn0) - 3.0103, nav_bits=gen_sym, periods_per_bit=1,
                       f_chip=10.23e6, f_carrier=1176.45e6, seed=89) for first in range(0, n_total, chunk))
    _device_vs_block_pilot(f"{signal} pilot", loop, blocks, chunks, nsat, doppler, 120 if signal == "5X" else 250)
    for b in blocks:
        b.close()
    loop.close()
    eng.close()


def C_create_string_buffer(n):
    import ctypes
    return ctypes.create_string_buffer(n)
