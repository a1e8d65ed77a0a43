"""Host C++ layer on the GPU: the acquisition adapter/block driven the way Channel + ChannelFsm drive the reference's
(set_doppler_*, set_threshold, init, set_local_code, set_state(1), then general_work over the stream), checked against the
oracle restatement of pcps_acquisition::acquisition_core on the same data-buffer contents; and the acq -> trk hand-over
(Acq_delay_samples / Acq_doppler_hz / Acq_samplestamp_samples into start_tracking, dll_pll_veml_tracking.cc:643-660,1757-1788).
"""
import os
import sys

import numpy as np
import pytest

import _oracle as O

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
from gen_if import quantise  # noqa: E402

pytestmark = pytest.mark.gpu
FS = 4e6

CONF = """
GNSS-SDR.internal_fs_sps=4000000
Acquisition_1C.implementation=GPS_L1_CA_PCPS_Acquisition_B200
Acquisition_1C.item_type={item}
Acquisition_1C.pfa={pfa}
Acquisition_1C.doppler_max=10000
Acquisition_1C.doppler_step=250
Acquisition_1C.max_dwells={dwells}
Acquisition_1C.blocking_on_standby=true
Tracking_1C.implementation=GPS_L1_CA_DLL_PLL_Tracking_B200
Tracking_1C.item_type=gr_complex
Tracking_1C.early_late_space_chips=0.5
Tracking_1C.pull_in_time_s=1
Tracking_1C.cn0_samples=20
Tracking_1C.cn0_min=25
Tracking_1C.max_lock_fail=50
Tracking_1C.max_carrier_lock_fail=5000
Tracking_1C.carrier_lock_th=0.5
"""


def _stream(n, prn, doppler, delay, cn0, seed=9):
    codes = np.stack([O.gps_l1_ca_code(prn)])
    return O.gen_if(n, 0, FS, codes, [doppler], [delay], [0.7], [cn0], seed=seed)


def _oracle(block, dwells=1, use_cfar=True):
    conf = O.acq_conf(FS, 4000, 4000, 10000, 250, float(np.float32(4e6) * np.float32(0.001)), 4, use_cfar=use_cfar, max_dwells=dwells)
    fc = O.acq_set_local_code(conf, O.gps_l1_ca_code_sampled(block["prn"], int(FS), 4000))
    grid = None
    r = None
    for d, s in enumerate(block["dwells"]):
        grid = O.acq_dwell(conf, s, fc, grid, d)
        r = O.acq_statistic(conf, grid, d + 1)
    return conf, r


@pytest.mark.parametrize("item", ["gr_complex", "cshort", "cbyte"])
def test_acquisition_block_positive_matches_oracle(item):
    from gnss_sim_receiver_b200 import _hostapi as H
    prn, doppler, delay = 10, 750.0, 600.0  # the reference's own synthetic acquisition case (gsoc2013 test :204-230)
    x = _stream(3 * 4000, prn, doppler, delay, 47.0)
    if item == "gr_complex":
        items, buf = x, x[:4000]
    else:
        q = quantise(x, "ishort" if item == "cshort" else "ibyte")
        items = q
        buf = (q[0:8000:2].astype(np.float32) + 1j * q[1:8000:2].astype(np.float32)).astype(np.complex64)
    a = H.AcquisitionBlock(CONF.format(item=item, pfa=0.01, dwells=1), "Acquisition_1C", prn)
    a.arm()
    consumed = a.work(items, chunk=1000)
    s = a.status()
    # the block buffered exactly the first consumed_samples items (state 1), then ran (state 2)
    conf, ro = _oracle(dict(prn=prn, dwells=[buf]))
    assert s.last_event == 1 and s.active == 0 and s.state == 0
    assert (s.indext, s.doppler_index) == (ro.indext, ro.doppler_index)
    assert s.acq_delay_samples == ro.acq_delay_samples and s.acq_doppler_hz == ro.acq_doppler_hz
    assert abs(s.test_statistics - ro.test_statistics) <= 1e-3 * ro.test_statistics
    assert s.acq_samplestamp_samples == 4000 and consumed == 4000
    assert abs(s.threshold - O.acq_threshold(conf, 0.01)) <= 1e-5 * s.threshold
    # truth: code start = delay / f_chip * fs samples, Doppler bin nearest to 750 Hz
    assert abs(s.acq_delay_samples - 600.0 / 1.023e6 * FS) <= 2 and s.acq_doppler_hz == 750.0
    a.close()


def test_acquisition_block_negative_after_max_dwells():
    from gnss_sim_receiver_b200 import _hostapi as H
    rng = np.random.default_rng(4)
    x = ((rng.standard_normal(5 * 4000) + 1j * rng.standard_normal(5 * 4000)) / np.sqrt(2)).astype(np.complex64)
    a = H.AcquisitionBlock(CONF.format(item="gr_complex", pfa=0.001, dwells=2), "Acquisition_1C", 5)
    a.arm()
    a.work(x, chunk=4000)
    s = a.status()
    conf, ro = _oracle(dict(prn=5, dwells=[x[:4000], x[4000:8000]]), dwells=2)
    assert s.last_event == 2 and s.active == 0 and s.state == 0  # negative acquisition after max_dwells (:803-818)
    assert (s.indext, s.doppler_index) == (ro.indext, ro.doppler_index)
    assert abs(s.test_statistics - ro.test_statistics) <= 1e-3 * ro.test_statistics
    assert s.test_statistics <= s.threshold
    a.close()


@pytest.mark.skipif(not O.ref_available(), reason="oracle/_ref not built")
def test_acquisition_to_tracking_handover():
    """Acquire on the first millisecond, start tracking from the Gnss_Synchro the block filled, reach narrow tracking."""
    from gnss_sim_receiver_b200 import _hostapi as H
    prn, doppler, delay, cn0 = 17, -2230.0, 311.4, 46.0
    bits = np.ones((1, 3000), dtype=np.int8)
    bits[0, ::2] = -1
    pre = np.array([1, -1, -1, -1, 1, -1, 1, 1], dtype=np.int8)
    for b in range(0, 3000, 50):
        bits[0, b:b + 8] = pre
    codes = np.stack([O.gps_l1_ca_code(prn)])
    n = int(5 * FS)  # pull-in ends 2 s after the acquisition stamp (integer seconds, strict '<'), bit sync within 1 s after
    x = O.gen_if(n, 0, FS, codes, [doppler], [delay], [1.1], [cn0], nav_bits=bits, seed=21)
    a = H.AcquisitionBlock(CONF.format(item="gr_complex", pfa=0.01, dwells=1), "Acquisition_1C", prn)
    a.arm()
    a.work(x[:12000], chunk=2000)
    s = a.status()
    assert s.last_event == 1 and s.acq_doppler_hz == -2250.0
    bank = H.TrackingBank(CONF.format(item="gr_complex", pfa=0.01, dwells=1), "Tracking_1C", 1)
    bank.start(0, prn, s.acq_delay_samples, s.acq_doppler_hz, s.acq_samplestamp_samples, s.acq_samplestamp_samples)
    recs = []
    for first in range(0, n, 400000):
        bank.push(x[first:first + 400000])
        recs.append(bank.step())
    st = bank.channel_state(0)
    r = np.concatenate(recs)
    assert st["state"] == 4, st
    assert abs(st["carrier_doppler_hz"] - doppler) < 5.0 and abs(st["cn0_db_hz"] - cn0) < 2.0
    assert len(r) > 50 and np.all(r["valid"] == 1) and np.all(r["prn"] == prn)
    bank.close()
    a.close()


def test_two_step_acquisition_refines_doppler():
    """make_two_steps (pcps_acquisition.cc:698-757,770-787): a positive coarse search arms a fine grid of second_nbins bins,
    second_doppler_step apart, centred on the coarse Doppler; it runs on the NEXT buffer with the second step's own threshold
    (pfa_second_step, second_nbins) and only then declares the acquisition."""
    from gnss_sim_receiver_b200 import _hostapi as H
    prn, doppler, delay = 7, 835.0, 311.0
    x = _stream(4 * 4000, prn, doppler, delay, 48.0, seed=13)
    conf_text = CONF.format(item="gr_complex", pfa=0.01, dwells=1) + "Acquisition_1C.make_two_steps=true\nAcquisition_1C.second_nbins=4\nAcquisition_1C.second_doppler_step=125\n"
    a = H.AcquisitionBlock(conf_text, "Acquisition_1C", prn)
    a.arm()
    consumed = a.work(x, chunk=1000)
    s = a.status()
    # oracle: coarse search on the first buffer, fine search on the second (the equivalent integer grid: -250 + c + 125 k)
    conf1, r1 = _oracle(dict(prn=prn, dwells=[x[:4000]]))
    assert r1.acq_doppler_hz == 750.0
    conf2 = O.acq_conf(FS, 4000, 4000, 250, 125, float(np.float32(4e6) * np.float32(0.001)), 4, doppler_center=int(r1.acq_doppler_hz))
    assert conf2.num_doppler_bins == 4
    fc = O.acq_set_local_code(conf2, O.gps_l1_ca_code_sampled(prn, int(FS), 4000))
    r2 = O.acq_statistic(conf2, O.acq_dwell(conf2, x[4000:8000], fc))
    assert s.last_event == 1 and s.active == 0 and s.state == 0
    assert consumed == 8000 and s.acq_samplestamp_samples == 8000
    assert (s.indext, s.doppler_index) == (r2.indext, r2.doppler_index)
    assert s.acq_doppler_hz == r2.acq_doppler_hz == 875.0          # nearest fine bin to 835 Hz
    assert s.acq_delay_samples == r2.acq_delay_samples
    # max_to_input_power_statistic estimates the input power in the first step only (:517-522 sit under `if (!d_step_two)`): the fine step's
    # statistic is the fine grid's maximum over the power the COARSE search left in d_input_power
    want_stat = r2.peak / r1.input_power
    assert abs(r2.test_statistics - want_stat) > 0.05 * want_stat  # (the two laws differ on this stream: the check below can tell them apart)
    assert abs(s.test_statistics - want_stat) <= 1e-4 * want_stat and abs(s.input_power - r1.input_power) <= 1e-5 * r1.input_power
    # the fine step was judged against its own threshold (4 bins, pfa_second_step); once it has declared the acquisition the block
    # is back on the coarse step's threshold (calculate_threshold after d_step_two = false, :775-786)
    assert s.test_statistics > O.acq_threshold(conf2, 0.01)
    assert abs(s.threshold - O.acq_threshold(conf1, 0.01)) <= 1e-5 * s.threshold
    a.close()
    # ... and the reference's OWN block (pcps_acquisition.cc compiled unmodified into oracle/_ref, oracle/ref_acq.cc) on the same .conf, replica
    # and stream: it takes the same two buffers, raises the same event and reports the same fine-step result
    if O.ref_available():
        b = O.RefAcq(conf_text.replace("_B200", ""), "Acquisition_1C", 1, 1.023e6, 2e6, prn, "G", "1C")
        b.init(10000, 250, 0.0)
        b.set_local_code(O.gps_l1_ca_code_sampled(prn, int(FS), 4000))
        b.arm()
        rc = b.work(x, 1, chunk=1000)  # the same scheduler calls as a.work(x, chunk=1000)
        rs = b.status()
        b.close()
        assert rc == consumed and list(rs.fsm_events[:rs.n_fsm_events]) == [1]
        assert (rs.acq_delay_samples, rs.acq_doppler_hz, rs.acq_samplestamp_samples) == (s.acq_delay_samples, s.acq_doppler_hz, s.acq_samplestamp_samples)
        assert abs(rs.test_statistics - s.test_statistics) <= 1e-5 * rs.test_statistics and abs(rs.threshold - s.threshold) <= 1e-5 * rs.threshold
        assert abs(rs.input_power - s.input_power) <= 1e-5 * rs.input_power


@pytest.mark.skipif(not O.ref_available(), reason="oracle/_ref not built")
def test_channel_manager_acquires_tracks_loses_and_reacquires():
    """the callers around the two engines.  Three satellites in the air, three channels, all 32 PRNs searched in
    one sweep per buffered block.  One satellite is switched off for 2.4 s: its channel loses lock (ChannelFsm state 2 -> 0),
    the sweeps resume for the idle channel, and the satellite is re-acquired when it comes back."""
    from gnss_sim_receiver_b200 import _hostapi as H
    prns = [4, 13, 22]
    dop = [1300.0, -2650.0, 420.0]
    dly = [100.7, 640.2, 901.5]
    cn0 = [47.0, 45.0, 46.0]
    codes = np.stack([O.gps_l1_ca_code(p) for p in prns])
    bits = np.random.default_rng(2).choice(np.array([-1, 1], dtype=np.int8), (3, 3000))
    pre = np.array([1, -1, -1, -1, 1, -1, 1, 1], dtype=np.int8)
    for b in range(0, 3000, 50):
        bits[:, b:b + 8] = pre
    # pfa: thousands of sweeps x 32 satellites must not raise a false alarm.  Lock test every 20 symbols x 20 ms = 0.4 s on a CN0
    # smoothed with the default alpha = 0.002 would coast through the outage: a fast smoother and no tolerance make the loss prompt
    conf_text = (CONF.format(item="gr_complex", pfa=1e-7, dwells=1).replace("max_lock_fail=50", "max_lock_fail=0").replace("cn0_min=25", "cn0_min=30")
                 + "Tracking_1C.cn0_smoother_alpha=0.5\n")
    rx = H.ChannelManager(conf_text, "Acquisition_1C", "Tracking_1C", 3, list(range(1, 33)))
    chunk = 400000
    off_from, off_to, total = int(4.0 * FS), int(6.4 * FS), int(10.4 * FS)  # pull-in 2 s + bit sync <= 1.2 s before and after
    recs = []
    history = []
    for first in range(0, total, chunk):
        gone = off_from <= first < off_to  # PRN 22 off the air for these chunks (same seed: the noise is unchanged)
        k = 2 if gone else 3
        x = O.gen_if(chunk, first, FS, codes[:k], dop[:k], dly[:k], [0.3, 1.1, 2.0][:k], cn0[:k], nav_bits=bits[:k], seed=31)
        recs.append(rx.push(x))
        history.append([rx.channel_info(c) for c in range(3)] + [rx.sweeps])
    r = np.concatenate(recs)
    info = [rx.channel_info(c) for c in range(3)]
    by_prn = {int(i["prn"]): i for i in info if i["state"] == 2}
    # all three satellites end up tracked, each on one channel, nothing else was declared
    assert sorted(by_prn) == prns, info
    assert sum(i["acquisitions"] for i in info) == 4 and sum(i["losses"] for i in info) == 1, info
    for p, f in zip(prns, dop):
        assert abs(by_prn[p]["acq_doppler_hz"] - f) <= 125.0 and by_prn[p]["trk_state"] >= 2
        assert abs(by_prn[p]["cn0_db_hz"] - cn0[prns.index(p)]) < 3.0
    # the two satellites that never went away were acquired by the first sweep and never lost
    first_sweep = history[0]
    assert sorted(int(i["prn"]) for i in first_sweep[:3] if i["state"] == 2) == prns and first_sweep[3] <= 3
    assert by_prn[22]["acquisitions"] == 2 and by_prn[22]["losses"] == 1 and by_prn[4]["losses"] == 0 and by_prn[13]["losses"] == 0
    assert by_prn[22]["acq_samplestamp_samples"] >= off_to
    # no idle channel, no sweep: the count stands still until the loss of lock, then one sweep per buffered millisecond
    assert history[int(3.9 * FS) // chunk][3] == first_sweep[3]
    # PRN 22's records stop while it is off the air and resume after the re-acquisition
    t22 = r["tracking_sample_counter"][(r["prn"] == 22) & (r["valid"] == 1)].astype(np.int64)
    assert np.any(t22 < off_from) and np.any(t22 > off_to + int(2.0 * FS))
    assert np.max(np.diff(t22)) > int(2.0 * FS)  # lost inside the outage, re-acquired after it, 2 s of pull-in
    assert 100 < rx.sweeps < 2500
    rx.close()


def test_acquisition_dump_file(tmp_path):
    """Acquisition_XX.dump=true: the .mat written after a positive acquisition (pcps_acquisition.cc:393-493) carries the search
    grid and the scalars under the reference's variable names; here for a two-step search (coarse + narrow grid)."""
    from scipy.io import loadmat
    from gnss_sim_receiver_b200 import _hostapi as H
    prn, doppler, delay = 7, 835.0, 311.0
    x = _stream(4 * 4000, prn, doppler, delay, 48.0, seed=13)
    conf_text = (CONF.format(item="gr_complex", pfa=0.01, dwells=1) + "Acquisition_1C.make_two_steps=true\nAcquisition_1C.second_nbins=4\n"
                 "Acquisition_1C.second_doppler_step=125\nAcquisition_1C.dump=true\nAcquisition_1C.dump_filename=" + str(tmp_path / "acq_dump.mat") + "\n")
    a = H.AcquisitionBlock(conf_text, "Acquisition_1C", prn)
    a.arm()
    a.work(x, chunk=1000)
    s = a.status()
    assert s.last_event == 1
    files = sorted(os.listdir(tmp_path))
    assert files == ["acq_dump_G_1C_ch_0_1_sat_7.mat"], files
    m = loadmat(str(tmp_path / files[0]))
    conf1 = O.acq_conf(FS, 4000, 4000, 10000, 250, float(np.float32(4e6) * np.float32(0.001)), 4)
    fc = O.acq_set_local_code(conf1, O.gps_l1_ca_code_sampled(prn, int(FS), 4000))
    g1 = O.acq_dwell(conf1, x[:4000], fc)
    conf2 = O.acq_conf(FS, 4000, 4000, 250, 125, float(np.float32(4e6) * np.float32(0.001)), 4, doppler_center=750)
    g2 = O.acq_dwell(conf2, x[4000:8000], fc)
    assert m["acq_grid"].shape == (4000, 80) and m["acq_grid_narrow"].shape == (4000, 4)
    assert np.max(np.abs(m["acq_grid"].T - g1)) <= 2e-5 * g1.max()
    assert np.max(np.abs(m["acq_grid_narrow"].T - g2)) <= 2e-5 * g2.max()
    assert int(m["doppler_max"][0, 0]) == 10000 and int(m["doppler_step"][0, 0]) == 250 and int(m["d_positive_acq"][0, 0]) == 1
    assert int(m["PRN"][0, 0]) == prn and int(m["sample_counter"][0, 0]) == 8000 and int(m["num_dwells"][0, 0]) == 1
    assert float(m["acq_doppler_hz"][0, 0]) == 875.0 and float(m["acq_delay_samples"][0, 0]) == s.acq_delay_samples
    assert float(m["doppler_step_narrow"][0, 0]) == 125.0 and float(m["doppler_grid_narrow_min"][0, 0]) == 500.0
    assert abs(float(m["test_statistic"][0, 0]) - s.test_statistics) <= 1e-6 * s.test_statistics
    a.close()


def test_two_bit_packed_stream_tracks_like_its_unpacked_bytes():
    """Two_Bit_Packed_File_Signal_Source shape (sample_type=iq): the packed bytes are unpacked on the device into an ibyte ring;
    the tracking records are identical to those of a bank fed the host-unpacked ibyte stream (same ring contents)."""
    from gnss_sim_receiver_b200 import _hostapi as H
    from gnss_sim_receiver_b200 import tracking as T
    prn, doppler, delay, cn0 = 9, 1480.0, 1203.7, 48.0
    codes = np.stack([O.gps_l1_ca_code(prn)])
    n = int(5 * FS)
    bits = np.ones((1, 3000), dtype=np.int8)
    bits[0, ::2] = -1
    for b in range(0, 3000, 50):
        bits[0, b:b + 8] = np.array([1, -1, -1, -1, 1, -1, 1, 1], dtype=np.int8)
    x = O.gen_if(n, 0, FS, codes, [doppler], [delay], [0.4], [cn0], nav_bits=bits, seed=33)
    # 2-bit quantiser: sign + magnitude against one sigma; field value s in {-2,-1,0,1} -> level 2 s + 1
    sigma = np.std(x.real)
    def field(v):
        lvl = np.where(np.abs(v) > sigma, 3, 1) * np.where(v < 0, -1, 1)
        return (((lvl - 1) // 2) & 3).astype(np.uint8)
    fi, fq = field(x.real), field(x.imag)
    packed = (fi[0::2] | (fq[0::2] << 2) | (fi[1::2] << 4) | (fq[1::2] << 6)).astype(np.uint8)
    unpacked = O.unpack_2bit(packed)  # int8 I,Q,I,Q...
    conf = CONF.format(item="gr_complex", pfa=0.01, dwells=1).replace("Tracking_1C.item_type=gr_complex", "Tracking_1C.item_type=ibyte")
    # the bank's TrackingMonitor (gnss_flowgraph.cc:241-262) streams every record as a protobuf datagram
    import socket
    from test_monitor_cpu import observables_class
    listener = socket.socket(socket.AF_INET, socket.SOCK_DGRAM)
    listener.bind(("127.0.0.1", 0))
    listener.settimeout(2.0)
    conf += f"TrackingMonitor.enable_monitor=true\nTrackingMonitor.udp_port={listener.getsockname()[1]}\n"
    delay_samples = delay / 1.023e6 * FS  # code start, like the acquisition tests above
    outs = []
    for packed_input in (False, True):
        bank = H.TrackingBank(conf, "Tracking_1C", 1)
        bank.start(0, prn, delay_samples, doppler + 30.0, 0, 0)
        recs = []
        for first in range(0, n, 400000):
            m = min(400000, n - first)
            if packed_input:
                assert bank.push_raw(packed[first // 2:(first + m) // 2], T.SRC_2BIT, T.TWOBIT_IQ) == m
            else:
                bank.push(unpacked[2 * first:2 * (first + m)])
            recs.append(bank.step())
        outs.append((np.concatenate(recs), bank.channel_state(0)))
        assert bank.monitor_datagrams == len(outs[-1][0])
        bank.close()
    (ra, sa), (rb, sb) = outs
    first = observables_class()()
    first.ParseFromString(listener.recv(4096))
    o = first.observable[0]
    assert (o.system, o.signal, o.prn, o.tracking_sample_counter) == ("G", "1C", prn, int(ra["tracking_sample_counter"][0]))
    assert o.prompt_i == ra["prompt_i"][0] and o.cn0_db_hz == ra["cn0_db_hz"][0] and o.fs == int(FS)
    listener.close()
    assert len(ra) > 50 and sa["state"] == 4 and abs(sa["carrier_doppler_hz"] - doppler) < 5.0, sa
    assert ra.tobytes() == rb.tobytes()
    assert sa == sb


MIXED = """
GNSS-SDR.internal_fs_sps=4000000
Tracking_1C.implementation=GPS_L1_CA_DLL_PLL_Tracking_B200
Tracking_1C.item_type=gr_complex
Tracking_1B.implementation=Galileo_E1_DLL_PLL_VEML_Tracking_B200
Tracking_1B.item_type=gr_complex
Tracking_1B.track_pilot=true
Tracking_L5.implementation=GPS_L5_DLL_PLL_Tracking_B200
Tracking_L5.item_type=gr_complex
Tracking_5X.implementation=Galileo_E5a_DLL_PLL_Tracking_B200
Tracking_5X.item_type=gr_complex
Tracking_5X.stream_id=7
"""


def test_engine_is_per_stream_and_sized_for_the_band():
    """the per-GPU engine is keyed on (device, item type, input stream) and sized for the band's longest code
    period, so (a) a GPS L1 channel that starts first no longer locks a Galileo E1 channel (4 x longer windows) out of the shared
    L1-band engine, and (b) an L5 / E5a channel never shares the L1 band's ring; a failed attach is an error, not a silent ch = -1."""
    from gnss_sim_receiver_b200 import _hostapi as H
    probe = lambda ra, pa, rb, pb: H.kv(H.lib().gb200h_engine_probe(MIXED.encode(), ra.encode(), pa, rb.encode(), pb))
    # (a) L1 first, then E1: one engine (same stream), both attached, windows of 2 x 16000 samples fit
    r = probe("Tracking_1C", 3, "Tracking_1B", 11)
    assert "error" not in r, r
    assert int(r["same"]) == 1 and int(r["stream_a"]) == int(r["stream_b"]) == 0
    assert int(r["slot_a"]) >= 0 and int(r["slot_b"]) >= 0 and int(r["slot_a"]) != int(r["slot_b"])
    assert int(r["max_block"]) >= 2 * 16000
    # (b) L1 band and L5 band: two engines, two rings
    r = probe("Tracking_1C", 3, "Tracking_L5", 5)
    assert "error" not in r, r
    assert int(r["same"]) == 0 and int(r["stream_a"]) == 0 and int(r["stream_b"]) == 1
    # an explicit <role>.stream_id wins over the band default
    r = probe("Tracking_L5", 5, "Tracking_5X", 9)
    assert "error" not in r, r
    assert int(r["same"]) == 0 and int(r["stream_b"]) == 7


def test_gr_block_faces_forward_to_the_blocks():
    """INTEGRATION.md §4 as code (host/b200_gr_blocks.{h,cc}, compiled here against the GNU Radio stub): the gr::block wrappers obey the
    reference's scheduler contract — forecast 2 * vector_length, one Gnss_Synchro per general_work, consume_each — and its message ports:
    acquisition publishes 1 on "events"; tracking gives the records the lock-step bank gives, publishes 3 on loss of lock, and a
    "telemetry_to_trk" event 1 forces that loss (dll_pll_veml_tracking.cc:617-641)."""
    from gnss_sim_receiver_b200 import _hostapi as H
    prn, doppler, delay, cn0 = 9, 1430.0, 455.2, 47.0
    bits = np.ones((1, 3000), dtype=np.int8)
    bits[0, ::2] = -1
    pre = np.array([1, -1, -1, -1, 1, -1, 1, 1], dtype=np.int8)
    for b in range(0, 3000, 50):
        bits[0, b:b + 8] = pre
    codes = np.stack([O.gps_l1_ca_code(prn)])
    n = int(3.5 * FS)
    x = O.gen_if(n, 0, FS, codes, [doppler], [delay], [0.4], [cn0], nav_bits=bits, seed=33)
    conf = CONF.format(item="gr_complex", pfa=0.01, dwells=1)
    # acquisition through the wrapper: same detection as through the block, event 1 on the port
    a = H.AcquisitionBlock(conf, "Acquisition_1C", prn)
    a.arm()
    consumed, ev = a.work_gr(x[:12000])
    s = a.status()
    assert ev == 1 and consumed == 4000 and s.acq_doppler_hz == 1500.0
    b = H.AcquisitionBlock(conf, "Acquisition_1C", prn)
    b.arm()
    b.work(x[:12000], chunk=2000)
    sb = b.status()
    assert (s.indext, s.doppler_index, s.test_statistics) == (sb.indext, sb.doppler_index, sb.test_statistics)
    a.close()
    b.close()
    # tracking through the wrapper against the lock-step bank
    recs, ev = H.gr_tracking_run(conf, "Tracking_1C", prn, s.acq_delay_samples, s.acq_doppler_hz, x)
    bank = H.TrackingBank(conf, "Tracking_1C", 1)
    bank.start(0, prn, s.acq_delay_samples, s.acq_doppler_hz, 0, 0)
    ref = []
    for first in range(0, n, 400000):
        bank.push(x[first:first + 400000])
        ref.append(bank.step())
    ref = np.concatenate(ref)
    assert ev["state"] == 4 and ev["n_events"] == 0 and len(recs) > 20
    m = min(len(recs), len(ref))
    assert abs(len(recs) - len(ref)) <= 2
    assert np.array_equal(recs["tracking_sample_counter"][:m], ref["tracking_sample_counter"][:m])
    assert np.array_equal(np.sign(recs["prompt_i"][:m]), np.sign(ref["prompt_i"][:m]))
    assert np.max(np.abs(recs["carrier_doppler_hz"][:m] - ref["carrier_doppler_hz"][:m])) < 1e-6
    bank.close()
    # a telemetry fault posted on "telemetry_to_trk" forces the loss of lock: event 3 on "events", block back in standby
    recs2, ev2 = H.gr_tracking_run(conf, "Tracking_1C", prn, s.acq_delay_samples, s.acq_doppler_hz, x, telemetry_fault_at=int(3.0 * FS))
    assert ev2["n_events"] == 1 and ev2["last_event"] == 3 and ev2["state"] == 0
    assert np.sum(recs2["valid"] == 0) == 1 and len(recs2) < len(recs)
