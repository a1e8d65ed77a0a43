"""The four acquisition adapters selected from a `.conf` (gnss_block_factory.cc GetAcqBlock chain) and driven the way Channel /
ChannelFsm drive the reference's (set_doppler_*, set_threshold, init, set_local_code, set_state(1), general_work over the stream):

  Galileo_E1_PCPS_Ambiguous_Acquisition_B200   adapters/galileo_e1_pcps_ambiguous_acquisition.cc:36-199 (4 ms code, acquire_pilot, cboc)
  GPS_L5i_PCPS_Acquisition_B200                adapters/gps_l5i_pcps_acquisition.cc:36-171
  Galileo_E5a_Pcps_Acquisition_B200            adapters/galileo_e5a_pcps_acquisition.cc:36-180 (acquire_pilot -> 5Q, acquire_iq -> 5X)
  (GPS_L1_CA_PCPS_Acquisition_B200 is covered in tests/test_host_gpu.py)

Checked: the replica the adapter hands to set_local_code is the reference generator's (oracle/_ref, bit for bit) repeated as the
reference adapter repeats it; the block's detection (index, bin, Acq_delay_samples, Acq_doppler_hz, statistic) equals the oracle
restatement of acquisition_core on the same buffered block; the satellite is found where it was put.
"""
import numpy as np
import pytest

import _oracle as O
import _signals as S

pytestmark = pytest.mark.gpu


def _conf(fs, role, impl, extra=""):
    return f"""
GNSS-SDR.internal_fs_sps={int(fs)}
{role}.implementation={impl}
{role}.item_type=gr_complex
{role}.doppler_max=5000
{role}.doppler_step=250
{role}.blocking_on_standby=true
{extra}
"""


def _oracle(fs, consumed, fft_size, spc, spchip, local, sigs, bit_transition=False, use_cfar=True, max_dwells=1):
    conf = O.acq_conf(fs, consumed, fft_size, 5000, 250, spc, spchip, bit_transition=bit_transition, use_cfar=use_cfar, max_dwells=max_dwells)
    fc = O.acq_set_local_code(conf, local)
    grid = None
    for d, s in enumerate(sigs):
        grid = O.acq_dwell(conf, s, fc, grid, d)
    return conf, O.acq_statistic(conf, grid, len(sigs))


def _ref_block(conf_text, role, ms_per_code, chip_rate, opt, prn, system, signal, local, x, dmax=5000, dstep=250):
    """the same .conf, replica and stream through the reference's OWN pcps_acquisition (oracle/ref_acq.cc): consumed items and status"""
    if not O.ref_available():
        return None
    b = O.RefAcq(conf_text.replace("_B200", ""), role, ms_per_code, chip_rate, opt, prn, system, signal)
    b.init(dmax, dstep, 0.0)
    code = np.zeros(b.status().fft_size, np.complex64)
    code[:len(local)] = local
    b.set_local_code(code)
    b.arm()
    consumed = b.work(np.ascontiguousarray(x, np.complex64), 1)
    st = b.status()
    b.close()
    return consumed, st


def _check_vs_ref_block(s, consumed, ref):
    """the product's block against the reference's block: same items consumed (same number of dwells), same event, same Gnss_Synchro"""
    if ref is None:
        return
    rc, rs = ref
    assert consumed == rc, (consumed, rc)
    assert list(rs.fsm_events[:rs.n_fsm_events]) == [1] and s.last_event == 1
    assert (s.acq_delay_samples, s.acq_doppler_hz, s.acq_samplestamp_samples) == (rs.acq_delay_samples, rs.acq_doppler_hz, rs.acq_samplestamp_samples)
    assert abs(s.threshold - rs.threshold) <= 1e-5 * rs.threshold and abs(s.test_statistics - rs.test_statistics) <= 1e-3 * rs.test_statistics


def _check(s, ro, consumed):
    assert s.last_event == 1 and s.active == 0 and s.state == 0, (s.last_event, s.active, s.state, s.test_statistics, s.threshold)
    assert (s.indext, s.doppler_index) == (ro.indext, ro.doppler_index)
    assert s.acq_delay_samples == ro.acq_delay_samples and s.acq_doppler_hz == ro.acq_doppler_hz
    assert abs(s.test_statistics - ro.test_statistics) <= 1e-3 * ro.test_statistics


@pytest.mark.parametrize("pilot,cboc", [(False, False), (True, False), (False, True)])
def test_galileo_e1_adapter_from_conf(pilot, cboc):
    from gnss_sim_receiver_b200 import _hostapi as H
    fs, prn, doppler, delay = 4e6, 11, 1375.0, 2731.4  # delay in table entries (2 per chip)
    extra = "Acquisition_1B.coherent_integration_time_ms=4\nAcquisition_1B.pfa=0.001\n"
    extra += "Acquisition_1B.acquire_pilot=true\n" if pilot else ""
    extra += "Acquisition0.cboc=true\n" if cboc else ""   # read as "Acquisition<channel>.cboc" by the reference (:151-152)
    conf_text = _conf(fs, "Acquisition_1B", "Galileo_E1_PCPS_Ambiguous_Acquisition_B200", extra)
    a = H.AcquisitionBlock(conf_text, "Acquisition_1B", prn, system="E", signal="1B")
    a.arm()
    local = a.local_code()
    sig = "1C" if pilot else "1B"
    assert len(local) == 16000
    if O.ref_available():
        want = np.zeros(2 * 16000, np.float32)
        O.ref().ref_galileo_e1_code_gen_complex_sampled(want.ctypes.data_as(O.f32p), 16000, sig.encode(), int(cboc), prn, int(fs), 0, 0)
        assert np.array_equal(local.view(np.float32), want)
    base = dict(f_chip=2.046e6, doppler=doppler, delay=delay, cn0=47.0, phase=0.9)
    x = S.synth_if(3 * 16000, fs, [dict(table=S.e1_table(prn, "1B"), weight=np.sqrt(0.5), **base), dict(table=S.e1_table(prn, "1C"), weight=-np.sqrt(0.5), **base)], seed=61)
    consumed = a.work(x, chunk=3000)
    s = a.status()
    conf, ro = _oracle(fs, 16000, 16000, float(np.float32(np.float32(fs) * np.float32(0.001)) * 4), 4, local, [x[:16000]])
    _check(s, ro, consumed)
    assert consumed == 16000 and s.acq_samplestamp_samples == 16000
    _check_vs_ref_block(s, consumed, _ref_block(conf_text, "Acquisition_1B", 4, 1.023e6, 4e6, prn, "E", "1B", local, x))
    assert abs(s.threshold - O.acq_threshold(conf, 0.001)) <= 1e-5 * s.threshold
    assert abs(s.acq_delay_samples - delay / 2.046e6 * fs) <= 2 and abs(s.acq_doppler_hz - doppler) <= 125
    a.close()


def test_gps_l5i_adapter_from_conf_with_dwells():
    from gnss_sim_receiver_b200 import _hostapi as H
    fs, prn, doppler, delay = 20e6, 7, -1820.0, 6111.7
    extra = "Acquisition_L5.max_dwells=3\nAcquisition_L5.pfa=0.0001\n"
    conf_text = _conf(fs, "Acquisition_L5", "GPS_L5i_PCPS_Acquisition_B200", extra)
    a = H.AcquisitionBlock(conf_text, "Acquisition_L5", prn, system="G", signal="L5")
    a.arm()
    local = a.local_code()
    assert len(local) == 20000
    if O.ref_available():
        want = np.zeros(2 * 20000, np.float32)
        O.ref().ref_gps_l5i_code_gen_complex_sampled(want.ctypes.data_as(O.f32p), 20000, prn, int(fs))
        assert np.array_equal(local.view(np.float32), want)
    base = dict(f_chip=10.23e6, doppler=doppler, delay=delay, phase=2.2)
    # too weak for one 1 ms dwell at pfa 1e-4 over 1.6 M cells (statistic ~ 2 d (1 + SNR) after d dwells against a threshold of ~72),
    # found once the dwells accumulate (pcps_acquisition.cc:795-815)
    x = S.synth_if(5 * 20000, fs, [dict(table=S.l5_table(prn), cn0=46.0, weight=np.sqrt(0.5), **base), dict(table=1j * S.l5_table(prn, q=True), cn0=46.0, weight=np.sqrt(0.5), **base)],
                   seed=62, f_carrier=1176.45e6)
    consumed = a.work(x, chunk=20000)
    s = a.status()
    n_dwells = consumed // 20000
    assert 1 <= n_dwells <= 3
    conf, ro = _oracle(fs, 20000, 20000, float(np.float32(np.float32(fs) * np.float32(0.001))), 2, local, [x[d * 20000:(d + 1) * 20000] for d in range(n_dwells)], max_dwells=3)
    _check(s, ro, consumed)
    assert abs(((s.acq_delay_samples - delay / 10.23e6 * fs + 10000) % 20000) - 10000) <= 2 and abs(s.acq_doppler_hz - doppler) <= 250  # either bin next to the truth
    _check_vs_ref_block(s, consumed, _ref_block(conf_text, "Acquisition_L5", 1, 10.23e6, 10e6, prn, "G", "L5", local, x))  # the same number of dwells
    a.close()


@pytest.mark.parametrize("opts,sig", [("", "5I"), ("Acquisition_5X.acquire_pilot=true\n", "5Q"), ("Acquisition_5X.acquire_iq=true\nAcquisition_5X.acquire_pilot=true\n", "5X")])
def test_galileo_e5a_adapter_from_conf(opts, sig):
    from gnss_sim_receiver_b200 import _hostapi as H
    fs, prn, doppler, delay = 20e6, 19, 2630.0, 901.2
    conf_text = _conf(fs, "Acquisition_5X", "Galileo_E5a_Pcps_Acquisition_B200", "Acquisition_5X.pfa=0.001\n" + opts)
    a = H.AcquisitionBlock(conf_text, "Acquisition_5X", prn, system="E", signal="5X")
    a.arm()
    local = a.local_code()
    assert len(local) == 20000
    if O.ref_available():
        want = np.zeros(2 * 20000, np.float32)
        O.ref().ref_galileo_e5_a_code_gen_complex_sampled(want.ctypes.data_as(O.f32p), 20000, prn, sig.encode(), int(fs), 0)
        assert np.array_equal(local.view(np.float32), want)
    x = S.synth_if(2 * 20000, fs, [dict(table=S.e5a_primary(prn, "5X"), f_chip=10.23e6, doppler=doppler, delay=delay, cn0=53.0, phase=0.4, weight=np.sqrt(0.5))], seed=63, f_carrier=1176.45e6)
    consumed = a.work(x, chunk=5000)
    s = a.status()
    conf, ro = _oracle(fs, 20000, 20000, float(np.float32(np.float32(fs) * np.float32(0.001))), 2, local, [x[:20000]])
    _check(s, ro, consumed)
    assert abs(((s.acq_delay_samples - delay / 10.23e6 * fs + 10000) % 20000) - 10000) <= 2 and abs(s.acq_doppler_hz - doppler) <= 250  # either bin next to the truth
    _check_vs_ref_block(s, consumed, _ref_block(conf_text, "Acquisition_5X", 1, 10.23e6, 10e6, prn, "E", "5X", local, x))
    a.close()
