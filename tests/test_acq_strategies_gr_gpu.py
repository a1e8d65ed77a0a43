"""The gr::block faces of the SURVEY 8f-4 acquisition blocks (host/b200_gr_blocks.h: b200_fine_doppler_acquisition_gr, b200_e5a_caf_acquisition_gr,
b200_quicksync_acquisition_gr) — what the adapters' get_left_block() / get_right_block() hand to a GNU Radio flowgraph inside GNSS-SDR.  Outside it they
compile against host/gnuradio_stub/ and are driven by its one-block scheduler (forecast -> general_work -> consume_each, one output item per call).
Each face must publish the block's event on its "events" message port and leave the results the block leaves when it is driven directly.
"""
import numpy as np
import pytest

import _signals as S
from test_acq_e5a_caf_refblock_cpu import e5a_stream
from test_acq_quicksync_refblock_cpu import gps_stream

pytestmark = pytest.mark.gpu


def test_fine_doppler_face():
    from gnss_sim_receiver_b200 import _hostapi as H
    fs, prn = 4e6, 7
    conf = (f"GNSS-SDR.internal_fs_sps={int(fs)}\nAcquisition_1C.implementation=GPS_L1_CA_PCPS_Acquisition_Fine_Doppler_B200\nAcquisition_1C.item_type=gr_complex\n"
            "Acquisition_1C.doppler_max=5000\nAcquisition_1C.max_dwells=2\n")
    x = S.synth_if(16 * 4000, fs, [dict(table=S.gps_l1_table(prn), f_chip=1.023e6, doppler=1830.0, delay=311.0, cn0=50.0, phase=1.0)], seed=3)
    a = H.FineDopplerAcquisitionBlock(conf, "Acquisition_1C", prn)
    a.arm(5000, 250, 2.5)
    a.work(x, 4000)
    direct = a.status()
    a.close()
    a = H.FineDopplerAcquisitionBlock(conf, "Acquisition_1C", prn)
    a.arm(5000, 250, 2.5)
    done, ev = a.work_gr(x)
    s = a.status()
    a.close()
    assert ev == 1 == direct.last_event and s.last_event == 1
    # forecast asks for fft_size items; with one output item per call the 10 ms buffer fills item by item (:489-496), the positive state takes one more, and
    # the idle block then swallows fft_size items per call while forecast is met (:459-470; the stub scheduler feeds until the stream is used up)
    assert 40001 <= done <= len(x) and (done - 40001) % 4000 == 0
    assert (s.acq_delay_samples, s.acq_doppler_hz, s.acq_samplestamp_samples, s.fine_index) == (direct.acq_delay_samples, direct.acq_doppler_hz, direct.acq_samplestamp_samples, direct.fine_index)


def test_e5a_caf_face():
    from gnss_sim_receiver_b200 import _hostapi as H
    fs, prn, r = 12e6, 11, "Acquisition_5X"
    conf = (f"GNSS-SDR.internal_fs_sps={int(fs)}\nChannel.signal=5X\n{r}.implementation=Galileo_E5a_Noncoherent_IQ_Acquisition_CAF_B200\n{r}.item_type=gr_complex\n{r}.doppler_max=5000\n"
            f"{r}.coherent_integration_time_ms=1\n{r}.max_dwells=1\n{r}.CAF_window_hz=0\n")
    x = e5a_stream(prn, 1830.0, 3333.0, 6)
    a = H.E5aCafAcquisitionBlock(conf, r, prn, "5X")
    a.arm(5000, 250, 1e-4)
    a.work(x, 12000)
    direct = a.status()
    a.close()
    a = H.E5aCafAcquisitionBlock(conf, r, prn, "5X")
    a.arm(5000, 250, 1e-4)
    done, ev = a.work_gr(x)
    s = a.status()
    a.close()
    assert ev == 1 == direct.last_event and done == len(x)  # the scheduler offers everything that is left: the positive state takes it all (:722-727)
    assert (s.acq_delay_samples, s.acq_doppler_hz, s.acq_samplestamp_samples, s.test_statistics) == (direct.acq_delay_samples, direct.acq_doppler_hz, direct.acq_samplestamp_samples, direct.test_statistics)


def test_quicksync_face():
    from gnss_sim_receiver_b200 import _hostapi as H
    fs, prn = 4e6, 7
    conf = (f"GNSS-SDR.internal_fs_sps={int(fs)}\nAcquisition_1C.implementation=GPS_L1_CA_PCPS_QuickSync_Acquisition_B200\nAcquisition_1C.item_type=gr_complex\n"
            "Acquisition_1C.doppler_max=5000\nAcquisition_1C.coherent_integration_time_ms=4\nAcquisition_1C.folding_factor=4\n")
    x = gps_stream(prn, 600.0, 977.0, 64)
    a = H.QuickSyncAcquisitionBlock(conf, "Acquisition_1C", prn)
    a.arm(5000, 250, 0.02)
    a.work(x, 1)
    direct = a.status()
    a.close()
    a = H.QuickSyncAcquisitionBlock(conf, "Acquisition_1C", prn)
    a.arm(5000, 250, 0.02)
    items, ev = a.work_gr(x)
    s = a.status()
    a.close()
    assert ev == 1 == direct.last_event and items == len(x) // 16000  # vector items of 16000 samples: one searched, the rest taken by the positive state
    assert (s.acq_delay_samples, s.acq_doppler_hz, s.test_statistics) == (direct.acq_delay_samples, direct.acq_doppler_hz, direct.test_statistics)
