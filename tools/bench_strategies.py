#!/usr/bin/env python
"""Search latency of the SURVEY 8f-4 acquisition strategies: the B200 block (through its C door, host buffers in, results out — what a channel
sees) against the reference's OWN block compiled into oracle/_ref on one host core (the reference runs one block per channel thread).

One JSON line per strategy:  {"strategy", "workload", "gpu_ms_per_search", "ref_cpu_ms_per_search", "speedup", "same_result"}.
A search = arm + every general_work call until the block raises its event.  GPU time is wall clock around the whole search after two
warm-up searches (kernel launches, host<->device copies and the host-side glue included), median of `--reps`.

    python tools/bench_strategies.py [--reps 9] > profiles/<round>_strategies.jsonl
"""
import argparse
import json
import os
import statistics
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import _oracle as O  # noqa: E402  (test infrastructure: the referee / CPU baseline, never the thing measured)
import _signals as S  # noqa: E402
from gnss_sim_receiver_b200 import _hostapi as H  # noqa: E402


def timed(make, arm, work, status, close, reps, warm):
    a = make()
    ts = []
    st = None
    for k in range(warm + reps):
        t0 = time.perf_counter()
        arm(a)
        work(a)
        st = status(a)
        t1 = time.perf_counter()
        if k >= warm:
            ts.append((t1 - t0) * 1e3)
    close(a)
    return statistics.median(ts), st


def line(name, workload, gpu, ref, same):
    print(json.dumps({"strategy": name, "workload": workload, "gpu_ms_per_search": round(gpu, 3), "ref_cpu_ms_per_search": round(ref, 2), "speedup": round(ref / gpu, 1),
                      "same_result": bool(same)}), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=9)
    ap.add_argument("--ref-reps", type=int, default=2)
    args = ap.parse_args()
    if not O.ref_available():
        raise SystemExit("oracle/_ref is not built: the reference blocks are the baseline of this tool")

    # ---- fine Doppler (GPS L1 C/A): 2 dwells + the 80 x samples-per-ms point fine estimate
    for fs in (4e6, 25e6):
        n_ms = int(fs / 1000)
        prn, dop = 7, 1830.0
        x = S.synth_if(14 * n_ms, fs, [dict(table=S.gps_l1_table(prn), f_chip=1.023e6, doppler=dop, delay=311.0, cn0=50.0, phase=1.0)], seed=3)
        conf = (f"GNSS-SDR.internal_fs_sps={int(fs)}\nAcquisition_1C.implementation=GPS_L1_CA_PCPS_Acquisition_Fine_Doppler_B200\nAcquisition_1C.item_type=gr_complex\n"
                "Acquisition_1C.doppler_max=5000\nAcquisition_1C.max_dwells=2\n")
        g, sa = timed(lambda: H.FineDopplerAcquisitionBlock(conf, "Acquisition_1C", prn), lambda a: a.arm(5000, 250, 2.5), lambda a: a.work(x, n_ms), lambda a: a.status(),
                      lambda a: a.close(), args.reps, 2)
        r, sb = timed(lambda: O.RefAcqFd(fs, 5000, 2, prn), lambda b: b.arm(5000, 250, 2.5), lambda b: b.work(x, n_ms), lambda b: b.status(), lambda b: b.close(), args.ref_reps, 0)
        line("fine_doppler", f"GPS L1 C/A {fs / 1e6:g} MS/s, 40 bins x 2 dwells + fine estimate on {80 * n_ms} points", g, r,
             (sa.acq_delay_samples, sa.acq_doppler_hz) == (sb.acq_delay_samples, sb.acq_doppler_hz))

    # ---- Galileo E5a non-coherent I/Q with CAF: 3 ms, both components (4 replicas per bin), 41 bins
    from test_acq_e5a_caf_refblock_cpu import e5a_stream
    for fs, ms in ((12e6, 3), (20e6, 1)):
        prn = 11
        x = e5a_stream(prn, 1830.0, 3333.0, 5 * ms + 2, sym_i=(1, -1, -1, 1, 1, -1), sym_q=(1, 1, -1, -1, 1, -1), fs=fs)
        r_ = "Acquisition_5X"
        conf = (f"GNSS-SDR.internal_fs_sps={int(fs)}\nChannel.signal=5X\n{r_}.implementation=Galileo_E5a_Noncoherent_IQ_Acquisition_CAF_B200\n{r_}.item_type=gr_complex\n{r_}.doppler_max=5000\n"
                f"{r_}.coherent_integration_time_ms={ms}\n{r_}.max_dwells=1\n{r_}.CAF_window_hz=3000\n")
        chunk = int(fs / 1000)
        g, sa = timed(lambda: H.E5aCafAcquisitionBlock(conf, r_, prn, "5X"), lambda a: a.arm(5000, 250, 0.0), lambda a: a.work(x, chunk), lambda a: a.status(), lambda a: a.close(),
                      args.reps, 2)

        r, sb = timed(lambda: O.RefAcqCaf(fs, ms, 1, 5000, prn, caf_window_hz=3000), lambda b: b.arm(5000, 250, 0.0), lambda b: b.work(x, chunk), lambda b: b.status(), lambda b: b.close(), args.ref_reps, 0)
        line("e5a_noncoherent_iq_caf", f"Galileo E5a {fs / 1e6:g} MS/s, {ms} ms, I + Q components ({4 if ms > 1 else 2} replicas per bin), 41 bins, CAF 3 kHz", g, r,
             (sa.acq_delay_samples, sa.acq_doppler_hz) == (sb.acq_delay_samples, sb.acq_doppler_hz))

    # ---- QuickSync (GPS L1 C/A, folding factor 4; Galileo E1, folding factor 2)
    for fs in (4e6, 16e6):
        prn = 7
        n_ms = int(fs / 1000)
        x = S.synth_if(3 * 4 * n_ms, fs, [dict(table=S.gps_l1_table(prn), f_chip=1.023e6, doppler=600.0, delay=977.0, cn0=50.0, phase=1.0)], seed=3)
        conf = (f"GNSS-SDR.internal_fs_sps={int(fs)}\nAcquisition_1C.implementation=GPS_L1_CA_PCPS_QuickSync_Acquisition_B200\nAcquisition_1C.item_type=gr_complex\n"
                "Acquisition_1C.doppler_max=5000\nAcquisition_1C.coherent_integration_time_ms=4\nAcquisition_1C.folding_factor=4\n")
        g, sa = timed(lambda: H.QuickSyncAcquisitionBlock(conf, "Acquisition_1C", prn), lambda a: a.arm(5000, 250, 0.02), lambda a: a.work(x, 1), lambda a: a.status(), lambda a: a.close(),
                      args.reps, 2)
        r, sb = timed(lambda: O.RefAcqQs(fs, 4, 4, 1, 5000, prn), lambda b: b.arm(5000, 250, 0.02), lambda b: b.work(x, 1), lambda b: b.status(), lambda b: b.close(), args.ref_reps, 0)
        line("quicksync", f"GPS L1 C/A {fs / 1e6:g} MS/s, folding factor 4 (4 ms item -> {n_ms // 4}-point transforms), 41 bins", g, r,
             (sa.acq_delay_samples, sa.acq_doppler_hz) == (sb.acq_delay_samples, sb.acq_doppler_hz))

    # ---- two-step PCPS (coarse 80 bins + fine 4 bins): the reference's pcps_acquisition with make_two_steps
    fs, prn = 4e6, 7
    x = S.synth_if(8 * 4000, fs, [dict(table=S.gps_l1_table(prn), f_chip=1.023e6, doppler=835.0, delay=311.0, cn0=48.0, phase=1.0)], seed=13)
    conf = (f"GNSS-SDR.internal_fs_sps={int(fs)}\nAcquisition_1C.implementation=GPS_L1_CA_PCPS_Acquisition_B200\nAcquisition_1C.item_type=gr_complex\nAcquisition_1C.doppler_max=10000\n"
            "Acquisition_1C.doppler_step=250\nAcquisition_1C.pfa=0.01\nAcquisition_1C.max_dwells=1\nAcquisition_1C.make_two_steps=true\nAcquisition_1C.second_nbins=4\n"
            "Acquisition_1C.second_doppler_step=125\nAcquisition_1C.blocking=true\n")
    code = O.gps_l1_ca_code_sampled(prn, int(fs), 4000)
    g, sa = timed(lambda: H.AcquisitionBlock(conf, "Acquisition_1C", prn), lambda a: a.arm(), lambda a: a.work(x, chunk=1000), lambda a: a.status(), lambda a: a.close(), args.reps, 2)

    def arm_ref(b):
        b.init(10000, 250, 0.0)
        b.set_local_code(code)
        b.arm()
    r, sb = timed(lambda: O.RefAcq(conf.replace("_B200", ""), "Acquisition_1C", 1, 1.023e6, 2e6, prn, "G", "1C"), arm_ref, lambda b: b.work(x, 1, chunk=1000), lambda b: b.status(), lambda b: b.close(),
                  args.ref_reps, 0)
    line("pcps_two_steps", "GPS L1 C/A 4 MS/s, 80 coarse bins then 4 fine bins (125 Hz) on the next buffer", g, r,
         sa.last_event == 1 and (sa.acq_delay_samples, sa.acq_doppler_hz) == (sb.acq_delay_samples, sb.acq_doppler_hz))


if __name__ == "__main__":
    main()
