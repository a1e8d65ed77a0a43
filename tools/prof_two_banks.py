#!/usr/bin/env python
"""How much of the closed loop's critical path (loop kernel + launch gaps, DESIGN.md §8 "open structural item") hides when the channels of
one GPU run as B independent banks — B engines (each with its own stream and ring) and B device loops of channels / B channels, submitted
back to back so that one bank's loop kernel runs under another bank's correlator.

Same workload as bench_closed_loop.py (GPS L1 C/A, 25 MS/s cf32, 11 taps, 10240 channels on the GPU), pipelined form
(gb200_loop_submit / gb200_loop_collect, two steps in flight per bank), host buffers in, records out; every bank is pushed the period's
IQ block itself (B x the H2D bytes).  One JSON line per B:

    python tools/prof_two_banks.py [--banks 1,2,4] [--periods 1000] [--lock 400] [--channels 1024,4096]
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def run(n_banks, periods, lock, stream, nch=None):
    import torch
    import bench
    import bench_closed_loop as C
    from gnss_sim_receiver_b200 import _capi
    from gnss_sim_receiver_b200.tracking import TrackingEngine, TrackingLoop, loop_config
    lib = _capi.lib()
    nch = nch or bench.CHANNELS_PER_GPU
    per = nch // n_banks
    codes = bench.gps_codes()
    pinned, sc = stream
    host = pinned.numpy()
    cfg = loop_config("1C", C.FS, pull_in_time_s=1, cn0_samples=20, cn0_min=25, max_code_lock_fail=50, max_carrier_lock_fail=5000, carrier_lock_th=0.5,
                      early_late_space_chips=0.5, n_correlators=bench.N_TAPS)
    rng = np.random.default_rng(7)
    sat = np.arange(nch) % C.NSAT
    acq_delay = np.mod(sc["delay"] / 1.023e6 * C.FS, C.FS / 1000.0)
    engs, loops = [], []
    for b in range(n_banks):
        eng = TrackingEngine(device=0, max_channels=per, max_code_len=1023, ring_samples=16 * C.N_BLK, max_block=C.N_BLK + 8, max_cmds=per)
        for c in range(per):
            eng.set_code(c, codes[sat[b * per + c]])
        loop = TrackingLoop(eng, cfg, per)
        for c in range(per):
            g = b * per + c
            loop.start(c, int(sat[g]) + 1, float(acq_delay[sat[g]] + rng.uniform(-2, 2)), float(sc["doppler"][sat[g]] + rng.uniform(-40, 40)), 0, 0)
        engs.append(eng)
        loops.append(loop)

    def block(p):
        o = (p * C.N_BLK) % len(host)
        return host[o:o + C.N_BLK]

    def pipelined(p0, k):
        cmds = recs = 0
        for lp in loops:
            lp.submit(block(p0))
        for p in range(p0 + 1, p0 + k):
            for lp in loops:
                lp.submit(block(p))
            for lp in loops:
                recs += len(lp.collect(copy=False))
                cmds += lp.last_step["commands"]
        for lp in loops:
            recs += len(lp.collect(copy=False))
            cmds += lp.last_step["commands"]
        return cmds, recs

    pipelined(0, lock)
    torch.cuda.synchronize()
    l0 = lib.gb200_launch_count()
    t0 = time.perf_counter()
    cmds, recs = pipelined(lock, periods)
    torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    launches = lib.gb200_launch_count() - l0
    tracking = sum(1 for lp in loops for c in range(0, per, 64) if lp.channel(c)["state"] >= 2)
    for lp in loops:
        lp.close()
    for e in engs:
        e.close()
    return {"banks": n_banks, "channels_per_bank": per, "periods": periods, "ms_per_period": round(wall / periods * 1e3, 4),
            "e2e_channel_msps_per_s": round(cmds * C.N_BLK / 1e6 / wall, 0), "commands": int(cmds), "records": int(recs), "gpu_launches": int(launches),
            "sampled_channels_tracking": tracking, "h2d_bytes_per_period": n_banks * C.N_BLK * 8}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--banks", default="1,2,4")
    ap.add_argument("--periods", type=int, default=1000)
    ap.add_argument("--lock", type=int, default=400)
    ap.add_argument("--channels", default="", help="comma-separated channel counts on the GPU (default: the bench's 10240)")
    a = ap.parse_args()
    import bench
    import bench_closed_loop as C
    stream = C._stream(0, bench.gps_codes(), 100)
    for nch in [int(x) for x in a.channels.split(",")] if a.channels else [None]:
        for b in [int(x) for x in a.banks.split(",")]:
            print(json.dumps(dict(run(b, a.periods, a.lock, stream, nch), channels=nch or bench.CHANNELS_PER_GPU)), flush=True)


if __name__ == "__main__":
    main()
