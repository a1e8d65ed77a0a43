"""bench.py's closed-loop leg: the device-resident DLL/PLL loop (gb200_loop_*) over CHANNELS_PER_GPU channels at the north_star
operating point (GPS L1 C/A, 25 MS/s gr_complex, one 25000-sample code period per step), >= 1000 consecutive periods.

Every step is what a receiver does each millisecond: hand the new IQ block (HOST memory) to the engine, correlate every channel's
next window with the commands the loop formed on the device, close every loop (discriminators, filters, NCO update, lock / CN0,
bit synchronisation), get the emitted Gnss_Synchro records back.  `e2e` = wall time of that call; `device` = CUDA-event time of its
kernels.  Parity: a subset of the channels is also tracked by the product's HOST loop (B200TrackingBank, whose loop closure is
tested bit-exact / within 0.1 dB against the reference's own objects) on the same stream, and lock state, CN0 and Doppler compared.

The stream is synthetic and cyclic (1 s of signal: 8 satellites with integer-Hz Doppler, 50 bit/s data carrying the telemetry
preamble, replayed seamlessly), generated with torch on the device and copied to pinned host memory before anything is timed.
"""
import time

import numpy as np

FS, N_BLK, NSAT = 25e6, 25000, 8


def _stream(local, codes, seed):
    """1 s of cyclic signal (25e6 samples) in pinned host memory + the scenario"""
    import torch
    dev = f"cuda:{local}"
    rng = np.random.default_rng(seed)
    dopp = rng.integers(-4000, 4001, NSAT).astype(np.float64)  # integer Hz: the carrier is periodic in 1 s
    delay = rng.uniform(0, 1023, NSAT)
    ph0 = rng.uniform(0, 2 * np.pi, NSAT)
    cn0 = rng.uniform(43, 49, NSAT)
    bits = rng.choice(np.array([-1.0, 1.0]), (NSAT, 50))
    pre = np.array([1, -1, -1, -1, 1, -1, 1, 1], dtype=np.float64)
    for b in (0, 25):
        bits[:, b:b + 8] = pre
    n_total = int(FS)
    out = torch.empty(n_total, dtype=torch.complex64).pin_memory()
    g = torch.Generator(device=dev)
    g.manual_seed(seed)
    tabs = torch.from_numpy(codes[:NSAT].astype(np.float64)).to(dev)
    bits_d = torch.from_numpy(bits).to(dev)
    blk = 1 << 22
    for first in range(0, n_total, blk):
        n = min(blk, n_total - first)
        i = torch.arange(first, first + n, device=dev, dtype=torch.float64)
        t = i / FS
        x = torch.randn(n, 2, device=dev, dtype=torch.float64, generator=g) / np.sqrt(2.0)
        x = torch.complex(x[:, 0], x[:, 1])
        for k in range(NSAT):
            chips = t * 1.023e6 - delay[k]  # no code Doppler: 1000 whole code periods per second, the stream wraps seamlessly
            ci = torch.floor(chips).to(torch.int64)
            c = tabs[k][torch.remainder(ci, 1023)]
            b = bits_d[k][torch.remainder(torch.div(ci, 1023 * 20, rounding_mode="floor"), 50)]
            ph = 2 * np.pi * torch.remainder(dopp[k] * t, 1.0) + ph0[k]
            amp = np.sqrt(10.0 ** (cn0[k] / 10.0) / FS)
            x = x + amp * c * b * torch.complex(torch.cos(ph), torch.sin(ph))
        out[first:first + n] = x.to(torch.complex64).cpu()
    torch.cuda.synchronize()
    return out, dict(doppler=dopp, delay=delay, cn0=cn0)


def run_two_banks(lib, codes, sat, acq, cfg, block, local, rank, dist, periods_lock, K, sub, one_bank_recs, n_banks=2):
    """The same channels as `n_banks` independent banks on the GPU (an engine with its own stream and ring + a device loop each, channels / n_banks
    channels per bank), submitted back to back: one bank's loop kernel (a dependent chain on two warps per SM) runs under the other bank's
    correlator instead of on the critical path.  Every bank is pushed the period's IQ block itself (n_banks x the H2D bytes).  The records of
    the sampled channels over the lock-in must be IDENTICAL to the one-bank run's: same kernels, same hand-over, same stream."""
    import torch
    import bench
    from gnss_sim_receiver_b200.tracking import TrackingEngine, TrackingLoop
    nch = len(sat)
    per = nch // n_banks
    engs, loops = [], []
    for b in range(n_banks):
        e = TrackingEngine(device=local, max_channels=per, max_code_len=1023, ring_samples=16 * N_BLK, max_block=N_BLK + 8, max_cmds=per)
        for c in range(per):
            e.set_code(c, codes[sat[b * per + c]])
        lp = TrackingLoop(e, cfg, per)
        for c in range(per):
            g = b * per + c
            lp.start(c, int(sat[g]) + 1, acq[g][0], acq[g][1], 0, 0)
        engs.append(e)
        loops.append(lp)
    local_sub = [[c - b * per for c in sub if b * per <= c < (b + 1) * per] for b in range(n_banks)]
    kept = []

    def pipelined(p0, k, keep):
        cmds = recs = 0
        for lp in loops:
            lp.submit(block(p0))
        for p in range(p0 + 1, p0 + k + 1):
            if p < p0 + k:
                for lp in loops:
                    lp.submit(block(p))
            for b, lp in enumerate(loops):
                r = lp.collect(copy=False)
                recs += len(r)
                cmds += lp.last_step["commands"]
                if keep and len(r):
                    m = r[np.isin(r["channel"], local_sub[b])].copy()
                    m["channel"] += b * per
                    kept.append(m)
        return cmds, recs

    pipelined(0, periods_lock, rank == 0)
    torch.cuda.synchronize()
    bench.barrier(dist)
    l0 = lib.gb200_launch_count()
    t0 = time.perf_counter()
    cmds, recs = pipelined(periods_lock, K, False)
    torch.cuda.synchronize()
    wall = bench.reduce_max(dist, local, time.perf_counter() - t0)
    launches = lib.gb200_launch_count() - l0
    locked = sum(1 for lp in loops for c in range(0, per, 64) if lp.channel(c)["state"] == 4)
    for lp in loops:
        lp.close()
    for e in engs:
        e.close()
    if rank != 0:
        return None
    one = np.concatenate(one_bank_recs) if one_bank_recs else np.zeros(0)
    got = np.concatenate(kept) if kept else np.zeros(0)
    same = len(one) == len(got) and len(one) > 0
    if same:
        for c in sub:
            a, b_ = one[one["channel"] == c], got[got["channel"] == c]
            same = same and len(a) == len(b_) and all(a[f].tobytes() == b_[f].tobytes() for f in a.dtype.names)
    return {"value": float(cmds) * N_BLK / 1e6 / wall, "unit": "channel-Msps/s", "ms_per_period": wall / K * 1e3, "banks": n_banks, "channels_per_bank": per,
            "h2d_bytes_per_step": n_banks * N_BLK * 8, "d2h_bytes_per_step": int(72 * recs / K) + 32 * n_banks, "gpu_launches": int(launches), "records": int(recs),
            "channels_in_narrow_tracking_of_sampled": f"{locked}/{n_banks * len(range(0, per, 64))}",
            "records_identical_to_one_bank": bool(same), "records_compared": int(len(got)),
            "form": f"{n_banks} banks of {per} channels on the GPU, gb200_loop_submit / gb200_loop_collect per bank, two steps in flight per bank"}


def run_closed_loop(args, rank, world, local, dist, flush):
    import torch
    import bench
    from gnss_sim_receiver_b200 import _capi
    from gnss_sim_receiver_b200.tracking import TrackingEngine, TrackingLoop, loop_config
    lib = _capi.lib()
    nch = bench.CHANNELS_PER_GPU
    codes = bench.gps_codes()
    pinned, sc = _stream(local, codes, 100 + rank)
    host = pinned.numpy()
    periods_lock = 2700                      # pull-in (2 s) + bit synchronisation on the preamble
    K = max(1000, args.steps * 50)           # timed consecutive periods
    eng = TrackingEngine(device=local, max_channels=nch, max_code_len=1023, ring_samples=16 * N_BLK, max_block=N_BLK + 8, max_cmds=nch)
    cfg = loop_config("1C", FS, pull_in_time_s=1, cn0_samples=20, cn0_min=25, max_code_lock_fail=50, max_carrier_lock_fail=5000, carrier_lock_th=0.5,
                      early_late_space_chips=0.5, n_correlators=bench.N_TAPS)  # 11 taps, 0.1 chip apart: E, P, L are taps 0, 5, 10
    rng = np.random.default_rng(7 + rank)
    sat = np.arange(nch) % NSAT
    for c in range(nch):
        eng.set_code(c, codes[sat[c]])
    loop = TrackingLoop(eng, cfg, nch)
    acq_delay = np.mod(sc["delay"] / 1.023e6 * FS, FS / 1000.0)
    acq = [(float(acq_delay[sat[c]] + rng.uniform(-2, 2)), float(sc["doppler"][sat[c]] + rng.uniform(-40, 40))) for c in range(nch)]
    for c in range(nch):
        loop.start(c, int(sat[c]) + 1, acq[c][0], acq[c][1], 0, 0)
    # ---- the same stream through the product's host loop for a subset (parity reference), rank 0 only
    sub = list(range(0, nch, nch // 8))[:8]
    bank = None
    if rank == 0:
        from gnss_sim_receiver_b200 import _hostapi as H
        conf = ("GNSS-SDR.internal_fs_sps=25000000\nTracking_1C.implementation=GPS_L1_CA_DLL_PLL_Tracking_B200\nTracking_1C.item_type=gr_complex\n"
                "Tracking_1C.early_late_space_chips=0.5\nTracking_1C.pull_in_time_s=1\nTracking_1C.cn0_samples=20\nTracking_1C.cn0_min=25\nTracking_1C.max_lock_fail=50\n"
                "Tracking_1C.max_carrier_lock_fail=5000\nTracking_1C.carrier_lock_th=0.5\n")
        bank = H.TrackingBank(conf, "Tracking_1C", len(sub))
        for j, c in enumerate(sub):
            bank.start(j, int(sat[c]) + 1, acq[c][0], acq[c][1], 0, 0)

    def block(p):
        o = (p * N_BLK) % len(host)
        return host[o:o + N_BLK]

    n_rec = 0
    dev_recs, host_recs = [], []
    for p in range(periods_lock):
        r = loop.step(block(p), copy=False)
        n_rec += len(r)
        if bank is not None:
            dev_recs.append(r[np.isin(r["channel"], sub)].copy())
            bank.push(block(p))
            host_recs.append(bank.step())
    bench.barrier(dist)
    # ---- timed: K consecutive periods, blocking form (gb200_loop_step: one call per period, returns that period's records)
    dev_ms = 0.0
    cmds = 0
    t0 = time.perf_counter()
    timed_recs = []
    for p in range(periods_lock, periods_lock + K):
        r = loop.step(block(p), copy=False)
        n_rec += len(r)
        if bank is not None and len(r):
            timed_recs.append(r.copy())  # (kept for the parity check after the timed region; ~36 KB)
        st = loop.last_step
        dev_ms += st["device_ms"]
        cmds += st["commands"]
    torch.cuda.synchronize()
    wall_blocking = time.perf_counter() - t0
    # ---- timed: K more consecutive periods, pipelined form (gb200_loop_submit / gb200_loop_collect, two steps in flight: the host reads
    # period k's records while the device runs period k + 1) — the form a real-time receiver uses; this is the e2e figure
    bench.barrier(dist)
    l0 = lib.gb200_launch_count()
    t0 = time.perf_counter()
    cmds2 = 0
    first = periods_lock + K
    loop.submit(block(first))
    for p in range(first + 1, first + K):
        loop.submit(block(p))
        r = loop.collect(copy=False)
        n_rec += len(r)
        cmds2 += loop.last_step["commands"]
        if bank is not None and len(r):
            timed_recs.append(r.copy())
    r = loop.collect(copy=False)
    n_rec += len(r)
    cmds2 += loop.last_step["commands"]
    if bank is not None and len(r):
        timed_recs.append(r.copy())
    torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    launches = lib.gb200_launch_count() - l0
    wall_blocking = bench.reduce_max(dist, local, wall_blocking)
    wall = bench.reduce_max(dist, local, wall)
    dev_ms = bench.reduce_max(dist, local, dev_ms)
    parity = None
    if bank is not None:
        for p in range(periods_lock, periods_lock + 2 * K):
            bank.push(block(p))
            host_recs.append(bank.step())
        dev = np.concatenate(dev_recs + [t[np.isin(t["channel"], sub)] for t in timed_recs])
        hst = np.concatenate(host_recs)
        d_cn0 = d_dop = 0.0
        same_state, same_signs, matched = True, True, 0
        for j, c in enumerate(sub):
            a, b = loop.channel(c), bank.channel_state(j)
            same_state = same_state and a["state"] == int(b["state"]) == 4
            g, h = dev[dev["channel"] == c], hst[hst["channel"] == j]
            # records of the same symbol carry the same sample counter (the host loop's may differ by one sample for a period)
            m = min(len(g), len(h))
            g, h = g[:m], h[:m]
            same = np.abs(g["tracking_sample_counter"].astype(np.int64) - h["tracking_sample_counter"].astype(np.int64)) <= 1
            same_signs = same_signs and bool(np.all(same)) and bool(np.array_equal(np.sign(g["prompt_i"]), np.sign(h["prompt_i"])))
            half = m // 2
            matched += m
            d_cn0 = max(d_cn0, float(np.max(np.abs(g["cn0_db_hz"][half:] - h["cn0_db_hz"][half:]))))
            d_dop = max(d_dop, float(np.sqrt(np.mean((g["carrier_doppler_hz"][half:] - h["carrier_doppler_hz"][half:]) ** 2))))
        parity = {"channels_compared_with_host_loop": len(sub), "records_matched": int(matched), "all_in_narrow_tracking_on_both": bool(same_state),
                  "same_symbols_and_sample_counters": bool(same_signs), "max_cn0_diff_db": d_cn0, "max_doppler_rms_diff_hz": d_dop,
                  "ok": bool(same_state and same_signs and d_cn0 < 0.1 and d_dop < 0.5),
                  "gates": "north_star's closed-loop bar: lock state, CN0 within 0.1 dB, Doppler within 0.5 Hz rms, identical telemetry symbols"}
        bank.close()
    locked = sum(1 for c in range(0, nch, 64) if loop.channel(c)["state"] == 4)
    kernel = loop.kernel
    loop.close()
    eng.close()
    two = run_two_banks(lib, codes, sat, acq, cfg, block, local, rank, dist, periods_lock, K, sub, dev_recs)
    if rank != 0:
        return None
    ch_samples = float(cmds) * N_BLK * world
    ch_samples2 = float(cmds2) * N_BLK * world
    return {
        "what": "device-resident DLL/PLL loop (gb200_loop_step): per period H2D of the IQ block, correlate kernel, loop-closure kernel, D2H of the emitted Gnss_Synchro records",
        "channels_per_gpu": nch, "consecutive_periods": K, "periods_before_timing": periods_lock, "kernel": kernel,
        "value": ch_samples / 1e6 / (dev_ms / 1e3), "unit": "channel-Msps/s", "device_ms_per_period": dev_ms / K,
        "e2e": {"value": ch_samples2 / 1e6 / wall, "unit": "channel-Msps/s", "ms_per_period": wall / K * 1e3, "h2d_bytes_per_step": N_BLK * 8,
                "d2h_bytes_per_step": int(72 * n_rec / (periods_lock + 2 * K)) + 32, "form": "gb200_loop_submit / gb200_loop_collect, two steps in flight"},
        "e2e_blocking": {"value": ch_samples / 1e6 / wall_blocking, "unit": "channel-Msps/s", "ms_per_period": wall_blocking / K * 1e3, "form": "gb200_loop_step, one blocking call per period"},
        "real_time_margin": (N_BLK / FS) / (wall / K),
        "gpu_launches": int(launches), "records": int(n_rec), "channels_in_narrow_tracking_of_sampled": f"{locked}/{len(range(0, nch, 64))}",
        "parity_vs_host_loop": parity,
        "e2e_two_banks": None if two is None else dict(two, value=two["value"] * world),
    }
