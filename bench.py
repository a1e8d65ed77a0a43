#!/usr/bin/env python
"""bench.py — headline benchmark of the two hot paths on B200 (contract: see DESIGN.md §Measurement).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--quick]

One "step" = one code period (1 ms of signal at 25 MS/s) correlated for every channel resident on the GPU,
11 taps each: the north_star operating point (>= 10^4 channels x 11 taps x 25 MS/s per B200), C3-shaped
(GPS L1 C/A, gr_complex IQ, N = 25000 samples per period, 1023-chip codes).  Channels are sharded over ranks
(weak scaling: CHANNELS_PER_GPU per rank, no data-path collective).  The same JSON line carries
  "acq"          the C3 acquisition sweep (32 PRN x 80 Doppler bins, fft 25000) with its own e2e / roofline / cpu_baseline,
                 and at N > 1 the PRN-sharded and the Doppler-tiled forms of ONE sweep (strong scaling of its latency),
  "closed_loop"  the device-resident DLL/PLL loop over the same 10240 channels (correlate -> loop closure -> next commands),
  "configs"      the other BASELINE.json shapes (C1, C2, C4, C5 tracking; T = 3 / 5; C3-E1 and C4 acquisition),
  "legacy_cuda"  the reference's own cuda_multicorrelator.cu, unmodified, on this GPU,
  "parity_checked"  what was compared with the oracle, untimed, on the very outputs of the timed workloads.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

FS = 25e6
N_BLK = 25000
L_CODE = 1023
N_TAPS = 11
CHANNELS_PER_GPU = 10240
METRIC = "tracked channel-Msps/s (11 taps)"


def flop_per_ch_sample(taps):
    return 12 + 4 * taps  # wipe-off 6 + phasor advance 6 + 2 FMA per tap


def ncu_traffic():
    """dram bytes per launch of the dominant kernels from the newest ncu --set full summary committed under profiles/
    (profiles/ncu_traffic.json, written by tools/ncu_summary.py from the capture of THIS tree's kernels); null when absent."""
    p = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if os.path.exists(p):
        try:
            return json.load(open(p))
        except Exception:
            pass
    return {}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d.get("hbm_gbs", 6650.0), d.get("sm_max_mhz", 1965.0), "measured"
    return 6650.0, 1965.0, "fallback"


class ClockSampler(threading.Thread):
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self._stop_ev = index, [], threading.Event()

    def run(self):
        while not self._stop_ev.is_set():
            try:
                o = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i", str(self.index)],
                                   capture_output=True, text=True, timeout=5).stdout.strip()
                if o:
                    self.rows.append([x.strip() for x in o.split(",")])
            except Exception:
                pass
            self._stop_ev.wait(0.2)

    def summary(self):
        self._stop_ev.set()
        self.join(timeout=6)
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows if len(r) >= 6 for n, v in zip(names, r[2:6]) if v.lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons, "samples": len(sm)}


# ----------------------------------------------------------------------------------------------------------- synthetic inputs
def gps_codes():
    """32 GPS L1 C/A tables (IS-GPS-200 G2 delays).  Product-side generator, kept local so that the GPU arm never touches oracle/."""
    g2_delay = [5, 6, 7, 8, 17, 18, 139, 140, 141, 251, 252, 254, 255, 256, 257, 258, 469, 470, 471, 472, 473, 474, 509, 512, 513, 514, 515, 516, 859, 860, 861, 862]
    g1 = np.zeros(1023, np.int8)
    g2 = np.zeros(1023, np.int8)
    r1 = [1] * 10
    r2 = [1] * 10
    for i in range(1023):
        g1[i], g2[i] = r1[9], r2[9]
        f1 = r1[2] ^ r1[9]
        f2 = r2[1] ^ r2[2] ^ r2[5] ^ r2[7] ^ r2[8] ^ r2[9]
        r1 = [f1] + r1[:9]
        r2 = [f2] + r2[:9]
    out = np.zeros((32, 1023), np.float32)
    for p in range(32):
        out[p] = np.where(g1 ^ np.roll(g2, g2_delay[p]), 1.0, -1.0)
    return out


def random_codes(n, L, seed=7):
    return np.random.default_rng(seed).choice([-1.0, 1.0], (n, L)).astype(np.float32)


def iq_stream(n, seed=1):
    rng = np.random.default_rng(seed)
    return (rng.standard_normal(n) + 1j * rng.standard_normal(n)).astype(np.complex64)


def synth_sats(n, fs, tables, f_chip, sats, seed, t0=0.0):
    """sum_k A_k c_k(floor(t f_chip (1 + fd/fc) - delay) mod L) exp(j(2 pi fd t + phi)) + CN(0, 1)   (SURVEY §8d model)"""
    rng = np.random.default_rng(seed)
    t = t0 + np.arange(n, dtype=np.float64) / fs
    x = (rng.standard_normal(n) + 1j * rng.standard_normal(n)) / np.sqrt(2.0)
    for s in sats:
        tab = tables[s["prn"] - 1]
        k = np.floor(t * f_chip * (1.0 + s["doppler"] / 1575.42e6) - s["delay"]).astype(np.int64) % len(tab)
        x += np.sqrt(10.0 ** (s["cn0"] / 10.0) / fs) * tab[k] * np.exp(1j * (2 * np.pi * s["doppler"] * t + s["phase"]))
    return x.astype(np.complex64)


def quantise(x, np_t):
    scale, lim = (400.0 * np.sqrt(2.0), 32767) if np_t == np.int16 else (25.0 * np.sqrt(2.0), 127)
    q = np.empty(2 * len(x), np_t)
    q[0::2] = np.clip(np.rint(x.real * scale), -lim, lim)
    q[1::2] = np.clip(np.rint(x.imag * scale), -lim, lim)
    return q


def tracking_workload(rank, nch, fs=FS, n_blk=N_BLK, L=L_CODE, chip_rate=1.023e6, taps=N_TAPS, spacing=0.1, seed=1):
    """Seeded synthetic tracking state: one IQ stream, nch channels each with its own window / NCO state."""
    rng = np.random.default_rng(seed + 1000 * rank)
    dopp = rng.uniform(-5e3, 5e3, nch)
    return {
        "prn": (np.arange(nch) % 32) + 1,
        "off": rng.integers(0, n_blk, nch),                    # window start inside the previous period
        "len": n_blk + rng.integers(-1, 2, nch),
        "phase_step": (2 * np.pi * dopp / fs).astype(np.float32),
        "rem_carr": rng.uniform(0, 2 * np.pi, nch).astype(np.float32),
        "code_step": (chip_rate * (1 + dopp / 1575.42e6) / fs).astype(np.float32),
        "rem_code": rng.uniform(-0.5, 0.5, nch).astype(np.float32),
        "shifts": np.tile(((np.arange(taps) - (taps - 1) / 2) * spacing).astype(np.float32), (nch, 1)),
    }


# ------------------------------------------------------------------------------------------------------------- CPU baselines
def cpu_reference_tracking(budget_s=12.0, n_blk=N_BLK, taps=N_TAPS, L=L_CODE, code_step=1.023e6 / FS):
    """The reference's own CPU path (Cpu_Multicorrelator_Real_Codes over its AVX protokernels, compiled from the
    reference sources into oracle/_ref), one std::thread per channel as in cpu_multicorrelator_real_codes_test.cc:116-143,
    on all host cores; falls back to the scalar oracle port when oracle/_ref was not prebuilt."""
    import _oracle as O
    cores = os.cpu_count() or 1
    sig = iq_stream(2 * n_blk, seed=99)
    code = random_codes(1, L)[0] if L != 1023 else gps_codes()[0]
    shifts = ((np.arange(taps) - (taps - 1) / 2) * 0.1).astype(np.float32)
    step, cstep, remc = np.float32(2 * np.pi * 1234.0 / FS), np.float32(code_step), np.float32(0.4)
    if O.ref_available():
        r = O.ref()
        f = C.c_float
        call = lambda it: r.ref_mc_real_codes_bench(1, cores, n_blk, taps, L, it, sig.ctypes.data_as(O.f32p), code.ctypes.data_as(O.f32p),
                                                    shifts.ctypes.data_as(O.f32p), f(step), f(cstep), f(remc), None)
        t1 = call(4)
        iters = max(4, int(budget_s / max(t1 / 4, 1e-6)))
        secs = call(iters)
        kind, sample = "reference", f"{cores} threads x {iters} periods of N={n_blk}, {taps} taps, volk_gnsssdr a_avx/u_avx protokernels"
    else:
        nch = cores * 4
        w = tracking_workload(0, nch, n_blk=n_blk, L=L, taps=taps)
        codes = np.tile(code, (nch, 1))
        ring = iq_stream(2 * n_blk + 8, seed=99)
        t0 = time.perf_counter()
        iters = 0
        while time.perf_counter() - t0 < budget_s:
            O.mc_batch(ring, w["off"], w["len"], codes, w["shifts"], w["rem_carr"], w["phase_step"], w["rem_code"], w["code_step"], nthreads=cores)
            iters += 1
        secs = time.perf_counter() - t0
        iters *= 4
        kind, sample = "port", f"{nch} channels x {iters // 4} periods of N={n_blk}, {taps} taps, scalar C port with OpenMP"
    value = cores * iters * n_blk / 1e6 / secs
    return {"value": value, "unit": "channel-Msps/s", "cores": cores, "kind": kind, "sample": sample}, secs


ACQ_DMAX, ACQ_DSTEP, ACQ_PRNS = 10000, 250, 32


def gps_code_sampled(chips, fs, n):
    """gps_l1_ca_code_gen_complex_sampled (src/algorithms/libs/gps_sdr_signal_replica.cc:145-187): fp32 index law
    aux = ts*(i+1)/tc, chip = int(aux + 1) - 1, last sample forced to the last chip; this fork puts the chips in the
    imaginary part (:127-139)."""
    f32 = np.float32
    ts, tc = f32(1.0) / f32(fs), f32(1.0) / f32(1.023e6)
    i = np.arange(n, dtype=np.float32)
    aux = (ts * (i + f32(1.0))) / tc
    k = (aux + f32(1.0)).astype(np.int64) - 1
    k[-1] = 1022
    return (1j * chips[k]).astype(np.complex64)


def acq_signal(n, seed):
    """the C3 acquisition input: 8 GPS satellites at 44-50 dB-Hz + noise (what the parity check and the timed sweep both see)"""
    rng = np.random.default_rng(seed)
    sats = [dict(prn=int(p), doppler=float(rng.uniform(-9500, 9500)), delay=float(rng.uniform(0, 1023)), cn0=float(rng.uniform(44, 50)), phase=float(rng.uniform(0, 6.28)))
            for p in (2, 5, 9, 13, 17, 22, 26, 31)]
    return synth_sats(n, FS, gps_codes(), 1.023e6, sats, seed), sats


def cpu_reference_acq_pocketfft(budget_s=10.0, n=N_BLK, bins=80, fs=FS):
    """BASELINE.md §4's acquisition CPU baseline: acquisition_core restated (pcps_acquisition.cc:600-697) with scipy.fft (pocketfft,
    complex64, workers = all cores) as the stand-in for FFTW (FFTW / VOLK / GNU Radio are not in the reference tree and not installable
    here), the wipe-off table from the oracle's restatement of the reference's sincos kernel (built once, as the reference does at
    init), element-wise ops in numpy.  One PRN per pass like the reference: per bin wipe-off, FFT, x code spectrum, IFFT, |.|^2, max."""
    import _oracle as O
    import scipy.fft as sfft
    cores = os.cpu_count() or 1
    sig, _ = acq_signal(n, 5)
    code = gps_code_sampled(gps_codes()[0], fs, n)
    fc = np.conj(sfft.fft(code)).astype(np.complex64)
    wipe = np.stack([O.acq_wipeoff(n, float(-ACQ_DMAX + ACQ_DSTEP * k), fs) for k in range(bins)])

    def one_prn():
        x = sig[None, :] * wipe                                           # volk_32fc_x2_multiply_32fc :643
        X = sfft.fft(x, axis=1, workers=cores)                            # d_fft_if->execute() :647
        X *= fc[None, :]                                                  # :650
        y = sfft.ifft(X, axis=1, workers=cores, norm="forward")           # d_ifft->execute() :653 (unnormalised)
        g = y.real * y.real + y.imag * y.imag                             # :659
        j = int(np.argmax(g))                                             # :505-514
        return j, float(g.reshape(-1)[j])

    one_prn()
    t0 = time.perf_counter()
    passes = 0
    while time.perf_counter() - t0 < budget_s:
        one_prn()
        passes += 1
    secs = time.perf_counter() - t0
    return {"value": passes * bins / secs, "unit": "cells/s", "cores": cores, "kind": "port",
            "sample": f"{passes} PRN passes x {bins} Doppler bins, fft {n}, scipy.fft pocketfft complex64 workers={cores} as the FFTW stand-in (BASELINE.md §4), numpy element-wise"}


def cpu_reference_acq_oracle(budget_s=8.0, n=N_BLK):
    """second CPU figure: the oracle's own C restatement (its mixed-radix C FFT), OpenMP over Doppler bins"""
    import _oracle as O
    cores = os.cpu_count() or 1
    conf = O.acq_conf(FS, n, n, ACQ_DMAX, ACQ_DSTEP, float(n), int(np.ceil(FS / 1.023e6)))
    sig = iq_stream(n, seed=5)
    fc = O.acq_set_local_code(conf, gps_code_sampled(gps_codes()[0], FS, n))
    O.acq_dwell(conf, sig, fc, nthreads=cores)
    t0 = time.perf_counter()
    sweeps = 0
    while time.perf_counter() - t0 < budget_s:
        O.acq_dwell(conf, sig, fc, nthreads=cores)
        sweeps += 1
    secs = time.perf_counter() - t0
    return {"value": sweeps * conf.num_doppler_bins / secs, "unit": "cells/s", "cores": cores, "kind": "port",
            "sample": f"{sweeps} PRN sweeps x {conf.num_doppler_bins} Doppler bins, fft {n}, oracle C FFT, OpenMP {cores} threads"}


# ------------------------------------------------------------------------------------------------------------------ plumbing
def dist_setup(n_gpus):
    """One process per GPU (torchrun env) — torch.distributed is plumbing for the barrier, the max over ranks and the small gathers."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    return rank, world, local, dist


def reduce_max(dist, local, x):
    from gnss_sim_receiver_b200.sharding import reduce_max as _rm
    return _rm(dist, x, device=f"cuda:{local}")


def barrier(dist):
    import torch
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()


# ---------------------------------------------------------------------------------------------------------------- acquisition
def acq_parity(res, sig, codes_sampled, n, bins):
    """the timed sweep's 32 results against the oracle restatement of acquisition_core on the same block (untimed)"""
    import _oracle as O
    conf = O.acq_conf(FS, n, n, ACQ_DMAX, ACQ_DSTEP, float(n), int(np.ceil(FS / 1.023e6)))
    same = 0
    worst = 0.0
    for p in range(ACQ_PRNS):
        fc = O.acq_set_local_code(conf, codes_sampled[p])
        ro = O.acq_statistic(conf, O.acq_dwell(conf, sig, fc), 1)
        r = res[p]
        ok = (r.indext, r.doppler_index) == (ro.indext, ro.doppler_index) and r.acq_delay_samples == ro.acq_delay_samples and r.acq_doppler_hz == ro.acq_doppler_hz
        same += int(ok)
        worst = max(worst, abs(r.test_statistics - ro.test_statistics) / ro.test_statistics)
    return {"acq_prns_identical_index_and_bin": same, "acq_prns": ACQ_PRNS, "acq_worst_statistic_rel_err": worst}


def run_acq(args, rank, world, local, dist, flush, parity):
    """C3 acquisition sweep on every rank (one antenna stream per GPU): 32 PRN x 80 Doppler bins, fft 25000, CFAR; at N > 1 also ONE
    sweep sharded over the ranks by PRN (p mod G) and by Doppler tile (north_star's split; merge of the per-bin cell reductions)."""
    import torch
    from gnss_sim_receiver_b200 import _capi, sharding
    from gnss_sim_receiver_b200.acquisition import AcquisitionEngine
    lib = _capi.lib()
    K, W = args.steps, max(args.warmup, 3)
    n = N_BLK
    spchip = int(np.ceil(FS / 1.023e6))
    eng = AcquisitionEngine(fs_in=FS, fft_size=n, samples_per_code=float(n), samples_per_chip=spchip, max_prns=ACQ_PRNS, device=local)
    eng.set_doppler_grid(ACQ_DMAX, ACQ_DSTEP)
    codes = gps_codes()
    sampled = [gps_code_sampled(codes[p], FS, n) for p in range(ACQ_PRNS)]
    for p in range(ACQ_PRNS):
        eng.set_local_code(p, sampled[p])
    bins = eng.num_doppler_bins
    stream, sats = acq_signal((K + W) * n, 3 + rank)
    pinned = torch.from_numpy(stream).pin_memory()
    host = pinned.numpy()
    e2e_t = 0.0
    res0 = None
    for i in range(K + W):
        flush.zero_()
        barrier(dist) if i == W else torch.cuda.synchronize()
        t0 = time.perf_counter()
        res = eng.run(host[i * n:(i + 1) * n])
        dt = time.perf_counter() - t0
        if i >= W:
            e2e_t += dt
        if i == 0:
            res0 = res
    if rank == 0 and parity is not None:
        parity.update(acq_parity(res0, host[:n], sampled, n, bins))
        parity["acq_present_prns_detected"] = int(sum(1 for s in sats if res0[s["prn"] - 1].test_statistics > eng.threshold(0.001)))
    e2e_t = reduce_max(dist, local, e2e_t)
    for _ in range(W):
        eng.replay(1)
    barrier(dist)
    l0 = lib.gb200_launch_count()
    dev_ms = 0.0
    for _ in range(K):
        flush.zero_()
        torch.cuda.synchronize()
        dev_ms += eng.replay(1)
    launches = lib.gb200_launch_count() - l0
    dev_ms = reduce_max(dist, local, dev_ms)
    single = [(r.indext, r.doppler_index, r.test_statistics) for r in eng.run(host[:n])]
    eng.close()

    sharded = None
    if world > 1:
        # ---- ONE sweep (rank 0's block, broadcast = each rank uploads the same 200 KB) split over the ranks; strong scaling of its latency
        blk = torch.from_numpy(stream[:n].copy()).to(f"cuda:{local}")
        dist.broadcast(blk, src=0)
        blk_host = blk.cpu().numpy()
        sharded = {}
        # (1) by PRN: slot p -> rank p mod G, whole grids stay local, one result struct per PRN is gathered
        mine = [p for p in range(ACQ_PRNS) if p % world == rank]
        e1 = AcquisitionEngine(fs_in=FS, fft_size=n, samples_per_code=float(n), samples_per_chip=spchip, max_prns=max(1, len(mine)), device=local)
        e1.set_doppler_grid(ACQ_DMAX, ACQ_DSTEP)
        for i, p in enumerate(mine):
            e1.set_local_code(i, sampled[p])
        t_sum = 0.0
        got = None
        for i in range(K + W):
            barrier(dist)
            t0 = time.perf_counter()
            r = e1.run(blk_host, n_slots=len(mine))
            rec = torch.tensor([[p, x.indext, x.doppler_index, x.test_statistics] for p, x in zip(mine, r)], dtype=torch.float64, device=f"cuda:{local}")
            allr = [torch.empty_like(rec) for _ in range(world)] if len(mine) * world == ACQ_PRNS else None
            if allr is not None:
                dist.all_gather(allr, rec)
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
            if i >= W:
                t_sum += dt
            got = allr
        t_sum = reduce_max(dist, local, t_sum)
        e1.close()
        merged = {}
        if got is not None:
            for t in got:
                for row in t.cpu().numpy():
                    merged[int(row[0])] = (int(row[1]), int(row[2]), float(np.float32(row[3])))
        sharded["by_prn"] = {"value": ACQ_PRNS * bins * K / t_sum, "unit": "cells/s", "ms_per_sweep": t_sum / K * 1e3, "scaling": "strong",
                             "identical_to_single_gpu": bool(rank != 0 or all(merged.get(p) == (single[p][0], single[p][1], float(np.float32(single[p][2]))) for p in range(ACQ_PRNS)))}
        # (2) by Doppler tile: rank k sweeps bins [kB/G, (k+1)B/G) for every PRN; the per-(PRN, bin) reductions are all-gathered (NCCL) and
        # merged with the reference's first-bin / first-index rule on every rank
        lo, hi = sharding.doppler_tiles(bins, world)[rank]
        tile_widths = [b1 - b0 for b0, b1 in sharding.doppler_tiles(bins, world)]
        e2 = AcquisitionEngine(fs_in=FS, fft_size=n, samples_per_code=float(n), samples_per_chip=spchip, max_prns=ACQ_PRNS, device=local)
        e2.set_doppler_bins([float(-ACQ_DMAX + ACQ_DSTEP * k) for k in range(lo, hi)])
        for p in range(ACQ_PRNS):
            e2.set_local_code(p, sampled[p])
        t_sum = 0.0
        out = None
        for i in range(K + W):
            barrier(dist)
            t0 = time.perf_counter()
            e2.run(blk_host)
            v, im, sm = e2.read_cells(ACQ_PRNS)
            v, im, sm = sharding.gather_cell_tiles(dist, v, im, sm, world, device=f"cuda:{local}", widths=tile_widths)
            out = sharding.merge_cells(v, im, sm, n, 1, -ACQ_DMAX, ACQ_DSTEP, float(n))
            dt = time.perf_counter() - t0
            if i >= W:
                t_sum += dt
        t_sum = reduce_max(dist, local, t_sum)
        e2.close()
        same = all((o["indext"], o["doppler_index"]) == (single[p][0], single[p][1]) and np.float32(o["test_statistics"]) == np.float32(single[p][2]) for p, o in enumerate(out))
        sharded["by_doppler_tile"] = {"value": ACQ_PRNS * bins * K / t_sum, "unit": "cells/s", "ms_per_sweep": t_sum / K * 1e3, "scaling": "strong",
                                      "identical_to_single_gpu": bool(same), "exchange": "one NCCL all-gather of 12 bytes per (PRN, bin)"}
    if rank != 0:
        return None
    cells = ACQ_PRNS * bins * world
    hbm, sm_max, which = peaks()
    per_gpu_cells_s = ACQ_PRNS * bins / (dev_ms / K / 1e3)
    bytes_per_cell = 16 * n  # fc32 input 8N + conj code spectrum 8N, single dwell
    achieved = per_gpu_cells_s * bytes_per_cell / 1e9
    flop_per_cell = 10 * n * np.log2(n) + 17 * n
    fp32_peak = 148 * 128 * 2 * sm_max * 1e6 / 1e12
    out = {
        "metric": "acq PRNxDoppler cells/s", "value": cells * K / (dev_ms / 1e3), "unit": "cells/s", "ms_per_sweep": dev_ms / K,
        "config": {"workload": f"C3 acquisition sweep: GPS L1 C/A {FS / 1e6:g} MS/s gr_complex, fft {n}, {ACQ_PRNS} PRN x {bins} Doppler bins (+-{ACQ_DMAX} Hz / {ACQ_DSTEP} Hz), CFAR, 1 dwell, one sweep per GPU, 8 satellites present"},
        "e2e": {"value": cells * K / e2e_t, "unit": "cells/s", "h2d_bytes_per_step": n * 8, "d2h_bytes_per_step": ACQ_PRNS * C.sizeof(_capi.AcqResult)},
        "gpu_launches": int(launches),
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": hbm, "unit": "GB/s", "frac": achieved / hbm, "traffic": ncu_traffic().get("acq_inv_kernel"),
                     "fp32_frac": per_gpu_cells_s * flop_per_cell / 1e12 / fp32_peak,
                     "note": f"algorithmic bytes per cell = 16*N = {bytes_per_cell} (SURVEY 8d), x {ACQ_PRNS * bins} cells per launch; peak = {which} HBM copy bandwidth; X and the code spectra are "
                             f"L2-resident, so the FP32 co-limit (10 N log2 N + 17 N = {flop_per_cell / 1e6:.2f} MFLOP per cell against {fp32_peak:.1f} TFLOP/s) is the binding one: fp32_frac"},
    }
    if sharded:
        out["sharded_one_sweep"] = sharded
    return out


# ------------------------------------------------------------------------------------------------------- per-config measurements
def trk_config_bench(local, name, fs, n_blk, L, chip_rate, taps, iq_dtype, nch, iters, flush, spacing=0.1):
    """device-resident replay of one tracking shape (kernel only, L2 flushed between launches), plus a parity sample vs float64"""
    import torch
    from gnss_sim_receiver_b200.tracking import TrackingEngine
    w = tracking_workload(0, nch, fs=fs, n_blk=n_blk, L=L, chip_rate=chip_rate, taps=taps, spacing=spacing, seed=11)
    codes = random_codes(32, L)
    eng = TrackingEngine(device=local, max_channels=nch, max_code_len=L, iq_dtype=iq_dtype, ring_samples=4 * n_blk, max_block=n_blk + 8, max_cmds=nch)
    for c in range(nch):
        eng.set_code(c, codes[w["prn"][c] - 1])
    x = iq_stream(2 * n_blk + 8, seed=12)
    raw = x if iq_dtype == 0 else quantise(x, np.int16 if iq_dtype == 1 else np.int8)
    eng.push_iq(raw)
    cmds = eng.make_cmds(np.arange(nch), w["off"], w["len"], w["shifts"], w["rem_carr"], w["phase_step"], w["rem_code"], w["code_step"])
    eng.enqueue(cmds)
    got = eng.collect()[:, :taps]
    kernel = eng.last_kernel
    for _ in range(3):
        eng.replay(1)
    ms = 0.0
    for _ in range(iters):
        flush.zero_()
        torch.cuda.synchronize()
        ms += eng.replay(1)
    ms /= iters
    eng.close()
    # parity sample (untimed): 16 channels against the float64 oracle
    import _oracle as O
    xq = x if iq_dtype == 0 else (raw[0::2].astype(np.float32) + 1j * raw[1::2].astype(np.float32)).astype(np.complex64)
    worst = 0.0
    for c in range(0, nch, max(1, nch // 16)):
        r64 = O.mc_real_codes_f64(xq[w["off"][c]:w["off"][c] + w["len"][c]], codes[w["prn"][c] - 1], w["shifts"][c], w["rem_carr"][c], w["phase_step"][c], w["rem_code"][c], w["code_step"][c], mimic_inc=True)
        worst = max(worst, float(np.max(np.abs(got[c] - r64)) / np.max(np.abs(r64))))
    rate = float(np.sum(w["len"])) / (ms / 1e3)
    _, sm_max, _ = peaks()
    fp32_peak = 148 * 128 * 2 * sm_max * 1e6
    return {"workload": name, "value": rate / 1e6, "unit": "channel-Msps/s", "ms_per_step": ms, "kernel": kernel, "channels": nch, "taps": taps,
            "real_time_channels": rate / fs, "fp32_frac": rate * flop_per_ch_sample(taps) / fp32_peak, "parity_worst_rel_err_vs_f64": worst}


def acq_config_bench(local, name, fs, n, consumed, prns, dmax, dstep, dwells, iq_dtype, iters, flush, bit_transition=False):
    import torch
    from gnss_sim_receiver_b200.acquisition import AcquisitionEngine
    eng = AcquisitionEngine(fs_in=fs, fft_size=n, consumed_samples=consumed, samples_per_code=float(consumed), samples_per_chip=int(np.ceil(fs / 1.023e6)),
                            bit_transition=bit_transition, max_dwells=dwells, max_prns=prns, iq_dtype=iq_dtype, device=local)
    eng.set_doppler_grid(dmax, dstep)
    rng = np.random.default_rng(21)
    n_code = n // 2 if bit_transition else consumed
    for p in range(prns):
        eng.set_local_code(p, rng.choice([-1.0, 1.0], n_code).astype(np.complex64))
    x = iq_stream(consumed, seed=22)
    raw = x if iq_dtype == 0 else quantise(x, np.int16 if iq_dtype == 1 else np.int8)
    eng.run(raw)
    for _ in range(2):
        eng.replay(1)
    ms = 0.0
    for _ in range(iters):
        flush.zero_()
        torch.cuda.synchronize()
        ms += eng.replay(1)
    ms /= iters
    bins = eng.num_doppler_bins
    eng.close()
    return {"workload": name, "value": prns * bins / (ms / 1e3), "unit": "cells/s", "ms_per_sweep_one_dwell": ms, "cells_per_sweep": prns * bins, "fft_size": n,
            "ms_all_dwells": ms * dwells, "dwells": dwells}


def legacy_cuda():
    """the reference's own GPU kernel (cuda_multicorrelator.cu unmodified, built into oracle/_ref/legacy_cuda_bench), driven like
    gpu_multicorrelator_test.cc: per-channel launch + sync; at the reference test's shape and at the product's shape"""
    exe = os.path.join(ROOT, "oracle", "_ref", "legacy_cuda_bench")
    if not os.path.exists(exe):
        return {"unavailable": "oracle/_ref/legacy_cuda_bench not prebuilt (needs /root/reference at build time)"}
    out = []
    for n, taps, threads, iters in [(8192, 3, 1, 2000), (8192, 3, 8, 1000), (25000, 11, 1, 1000), (25000, 11, 16, 500)]:
        try:
            o = subprocess.run([exe, str(n), str(taps), str(threads), str(iters)], capture_output=True, text=True, timeout=120)
            line = [l for l in o.stdout.splitlines() if l.startswith("{")]
            out.append(json.loads(line[-1]) if line else {"n_samples": n, "n_taps": taps, "threads": threads, "error": (o.stderr or o.stdout)[-300:]})
        except Exception as e:  # a crash of the legacy kernel is a result too
            out.append({"n_samples": n, "n_taps": taps, "threads": threads, "error": repr(e)})
    return out


# -------------------------------------------------------------------------------------------------------------------- ours
def run_ours(args):
    import torch
    from gnss_sim_receiver_b200 import _capi
    from gnss_sim_receiver_b200.tracking import TrackingEngine
    rank, world, local, dist = dist_setup(args.gpus)
    torch.cuda.set_device(local)
    lib = _capi.lib()
    nch = CHANNELS_PER_GPU
    K, W = args.steps, max(args.warmup, 3)
    w = tracking_workload(rank, nch)
    codes = gps_codes()
    eng = TrackingEngine(device=local, max_channels=nch, max_code_len=L_CODE, ring_samples=16 * N_BLK, max_block=N_BLK + 8, max_cmds=nch)
    for c in range(nch):
        eng.set_code(c, codes[w["prn"][c] - 1])
    # pinned host staging for the per-step IQ block (what a signal source hands over every millisecond)
    stream = iq_stream((K + W + 4) * N_BLK, seed=1)
    pinned = torch.from_numpy(stream).pin_memory()
    host_iq = pinned.numpy()
    eng.push_iq(host_iq[: 2 * N_BLK])
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=f"cuda:{local}")  # > 126 MB L2

    def cmds_for(step_i):
        return eng.make_cmds(np.arange(nch), w["off"] + step_i * N_BLK, w["len"], w["shifts"], w["rem_carr"], w["phase_step"], w["rem_code"], w["code_step"])

    cmd_sets = [cmds_for(i) for i in range(K + W)]
    # ---- open-loop e2e: public call with HOST buffers, H2D of the IQ block and the commands, D2H of the correlator outputs inside the timed region
    out_bytes = nch * _capi.GB200_MAX_TAPS * 8
    corr_out = eng.host_outputs(nch)  # caller-owned, kept across calls like the reference's d_correlator_outs
    step_cmds = type(cmd_sets[0])()
    step_cmds._n_taps = cmd_sets[0]._n_taps
    e2e_t = 0.0
    first_out = None
    for i in range(K + W):
        flush.zero_()
        C.memmove(step_cmds, cmd_sets[i], C.sizeof(step_cmds))
        barrier(dist) if i == W else torch.cuda.synchronize()
        t0 = time.perf_counter()
        eng.push_iq(host_iq[(2 + i) * N_BLK:(3 + i) * N_BLK])
        res = eng.correlate(step_cmds, out=corr_out)
        dt = time.perf_counter() - t0
        if i >= W:
            e2e_t += dt
        if i == 0:
            first_out = np.array(res)
    e2e_t = reduce_max(dist, local, e2e_t)
    parity = {} if rank == 0 else None
    if rank == 0:
        # the outputs of timed step 0 against the oracle (untimed): a sample of the 10240 channels vs the float64 evaluation of the
        # reference's law, all of them vs the fp32 restatement that is bit-pinned to the reference's compiled objects
        import _oracle as O
        ring = stream[: 3 * N_BLK]
        cd = np.stack([codes[p - 1] for p in w["prn"]])
        ref32 = O.mc_batch(ring, w["off"], w["len"], cd, w["shifts"], w["rem_carr"], w["phase_step"], w["rem_code"], w["code_step"])
        e32 = np.max(np.abs(first_out - ref32), axis=1) / np.max(np.abs(ref32), axis=1)
        worst64 = 0.0
        sample = np.concatenate([np.argsort(e32)[-64:], np.arange(0, nch, nch // 192)])
        for c in sample:
            r64 = O.mc_real_codes_f64(ring[w["off"][c]:w["off"][c] + w["len"][c]], cd[c], w["shifts"][c], w["rem_carr"][c], w["phase_step"][c], w["rem_code"][c], w["code_step"][c], mimic_inc=True)
            worst64 = max(worst64, float(np.max(np.abs(first_out[c] - r64)) / np.max(np.abs(r64))))
        parity.update({"trk_channels_vs_fp32_oracle": nch, "trk_frac_within_1e-4_of_fp32_oracle": float(np.mean(e32 < 1e-4)), "trk_worst_rel_err_vs_fp32_oracle": float(np.max(e32)),
                       "trk_channels_vs_f64": int(len(sample)), "trk_worst_rel_err_vs_f64": worst64, "trk_ok": bool(worst64 < 1e-4),
                       "note": "noise-only input: the reference's sequential fp32 rotator is itself up to ~1e-3 from the float64 truth; the 64 worst channels vs fp32 are re-judged against float64"})
    # ---- device-resident: inputs and commands in HBM, CUDA events on the engine's stream, L2 flushed between steps
    eng.enqueue(cmd_sets[-1])
    eng.collect()
    kernel_name = eng.last_kernel
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    for _ in range(W):
        eng.replay(1)
    barrier(dist)
    l0 = lib.gb200_launch_count()
    dev_ms = 0.0
    for _ in range(K):
        flush.zero_()
        torch.cuda.synchronize()
        dev_ms += eng.replay(1)
    launches = lib.gb200_launch_count() - l0
    barrier(dist)
    # sustained (no flush, back to back) for the clock record: keeps the GPU under load long enough for nvidia-smi to see it
    sustained_ms = eng.replay(max(K, 50, int(1500.0 * K / max(dev_ms, 1e-3))))  # ~1.5 s: several nvidia-smi samples
    dev_ms = reduce_max(dist, local, dev_ms)
    sustained_ms = reduce_max(dist, local, sustained_ms)
    clocks = sampler.summary() if rank == 0 else None
    eng.close()
    closed = None
    try:
        from bench_closed_loop import run_closed_loop  # the device-resident DLL/PLL loop (gb200_bank_*)
        closed = run_closed_loop(args, rank, world, local, dist, flush)
    except ImportError:
        closed = None
    acq = run_acq(args, rank, world, local, dist, flush, parity)
    ch_samples_per_step = float(np.sum(w["len"])) * world
    value = ch_samples_per_step * K / 1e6 / (dev_ms / 1e3)
    e2e_value = ch_samples_per_step * K / 1e6 / e2e_t
    if rank != 0:
        return
    hbm, sm_max, which = peaks()
    fp32_peak = 148 * 128 * 2 * sm_max * 1e6 / 1e12
    per_gpu_rate = float(np.sum(w["len"])) / (dev_ms / K / 1e3)
    achieved = per_gpu_rate * flop_per_ch_sample(N_TAPS) / 1e12
    configs = None
    if not args.quick:
        it = 5
        configs = {"tracking": [
            trk_config_bench(local, "C1 GPS L1 4 MS/s gr_complex, 3 taps (E,P,L)", 4e6, 4000, 1023, 1.023e6, 3, 0, 8192, it, flush, spacing=0.5),
            trk_config_bench(local, "C2 Galileo E1 4 MS/s ishort, 11 taps, 4 ms codes (8184 entries, 2 per chip)", 4e6, 16000, 8184, 2.046e6, 11, 1, 4096, it, flush, spacing=0.15),
            trk_config_bench(local, "C3 GPS L1 25 MS/s gr_complex, 3 taps", FS, N_BLK, 1023, 1.023e6, 3, 0, nch, it, flush, spacing=0.5),
            trk_config_bench(local, "C3 GPS L1 25 MS/s gr_complex, 5 taps (VEML)", FS, N_BLK, 1023, 1.023e6, 5, 0, nch, it, flush, spacing=0.3),
            trk_config_bench(local, "C3 Galileo E1 25 MS/s gr_complex, 11 taps, 100000 samples per period", FS, 100000, 8184, 2.046e6, 11, 0, 2048, it, flush, spacing=0.15),
            trk_config_bench(local, "C4 GPS L5 / Galileo E5a 20 MS/s gr_complex, 11 taps, 10230-chip codes", 20e6, 20000, 10230, 10.23e6, 11, 0, 4096, it, flush),
            trk_config_bench(local, "C5 GPS L1 100 MS/s int8 IQ, 11 taps, 1024 channels per GPU", 100e6, 100000, 1023, 1.023e6, 11, 2, 1024, it, flush, spacing=0.05),
        ], "acquisition": [
            acq_config_bench(local, "C1 GPS L1 4 MS/s, fft 4000 x 80 bins x 32 PRN", 4e6, 4000, 4000, 32, 10000, 250, 1, 0, it, flush),
            acq_config_bench(local, "C2 Galileo E1 4 MS/s cshort, bit_transition, fft 32000 x 80 bins x 36 PRN", 4e6, 32000, 32000, 36, 5000, 125, 1, 1, it, flush, bit_transition=True),
            acq_config_bench(local, "C3 Galileo E1 25 MS/s, fft 100000 x 80 bins x 36 PRN", FS, 100000, 100000, 36, 5000, 125, 1, 0, 3, flush),
            acq_config_bench(local, "C4 GPS L5 / Galileo E5a 20 MS/s, fft 20000 x 80 bins x 32 PRN x 20 dwells", 20e6, 20000, 20000, 32, 10000, 250, 20, 0, it, flush),
            acq_config_bench(local, "C5 GPS L1 100 MS/s int8, fft 100000 x 80 bins x 32 PRN", 100e6, 100000, 100000, 32, 10000, 250, 1, 2, 3, flush),
        ]}
    cpu, _ = cpu_reference_tracking(12.0)
    # e2e = what a receiver calls every millisecond.  With the device-resident loop that is gb200_loop_step (host IQ block in, Gnss_Synchro
    # records out, commands formed on the device); the open-loop form (commands formed on the host and uploaded every step, all tap
    # outputs downloaded: gb200_trk_push_iq + gb200_trk_correlate) is kept beside it.
    e2e_open = {"value": e2e_value, "unit": "channel-Msps/s", "h2d_bytes_per_step": N_BLK * 8 + nch * C.sizeof(_capi.TrkCmd), "d2h_bytes_per_step": out_bytes,
                "what": "open loop through gb200_trk_push_iq + gb200_trk_correlate with host buffers (commands formed on the host every step)"}
    if closed is not None and closed.get("parity_vs_host_loop", {}).get("ok"):
        e2e_obj = dict(closed["e2e"], what="closed loop through gb200_loop_step with a host IQ block per code period (see closed_loop)")
        e2e_obj.pop("ms_per_period", None)
        # the same channels as two banks on the GPU (one bank's loop kernel under the other's correlator): the headline when its records
        # are identical to the one-bank run's and it is the faster form; the one-bank figure stays in closed_loop.e2e
        two = closed.get("e2e_two_banks")
        if two and two["records_identical_to_one_bank"] and two["value"] > e2e_obj["value"]:
            e2e_obj = {k: two[k] for k in ("value", "unit", "h2d_bytes_per_step", "d2h_bytes_per_step", "form")}
            e2e_obj["what"] = "closed loop, the GPU's channels as two banks, a host IQ block per code period into each (see closed_loop.e2e_two_banks; one bank: closed_loop.e2e)"
    else:
        e2e_obj = e2e_open
    if acq is not None:
        acq["cpu_baseline"] = cpu_reference_acq_pocketfft(8.0)
        acq["cpu_baseline_oracle_c_fft"] = cpu_reference_acq_oracle(5.0)
    line = {
        "metric": METRIC, "value": value, "unit": "channel-Msps/s", "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": dev_ms / K,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"C3-shaped tracking: GPS L1 C/A {FS / 1e6:g} MS/s gr_complex, N={N_BLK}/period, {N_TAPS} taps, {nch} channels per GPU",
                   "channels_per_gpu": nch, "l2": "256 MiB flush write between timed steps", "parallelism": f"channels sharded x{world}, no collective", "kernel": kernel_name},
        "e2e": e2e_obj, "e2e_open_loop": e2e_open,
        "gpu_launches": int(launches),
        "roofline": {"bound": "fp32", "achieved": achieved, "peak": fp32_peak, "unit": "TFLOP/s", "frac": achieved / fp32_peak, "traffic": ncu_traffic().get("trk_run_kernel"),
                     "note": f"FP32-FMA pipe roofline (SURVEY 8d): {flop_per_ch_sample(N_TAPS)} flop per channel-sample; peak = 148 SM x 128 lanes x 2 x sm_max_mhz ({which}); kernel moves ~0 HBM bytes per unit",
                     "sustained_ms_per_step": sustained_ms},
        "cpu_baseline": cpu, "clocks": clocks, "parity_checked": parity, "acq": acq, "closed_loop": closed, "configs": configs,
        "legacy_cuda": None if args.quick else legacy_cuda(),
    }
    if acq is not None:
        line["acq_value"] = acq["value"]  # top-level so that the per-N scaling record sees the acquisition metric too
        if "sharded_one_sweep" in acq:
            line["acq_one_sweep_doppler_tiled_cells_s"] = acq["sharded_one_sweep"]["by_doppler_tile"]["value"]
            line["acq_one_sweep_prn_sharded_cells_s"] = acq["sharded_one_sweep"]["by_prn"]["value"]
    emit(line)


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    K, W = args.steps, args.warmup
    per_step = min(8.0, max(0.5, 100.0 / max(1, K + W)))
    for _ in range(W):
        cpu_reference_tracking(per_step)
    vals, secs = [], 0.0
    for _ in range(K):
        cpu, s = cpu_reference_tracking(per_step)
        vals.append(cpu["value"])
        secs += s
    v = float(np.mean(vals))
    cpu["value"] = v
    acq_cpu = cpu_reference_acq_pocketfft(10.0)
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": "channel-Msps/s", "n_gpus": args.gpus, "steps": K, "warmup": W,
            "ms_per_step": secs / K * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"C3-shaped tracking: GPS L1 C/A {FS / 1e6:g} MS/s gr_complex, N={N_BLK}/period, {N_TAPS} taps; CPU: one thread per channel on all host cores"},
            "cpu_baseline": cpu, "e2e": {"value": v, "unit": "channel-Msps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0,
            # the metric's second half on the same host cores: restated acquisition_core over pocketfft, kind "port"
            "acq": {"metric": "acq PRNxDoppler cells/s", "value": acq_cpu["value"], "unit": "cells/s", "cpu_baseline": acq_cpu,
                    "e2e": {"value": acq_cpu["value"], "unit": "cells/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}},
            "acq_value": acq_cpu["value"]}
    emit(line)


_JSON_FD = None


def emit(line):
    """The one JSON line, on the process's ORIGINAL stdout (see main: libraries such as NCCL print banners to fd 1)."""
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_JSON_FD, data)


def main():
    global _JSON_FD
    # keep stdout clean for the contract's single JSON line: everything else written to fd 1 during the run (e.g. the
    # "NCCL version ..." banner of the first communicator) goes to stderr
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--quick", action="store_true", help="headline + acquisition only (skip the per-config table and the legacy kernel)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        try:
            run_ours(args)
        finally:
            import torch.distributed as dist
            if dist.is_available() and dist.is_initialized():
                dist.destroy_process_group()


if __name__ == "__main__":
    main()
