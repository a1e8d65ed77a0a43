#!/usr/bin/env python
"""bench.py — headline benchmark of the two hot paths on B200 (contract: see DESIGN.md §Measurement).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

One "step" = one code period (1 ms of signal at 25 MS/s) correlated for every channel resident on the GPU,
11 taps each: the north_star operating point (>= 10^4 channels x 11 taps x 25 MS/s per B200), C3-shaped
(GPS L1 C/A, gr_complex IQ, N = 25000 samples per period, 1023-chip codes).  Channels are sharded over ranks
(weak scaling: CHANNELS_PER_GPU per rank, no data-path collective).  The acquisition sweep (C3: 32 PRN x 80
Doppler bins, fft 25000) is measured on rank 0 and reported in the "acq" object of the same JSON line.
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
FLOP_PER_CH_SAMPLE = 12 + 4 * N_TAPS  # wipe-off 6 + phasor advance 6 + 2 FMA per tap
METRIC = "tracked channel-Msps/s (11 taps)"
# dram__bytes_read.sum + dram__bytes_write.sum of ONE launch of the dominant kernels at this workload, from the ncu --set full
# captures committed under profiles/ (r01_trk_v3_details / r01_acq_v3_details); a profiler figure, not measured by this script
TRK_NCU_DRAM_BYTES = 44993024 + 61952   # trk_run_kernel<11,CF32,9,1>, 10240 commands: code tables + IQ ring + commands
ACQ_NCU_DRAM_BYTES = 22503168 + 0       # acq_inv_kernel<0>, 2560 cells: X and the code spectra stay in L2


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


def tracking_workload(rank, nch, seed=1):
    """Seeded synthetic C3-shaped tracking state: one IQ stream, nch channels with their own window/NCO state."""
    rng = np.random.default_rng(seed + 1000 * rank)
    dopp = rng.uniform(-5e3, 5e3, nch)
    w = {
        "prn": (np.arange(nch) % 32) + 1,
        "off": rng.integers(0, N_BLK, nch),                    # window start inside the previous period
        "len": N_BLK + rng.integers(-1, 2, nch),
        "phase_step": (2 * np.pi * dopp / FS).astype(np.float32),
        "rem_carr": rng.uniform(0, 2 * np.pi, nch).astype(np.float32),
        "code_step": (1.023e6 * (1 + dopp / 1575.42e6) / FS).astype(np.float32),
        "rem_code": rng.uniform(-0.5, 0.5, nch).astype(np.float32),
        "shifts": np.tile(((np.arange(N_TAPS) - (N_TAPS - 1) / 2) * 0.1).astype(np.float32), (nch, 1)),
    }
    return w


def iq_stream(n, seed=1):
    rng = np.random.default_rng(seed)
    return (rng.standard_normal(n) + 1j * rng.standard_normal(n)).astype(np.complex64)


def gps_codes():
    """32 GPS L1 C/A tables.  Product-side generator (host C++ restated in python would be overkill here): the LFSR
    is 20 lines, kept local so that bench's GPU arm never touches oracle/."""
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
        d = g2_delay[p]
        out[p] = np.where(g1 ^ np.roll(g2, d), 1.0, -1.0)
    return out


def cpu_reference_tracking(budget_s=12.0):
    """The reference's own CPU path (Cpu_Multicorrelator_Real_Codes over its AVX protokernels, compiled from the
    reference sources into oracle/_ref), one std::thread per channel as in cpu_multicorrelator_real_codes_test.cc:116-143,
    on all host cores; falls back to the scalar oracle port when oracle/_ref was not prebuilt."""
    import _oracle as O
    cores = os.cpu_count() or 1
    sig = iq_stream(2 * N_BLK, seed=99)
    code = gps_codes()[0]
    shifts = ((np.arange(N_TAPS) - (N_TAPS - 1) / 2) * 0.1).astype(np.float32)
    step, cstep, remc = np.float32(2 * np.pi * 1234.0 / FS), np.float32(1.023e6 / FS), np.float32(0.4)
    if O.ref_available():
        r = O.ref()
        f = C.c_float
        call = lambda it: r.ref_mc_real_codes_bench(1, cores, N_BLK, N_TAPS, L_CODE, it, sig.ctypes.data_as(O.f32p), code.ctypes.data_as(O.f32p),
                                                    shifts.ctypes.data_as(O.f32p), f(step), f(cstep), f(remc), None)
        t1 = call(4)
        iters = max(4, int(budget_s / max(t1 / 4, 1e-6)))
        secs = call(iters)
        kind, sample = "reference", f"{cores} threads x {iters} periods of N={N_BLK}, {N_TAPS} taps, volk_gnsssdr a_avx/u_avx protokernels"
    else:
        nch = cores * 4
        w = tracking_workload(0, nch)
        codes = gps_codes()[(w["prn"] - 1)]
        ring = iq_stream(2 * N_BLK + 8, seed=99)
        t0 = time.perf_counter()
        iters = 0
        while time.perf_counter() - t0 < budget_s:
            O.mc_batch(ring, w["off"], w["len"], codes, w["shifts"], w["rem_carr"], w["phase_step"], w["rem_code"], w["code_step"], nthreads=cores)
            iters += 1
        secs = time.perf_counter() - t0
        iters *= 4
        kind, sample = "port", f"{nch} channels x {iters // 4} periods of N={N_BLK}, {N_TAPS} taps, scalar C port with OpenMP"
    value = cores * iters * N_BLK / 1e6 / secs
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


def cpu_reference_acq(budget_s=10.0):
    """CPU arm for acquisition: the oracle's restatement of acquisition_core (FFTW/VOLK are not in the reference tree, so the
    FFT is the oracle's own mixed-radix C; kind = "port"), OpenMP over Doppler bins on all host cores."""
    import _oracle as O
    cores = os.cpu_count() or 1
    n = N_BLK
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
            "sample": f"{sweeps} PRN sweeps x {conf.num_doppler_bins} Doppler bins, fft {n}, oracle C FFT (FFTW stand-in), OpenMP {cores} threads"}


def run_acq(args, rank, world, local, dist, flush):
    """C3 acquisition sweep on every rank (one antenna stream per GPU): 32 PRN x 80 Doppler bins, fft 25000, CFAR."""
    import torch
    from gnss_sim_receiver_b200 import _capi
    from gnss_sim_receiver_b200.acquisition import AcquisitionEngine
    lib = _capi.lib()
    K, W = args.steps, max(args.warmup, 3)
    n = N_BLK
    eng = AcquisitionEngine(fs_in=FS, fft_size=n, samples_per_code=float(n), samples_per_chip=int(np.ceil(FS / 1.023e6)), max_prns=ACQ_PRNS, device=local)
    eng.set_doppler_grid(ACQ_DMAX, ACQ_DSTEP)
    codes = gps_codes()
    for p in range(ACQ_PRNS):
        eng.set_local_code(p, gps_code_sampled(codes[p], FS, n))
    bins = eng.num_doppler_bins
    pinned = torch.from_numpy(iq_stream((K + W) * n, seed=3 + rank)).pin_memory()
    host = pinned.numpy()
    e2e_t = 0.0
    for i in range(K + W):
        flush.zero_()
        barrier(dist) if i == W else torch.cuda.synchronize()
        t0 = time.perf_counter()
        res = eng.run(host[i * n:(i + 1) * n])
        dt = time.perf_counter() - t0
        if i >= W:
            e2e_t += dt
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
    eng.close()
    if rank != 0:
        return None
    cells = ACQ_PRNS * bins * world
    hbm, _, which = peaks()
    per_gpu_cells_s = ACQ_PRNS * bins / (dev_ms / K / 1e3)
    bytes_per_cell = 16 * n  # fc32 input 8N + conj code spectrum 8N, single dwell
    achieved = per_gpu_cells_s * bytes_per_cell / 1e9
    return {
        "metric": "acq PRNxDoppler cells/s", "value": cells * K / (dev_ms / 1e3), "unit": "cells/s", "ms_per_sweep": dev_ms / K,
        "config": {"workload": f"C3 acquisition sweep: GPS L1 C/A {FS / 1e6:g} MS/s gr_complex, fft {n}, {ACQ_PRNS} PRN x {bins} Doppler bins (+-{ACQ_DMAX} Hz / {ACQ_DSTEP} Hz), CFAR, 1 dwell, one sweep per GPU"},
        "e2e": {"value": cells * K / e2e_t, "unit": "cells/s", "h2d_bytes_per_step": n * 8, "d2h_bytes_per_step": ACQ_PRNS * C.sizeof(_capi.AcqResult)},
        "gpu_launches": int(launches),
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": hbm, "unit": "GB/s", "frac": achieved / hbm, "traffic": ACQ_NCU_DRAM_BYTES,
                     "note": f"algorithmic bytes per cell = 16*N = {bytes_per_cell} (SURVEY 8d), x {ACQ_PRNS * bins} cells per launch; peak = {which} HBM copy bandwidth; "
                             "traffic = dram bytes of one acq_inv_kernel launch from the ncu --set full capture under profiles/ (spectra are L2-resident)"},
        "cpu_baseline": cpu_reference_acq(8.0),
    }


def dist_setup(n_gpus):
    """One process per GPU (torchrun env) — torch.distributed is plumbing for the barrier and the max over ranks."""
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
    # ---- e2e: public call with HOST buffers, H2D of the IQ block and D2H of the correlator outputs inside the timed region
    out_bytes = nch * _capi.GB200_MAX_TAPS * 8
    corr_out = eng.host_outputs(nch)  # caller-owned, kept across calls like the reference's d_correlator_outs
    # the step's commands are written into ONE caller-owned array just before the call, as the loop closure does every
    # period (dll_pll_veml_tracking.cc:1037-1062 fills the arguments right before do_correlation_step): untimed here, it
    # only puts the 1.1 MB of commands where a closed loop would have left them (host caches) instead of cold in DRAM
    step_cmds = type(cmd_sets[0])()
    step_cmds._n_taps = cmd_sets[0]._n_taps
    e2e_t = 0.0
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
    e2e_t = reduce_max(dist, local, e2e_t)
    # ---- device-resident: inputs and commands in HBM, CUDA events on the engine's stream, L2 flushed between steps
    eng.enqueue(cmd_sets[-1])
    eng.collect()
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
    acq = run_acq(args, rank, world, local, dist, flush)
    ch_samples_per_step = float(np.sum(w["len"])) * world
    value = ch_samples_per_step * K / 1e6 / (dev_ms / 1e3)
    e2e_value = ch_samples_per_step * K / 1e6 / e2e_t
    if rank != 0:
        return
    hbm, sm_max, which = peaks()
    fp32_peak = 148 * 128 * 2 * sm_max * 1e6 / 1e12
    per_gpu_rate = float(np.sum(w["len"])) / (dev_ms / K / 1e3)
    achieved = per_gpu_rate * FLOP_PER_CH_SAMPLE / 1e12
    cpu, _ = cpu_reference_tracking(12.0)
    line = {
        "metric": METRIC, "value": value, "unit": "channel-Msps/s", "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": dev_ms / K,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"C3-shaped tracking: GPS L1 C/A {FS / 1e6:g} MS/s gr_complex, N={N_BLK}/period, {N_TAPS} taps, {nch} channels per GPU",
                   "channels_per_gpu": nch, "l2": "256 MiB flush write between timed steps", "parallelism": f"channels sharded x{world}, no collective"},
        "e2e": {"value": e2e_value, "unit": "channel-Msps/s", "h2d_bytes_per_step": N_BLK * 8 + nch * C.sizeof(_capi.TrkCmd), "d2h_bytes_per_step": out_bytes},
        "gpu_launches": int(launches),
        "roofline": {"bound": "fp32", "achieved": achieved, "peak": fp32_peak, "unit": "TFLOP/s", "frac": achieved / fp32_peak, "traffic": TRK_NCU_DRAM_BYTES,
                     "note": f"FP32-FMA pipe roofline (SURVEY 8d): {FLOP_PER_CH_SAMPLE} flop per channel-sample; peak = 148 SM x 128 lanes x 2 x sm_max_mhz ({which}); kernel moves ~0 HBM bytes per unit",
                     "sustained_ms_per_step": sustained_ms},
        "cpu_baseline": cpu, "clocks": clocks, "acq": acq,
    }
    emit(line)


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    K, W = args.steps, args.warmup
    per_step = min(10.0, max(0.5, 120.0 / max(1, K + W)))
    for _ in range(W):
        cpu_reference_tracking(per_step)
    vals, secs = [], 0.0
    for _ in range(K):
        cpu, s = cpu_reference_tracking(per_step)
        vals.append(cpu["value"])
        secs += s
    v = float(np.mean(vals))
    cpu["value"] = v
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": "channel-Msps/s", "n_gpus": args.gpus, "steps": K, "warmup": W,
            "ms_per_step": secs / K * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"C3-shaped tracking: GPS L1 C/A {FS / 1e6:g} MS/s gr_complex, N={N_BLK}/period, {N_TAPS} taps; CPU: one thread per channel on all host cores"},
            "cpu_baseline": cpu, "e2e": {"value": v, "unit": "channel-Msps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
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
