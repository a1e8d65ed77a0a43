"""Breakdown of the tracking e2e step (host buffers): push_iq / enqueue / collect, fresh vs reused output buffer.
usage: python tools/e2e_probe.py [steps]"""
import sys, time, os
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from gnss_sim_receiver_b200 import _capi
from gnss_sim_receiver_b200.tracking import TrackingEngine

K = int(sys.argv[1]) if len(sys.argv) > 1 else 30
nch, N = bench.CHANNELS_PER_GPU, bench.N_BLK
w = bench.tracking_workload(0, nch)
codes = bench.gps_codes()
eng = TrackingEngine(device=0, max_channels=nch, max_code_len=bench.L_CODE, ring_samples=16 * N, max_block=N + 8, max_cmds=nch)
for c in range(nch):
    eng.set_code(c, codes[w["prn"][c] - 1])
stream = bench.iq_stream((2 * K + 16) * N, seed=1)
host_iq = torch.from_numpy(stream).pin_memory().numpy()
eng.push_iq(host_iq[:2 * N])
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda:0")
cmds = [eng.make_cmds(np.arange(nch), w["off"] + i * N, w["len"], w["shifts"], w["rem_carr"], w["phase_step"], w["rem_code"], w["code_step"]) for i in range(K + 4)]
lib = _capi.lib()
for mode in ("fresh", "reused", "pinned"):
    if mode == "reused":
        out = np.empty((nch, _capi.GB200_MAX_TAPS), dtype=np.complex64); out[:] = 0
    elif mode == "pinned":
        out = eng.host_outputs(nch)
    t = np.zeros(4)
    for i in range(K + 4):
        flush.zero_(); torch.cuda.synchronize()
        t0 = time.perf_counter()
        eng.push_iq(host_iq[(2 + i) * N:(3 + i) * N]) if mode == 'fresh' else None
        ci = i if mode == 'fresh' else K + 3
        t1 = time.perf_counter()
        if mode == "fresh":
            out = np.empty((nch, _capi.GB200_MAX_TAPS), dtype=np.complex64)
        _capi.check(lib.gb200_trk_enqueue(eng._h, cmds[ci], nch))
        t2 = time.perf_counter()
        _capi.check(lib.gb200_trk_collect(eng._h, out.ctypes.data))
        t3 = time.perf_counter()
        if i >= 4:
            t += [t1 - t0, t2 - t1, t3 - t2, t3 - t0]
    t *= 1e3 / K
    print(f"{mode:7s}: push {t[0]:.3f}  enqueue {t[1]:.3f}  collect {t[2]:.3f}  total {t[3]:.3f} ms")
# push_iq alone
tp = 0.0
for i in range(K):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    eng.push_iq(host_iq[(6 + K + i) * N:(7 + K + i) * N]); tp += time.perf_counter() - t0
print(f"push_iq alone {tp / K * 1e3:.3f} ms")
