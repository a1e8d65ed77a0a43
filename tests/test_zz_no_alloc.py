"""gb200_trk_correlate_noalloc (include/gnss_b200.h): the allocation-free variant of the general correlator kernel, whose scratch
(device copy of the commands, per-chunk partial sums, arrival counters) lives in a file-scope device array inside the library.
No counterpart in the reference — so the checker is the EXISTING entry point
(gb200_trk_correlate, itself parity-tested against the oracle in tests/test_trk_gpu.py): same inputs, same commands, outputs within the
tracking tests' tolerance (1e-4 of max_t |out_t| per command), on calls in a row that differ in shape so that stale scratch contents
(commands, counters, partial sums of the previous call) would show.

Named test_zz_* so that pytest collects it last: a failure here must not hide other tests.
Device buffers of the test come from torch (the TEST may allocate; the entry point does not).
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
TOL = 1e-4
MAX_TAPS = 16


def _taps(nt, spacing):
    return ((np.arange(nt) - (nt - 1) / 2.0) * spacing).astype(np.float32)


def _scene(seed, n_cmds, n, L, nt, spacing, step_nom, n_codes):
    rng = np.random.default_rng(seed)
    total = 3 * n + 64
    iq = (rng.standard_normal(total) + 1j * rng.standard_normal(total)).astype(np.complex64)
    codes = rng.choice([-1.0, 1.0], (n_codes, L)).astype(np.float32)
    ch = rng.integers(0, n_codes, n_cmds)
    offs = rng.integers(0, 2 * n, n_cmds)
    lens = n + rng.integers(-3, 4, n_cmds)
    dopp = rng.uniform(-5e3, 5e3, n_cmds)
    return dict(iq=iq, codes=codes, ch=ch, offs=offs, lens=lens, nt=nt,
                shifts=np.tile(_taps(nt, spacing), (n_cmds, 1)),
                remc=rng.uniform(0, 2 * np.pi, n_cmds).astype(np.float32), step=(2 * np.pi * dopp / 25e6).astype(np.float32),
                remcode=rng.uniform(-1.0, 1.0, n_cmds).astype(np.float32), cstep=(step_nom * (1 + dopp / 1575.42e6)).astype(np.float32))


def _existing(s):
    """the existing entry point: gb200_trk_correlate on an engine handle (few commands -> trk_correlate_kernel, the kernel the new entry shares)"""
    from gnss_sim_receiver_b200.tracking import TrackingEngine
    n_codes, L = s["codes"].shape
    eng = TrackingEngine(max_channels=n_codes, max_code_len=L, ring_samples=2 * len(s["iq"]), max_block=int(max(s["lens"])) + 8, max_cmds=len(s["ch"]))
    for c in range(n_codes):
        eng.set_code(c, s["codes"][c])
    first = eng.push_iq(s["iq"])
    cmds = eng.make_cmds(s["ch"], s["offs"] + first, s["lens"], s["shifts"], s["remc"], s["step"], s["remcode"], s["cstep"])
    out = eng.correlate(cmds).copy()
    assert eng.last_kernel.startswith("trk_correlate_kernel"), eng.last_kernel
    eng.close()
    return out


def _noalloc(s, stream=None):
    import torch
    from gnss_sim_receiver_b200 import _capi
    from gnss_sim_receiver_b200.tracking import TrackingEngine
    lib = _capi.lib()
    dev = torch.device("cuda", 0)
    iq = torch.from_numpy(s["iq"].view(np.float32).copy()).to(dev)
    codes = torch.from_numpy(s["codes"].copy()).to(dev)
    n = len(s["ch"])
    out = torch.full((n, MAX_TAPS, 2), float("nan"), dtype=torch.float32, device=dev)
    cmds = TrackingEngine.make_cmds(s["ch"], s["offs"], s["lens"], s["shifts"], s["remc"], s["step"], s["remcode"], s["cstep"])
    torch.cuda.synchronize()
    rc = lib.gb200_trk_correlate_noalloc(0, iq.data_ptr(), len(s["iq"]), codes.data_ptr(), s["codes"].shape[1], s["codes"].shape[0], cmds, n, out.data_ptr(),
                                         None if stream is None else C.c_void_p(stream.cuda_stream))
    assert rc == 0, lib.gb200_last_error().decode()
    o = out.cpu().numpy()
    return (o[..., 0] + 1j * o[..., 1]).astype(np.complex64), cmds, (iq, codes, out)


def _check(got, want, nt):
    assert np.all(np.isfinite(got[:, :nt].view(np.float32)))
    scale = np.max(np.abs(want), axis=1)
    err = np.max(np.abs(got[:, :nt] - want), axis=1) / scale
    assert np.max(err) < TOL, (float(np.max(err)), int(np.argmax(err)))
    return float(np.max(err))


def test_noalloc_entry_point_matches_the_existing_one_on_calls_in_a_row():
    import torch
    a = _scene(11, n_cmds=8, n=25000, L=1023, nt=11, spacing=0.1, step_nom=1.023e6 / 25e6, n_codes=4)      # C3 shape: 48 chunks per command
    b = _scene(12, n_cmds=5, n=4000, L=1023, nt=3, spacing=0.5, step_nom=1.023e6 / 4e6, n_codes=2)        # C1 shape: fewer commands, fewer chunks, other taps
    c = _scene(13, n_cmds=64, n=20000, L=10230, nt=5, spacing=0.15, step_nom=10.23e6 / 20e6, n_codes=3)   # C4 shape at the command limit
    want = {k: _existing(s) for k, s in (("a", a), ("b", b), ("c", c))}
    # two calls in a row on the shared scratch, then the first shape again and the largest one; a side stream for one of them
    got_a1, _, _ = _noalloc(a)
    got_b, _, _ = _noalloc(b)
    got_a2, _, _ = _noalloc(a, stream=torch.cuda.Stream())
    got_c, _, _ = _noalloc(c)
    got_b2, _, _ = _noalloc(b)
    worst = [_check(got_a1, want["a"], 11), _check(got_b, want["b"], 3), _check(got_a2, want["a"], 11), _check(got_c, want["c"], 5), _check(got_b2, want["b"], 3)]
    print(f"[noalloc] worst relative difference to gb200_trk_correlate per call: {worst}")
    # the same inputs give the same bits whatever ran in between (fixed-order reduce, counters cleared per call)
    assert np.array_equal(got_a1[:, :11].view(np.uint32), got_a2[:, :11].view(np.uint32))
    assert np.array_equal(got_b[:, :3].view(np.uint32), got_b2[:, :3].view(np.uint32))
    # only the padded tap slots are written: b has 3 taps -> slots 3.. keep the caller's contents
    assert np.all(np.isnan(got_b[:, 3:].real))


def test_noalloc_entry_point_rejects_what_the_fixed_scratch_cannot_hold():
    import torch
    from gnss_sim_receiver_b200 import _capi
    from gnss_sim_receiver_b200.tracking import TrackingEngine
    lib = _capi.lib()
    s = _scene(21, n_cmds=65, n=4000, L=1023, nt=3, spacing=0.5, step_nom=0.25575, n_codes=1)
    _, cmds, (iq, codes, out) = _noalloc({**s, **{k: s[k][:4] for k in ("ch", "offs", "lens", "shifts", "remc", "step", "remcode", "cstep")}})
    big = TrackingEngine.make_cmds(s["ch"], s["offs"], s["lens"], s["shifts"], s["remc"], s["step"], s["remcode"], s["cstep"])
    args = (0, iq.data_ptr(), len(s["iq"]), codes.data_ptr(), 1023, 1)
    out65 = torch.zeros((65, MAX_TAPS, 2), dtype=torch.float32, device="cuda:0")
    assert lib.gb200_trk_correlate_noalloc(*args, big, 65, out65.data_ptr(), None) == -1            # GB200_EINVAL
    assert b"GB200_NOALLOC_MAX_CMDS" in lib.gb200_last_error()
    cmds[0].n_samples = 64 * 16384 + 1
    assert lib.gb200_trk_correlate_noalloc(*args, cmds, 4, out.data_ptr(), None) == -1
    assert b"GB200_NOALLOC_MAX_SAMPLES" in lib.gb200_last_error()
    cmds[0].n_samples = 4000
    cmds[0].sample_offset = len(s["iq"]) - 100
    assert lib.gb200_trk_correlate_noalloc(*args, cmds, 4, out.data_ptr(), None) == -6              # GB200_ERANGE: window outside the samples passed
    cmds[0].sample_offset = 0
    cmds[0].flags = 1
    assert lib.gb200_trk_correlate_noalloc(*args, cmds, 4, out.data_ptr(), None) == -1              # GB200_TRK_HIGH_DYN is not served here
    cmds[0].flags = 0
    assert lib.gb200_trk_correlate_noalloc(*args, None, 4, out.data_ptr(), None) == -1
    assert lib.gb200_trk_correlate_noalloc(*args, cmds, 4, out.data_ptr(), None) == 0               # and the handle-free path still works afterwards
