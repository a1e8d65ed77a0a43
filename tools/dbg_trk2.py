import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, _oracle as O
import test_trk_gpu as T
fs, L, chip_rate, nt, spacing, iq_dtype, nch = 100e6, 1023, 1.023e6, 11, 0.05, 2, 1280
rng = np.random.default_rng(int(fs / 1e6) + nt)
n = int(round(fs * L / chip_rate)); spread = n; total = spread + n + 64
rawi = np.clip(np.rint(rng.standard_normal((total, 2)) * 25.0), -127, 127).astype(np.int8)
ring = (rawi[:, 0].astype(np.float32) + 1j * rawi[:, 1].astype(np.float32)).astype(np.complex64)
codes = rng.choice([-1.0, 1.0], (nch, L)).astype(np.float32)
eng = T._engine(max_channels=nch, max_code_len=L, iq_dtype=2, ring_samples=2 * total, max_block=n + 8, max_cmds=nch)
for c in range(nch): eng.set_code(c, codes[c])
eng.push_iq(np.zeros(10, np.int8)); first = eng.push_iq(rawi.reshape(-1))
offs = rng.integers(0, spread, nch); lens = n + rng.integers(-1, 2, nch); dopp = rng.uniform(-5e3, 5e3, nch)
step = (2 * np.pi * dopp / fs).astype(np.float32); remc = rng.uniform(0, 2 * np.pi, nch).astype(np.float32)
cstep = ((chip_rate * (1 + dopp / 1575.42e6)) / fs).astype(np.float32); remcode = rng.uniform(-1.0, 1.0, nch).astype(np.float32)
shifts = np.tile(T._taps(nt, spacing), (nch, 1))
cmds = eng.make_cmds(np.arange(nch), offs + first, lens, shifts, remc, step, remcode, cstep)
got = eng.correlate(cmds); print(eng.last_kernel)
ref = O.mc_batch(ring, offs, lens, codes, shifts, remc, step, remcode, cstep)
err = np.max(np.abs(got - ref), axis=1) / np.max(np.abs(ref), axis=1)
bad = np.argsort(err)[-5:]
for c in bad:
    r64 = O.mc_real_codes_f64(ring[offs[c]:offs[c] + lens[c]], codes[c], shifts[c], remc[c], step[c], remcode[c], cstep[c], mimic_inc=True)
    r64n = O.mc_real_codes_f64(ring[offs[c]:offs[c] + lens[c]], codes[c], shifts[c], remc[c], step[c], remcode[c], cstep[c], mimic_inc=False)
    s = np.max(np.abs(r64))
    print(c, "got-ref32 %.2e  got-f64 %.2e  ref32-f64 %.2e  f64mimic-f64plain %.2e  dopp %.0f step %.3e" % (err[c], np.max(np.abs(got[c]-r64))/s, np.max(np.abs(ref[c]-r64))/s, np.max(np.abs(r64n-r64))/s, dopp[c], step[c]))
# v1 on the same commands (few at a time -> chunked kernel)
sub = eng.make_cmds(bad, offs[bad] + first, lens[bad], shifts[bad], remc[bad], step[bad], remcode[bad], cstep[bad])
g1 = eng.correlate(sub); print(eng.last_kernel)
for i, c in enumerate(bad):
    print(c, "v1-ref32 %.2e  v1-v2 %.2e" % (np.max(np.abs(g1[i]-ref[c]))/np.max(np.abs(ref[c])), np.max(np.abs(g1[i]-got[c]))/np.max(np.abs(ref[c]))))
