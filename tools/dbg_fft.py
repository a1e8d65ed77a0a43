import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from gnss_sim_receiver_b200.acquisition import AcquisitionEngine
n = int(sys.argv[1]) if len(sys.argv) > 1 else 8
rng = np.random.default_rng(n)
x = (rng.standard_normal(n) + 1j * rng.standard_normal(n)).astype(np.complex64)
eng = AcquisitionEngine(fs_in=4e6, fft_size=n, max_prns=1)
X = eng.fft(x)
ref = np.fft.fft(x.astype(np.complex128))
print("fwd err", np.linalg.norm(X - ref) / np.linalg.norm(ref))
if n <= 20:
    print(np.round(X, 3)); print(np.round(ref, 3))
y = eng.fft(ref.astype(np.complex64), inverse=True)
print("inv err", np.linalg.norm(y - n * x) / (n * np.linalg.norm(x)))
