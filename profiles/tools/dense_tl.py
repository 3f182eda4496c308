import os, sys, ctypes as C
os.environ["MCMCPP_B200_DENSE_TIMELINE"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
import mcmcpp_b200 as m
from mcmcpp_b200 import _capi
D, W = 100, 65536
idx = np.arange(D)
P = np.linalg.inv(0.9 ** np.abs(idx[:, None] - idx[None, :]))
init = np.random.default_rng(1).standard_normal((W, D))
lp = -0.5 * np.einsum("wi,ij,wj->w", init, P, init)
s = m.EnsembleSampler(0, W, D, m.StretchMove(D, 0, m.DenseGaussian(P)), maxChainSizeBytes=0)
s.setInitialWalkerPos(init, lp)
s.runMCMC(20)
buf = (C.c_longlong * 256)()
lib = C.CDLL(os.path.join(os.path.dirname(_capi.__file__), "libmcmcpp_b200.so"))
print("rc", lib.mcmcpp_dense_timeline(buf))
t = np.array(buf[:]).reshape(2, 16, 8)
t0 = t[t > 0].min()
for g in range(2):
    for k in range(16):
        if t[g, k, 0] > 0:
            r = (t[g, k, :8] - t0) / 1.965e3
            print(f"group {g} tile {k}: load_start {r[0]:7.2f} loads_done {r[1]:7.2f} token {r[2]:7.2f} mma_begin {r[6]:7.2f} mma_end {r[7]:7.2f} decided {r[3]:7.2f} sync {r[4]:7.2f} commit_done {r[5]:7.2f}  (us)")
