"""Small invocations of every kernel family for compute-sanitizer (memcheck / racecheck), round 2:
  compute-sanitizer --tool memcheck  python profiles/tools/sanitize_small.py
Each case is also checked against the oracle, so a sanitizer-clean run that computes garbage would still fail."""
import os
import sys
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import mcmcpp_b200 as m                      # noqa: E402
from mcmcpp_b200 import parallel             # noqa: E402
from oracle.pyoracle import Port             # noqa: E402

port = Port()
rng = np.random.default_rng(0)


def run(kind, calc, params, W, D, steps, **kw):
    init = rng.standard_normal((W, D))
    lp = port.logp(kind, params, init)
    s = m.EnsembleSampler(1, W, D, m.StretchMove(D, 1, calc), **kw)
    s.setInitialWalkerPos(init, lp)
    assert s.runMCMC(steps)
    pw = {"per_walker": True, "per_walker32": 32}.get(kw.get("rngMode", "sequential"), False)
    o = port.run(kind, params, init, lp, steps, seed=1, per_walker=pw)
    pos, aux = s.getWalkerState()
    ok = np.array_equal(pos, o["pos"]) and np.allclose(aux, o["logp"], rtol=1e-12, atol=0)
    print("%-16s W=%-5d D=%-4d %s: %s" % (kind, W, D, kw, "ok" if ok else "MISMATCH"), flush=True)
    s.close()
    return ok


iv = lambda D: 1.0 / (1.0 + (np.arange(D) % 10))
good = True
good &= run("skewed_gauss_2d", m.SkewedGaussianTwoDim(0.13), [0.13], 32, 2, 50)                      # stretchFusedSmall
os.environ["MCMCPP_B200_FUSED"] = "0"
good &= run("rosenbrock_2d", m.RosenbrockTwoDim(), None, 4096, 2, 5)                                # stretchHalfStepThread, PDL
good &= run("rosenbrock_2d", m.RosenbrockTwoDim(), None, 4096, 2, 5, rngMode="per_walker32")
good &= run("diag_gauss", m.DiagGaussian(iv(37)), iv(37), 256, 37, 4)                                # stretchHalfStepWarp
good &= run("diag_gauss", m.DiagGaussian(iv(1000)), iv(1000), 2048, 1000, 2, maxChainSizeBytes=0)    # stretchDrawKernel + stretchHalfStepBulk (TMA)
good &= run("diag_gauss", m.DiagGaussian(iv(200)), iv(200), 512, 200, 3, deviceChainBytes=1 << 24)   # device-resident chain
P = np.linalg.inv(0.9 ** np.abs(np.subtract.outer(np.arange(24), np.arange(24))))
good &= run("dense_gauss", m.DenseGaussian(P), P.reshape(-1), 256, 24, 4)                            # denseDrawKernel + denseGaussHalfStep (DMMA, token barriers)
good &= run("diag_gauss", m.DiagGaussian(iv(126)), iv(126), 2048, 126, 3, shardMode="dim")           # dimShardDraw + dimShardFused, world 1
# world 2 and 3 in one launch (slot protocol between ranks)
for world, W, D in ((2, 1024, 250), (3, 384, 37)):
    init = rng.standard_normal((W, D))
    lp = port.logp("diag_gauss", iv(D), init)
    o = port.run("diag_gauss", iv(D), init, lp, 3, seed=6)
    ranks = [m.EnsembleSampler(6, W, D, m.StretchMove(D, 6, m.DiagGaussian(iv(D))), rank=r, world=world, shardMode="dim") for r in range(world)]
    g = parallel.RankGroup(ranks)
    for s in ranks:
        s.setInitialWalkerPos(init, lp)
    assert g.runMCMC(3)
    ok = True
    for r, s in enumerate(ranks):
        b, e = parallel.dim_slice(D, r, world)
        pos, aux = s.getWalkerState()
        ok &= np.array_equal(pos[:, b:e], o["pos"][:, b:e]) and np.allclose(aux, o["logp"], rtol=1e-12, atol=0)
        s.close()
    print("rank group world=%d W=%d D=%d: %s" % (world, W, D, "ok" if ok else "MISMATCH"), flush=True)
    good &= ok
print("SANITIZE_SMALL_OK" if good else "SANITIZE_SMALL_FAILED", flush=True)
sys.exit(0 if good else 1)
