"""GPU (-m gpu): Mover::DifferentialEvolution on the device (include/mcmcpp_b200/diff_evo_kernels.cuh) against the
restatement oracle_de_run (pinned bit-exact to the unmodified reference mover by tests/test_movers_cpu.py), one pcg64
stream per walker.  Bar: positions, log-probs, both partner indices and accept bits bit-exact; log(1-u) within 4 ulp."""
import numpy as np
import pytest
from test_gpu_parity import CALC, ulp_close

pytestmark = pytest.mark.gpu


def run_both(port, kind, params, W, D, steps, seed, interval=1, gamma=None, rand=None, trace=0):
    import mcmcpp_b200 as m
    params = np.asarray(params, dtype=np.float64)
    init = np.random.default_rng(W + D).standard_normal((W, D))
    lp = port.logp(kind, params if params.size else None, init)
    mover = m.DifferentialEvolution(D, seed, CALC[kind](m, params, D))
    if gamma is not None:
        mover.overrideGamma(gamma)
    if rand is not None:
        mover.overrideRandBounds(rand)
    s = m.EnsembleSampler(seed, W, D, mover, traceUpdates=trace)
    s.setSlicingMode(interval > 1, interval)
    s.setInitialWalkerPos(init, lp)
    assert s.runMCMC(steps)
    kw = {}
    if gamma is not None:
        kw["gamma"] = gamma
    if rand is not None:
        kw["rand"] = rand
    o = port.run_de(kind, params if params.size else None, init, lp, steps, interval=interval, seed=seed, per_walker=True,
                    want_trace=trace > 0, **kw)
    return s, o


@pytest.mark.parametrize("kind,W,D,steps,interval", [("skewed_gauss_2d", 16, 2, 200, 1), ("skewed_gauss_2d", 4096, 2, 30, 2),
                                                     ("rosenbrock_2d", 1024, 2, 40, 1), ("diag_gauss", 64, 5, 50, 1),
                                                     ("diag_gauss", 256, 8, 20, 1)])
def test_diff_evo_thread_kernel_bit_exact(port, kind, W, D, steps, interval):
    params = {"skewed_gauss_2d": [0.13], "rosenbrock_2d": [], "diag_gauss": 1.0 / (1.0 + (np.arange(D) % 10))}[kind]
    s, o = run_both(port, kind, params, W, D, steps, seed=3, interval=interval, trace=W * 2)
    assert np.array_equal(s.getChain()[1:], o["chain"])
    pos, aux = s.getWalkerState()
    assert np.array_equal(pos, o["pos"]) and np.array_equal(aux, o["logp"])
    assert (s.getAcceptedSteps(), s.getTotalSteps()) == (int(o["acc"].sum()), int(o["tot"].sum()))
    tr, otr = s.getTrace(W * 2), o["trace"][:W * 2]
    assert np.array_equal(tr["j"], otr["j"]) and np.array_equal(tr["z"], otr["z"]) and np.array_equal(tr["accept"], otr["accept"])
    assert np.array_equal(tr["newp"], otr["newp"]) and ulp_close(tr["e"], otr["e"])
    s.close()


@pytest.mark.parametrize("W,D,steps", [(64, 9, 20), (256, 64, 10), (256, 65, 10), (512, 100, 8), (512, 127, 6), (2100, 1000, 3),
                                       (2100, 1001, 2)])
def test_diff_evo_warp_kernel_lane_parallel_draws_bit_exact(port, W, D, steps):
    """one warp per walker: every lane jumps its own copy of the walker's engine to its two uniforms of each 64-element
    chunk; the separable log-posterior is reduced in the canonical order -> everything bit-exact, ragged tails included."""
    iv = 1.0 / (1.0 + (np.arange(D) % 10))
    s, o = run_both(port, "diag_gauss", iv, W, D, steps, seed=5, trace=W)
    assert np.array_equal(s.getChain()[1:], o["chain"])
    pos, aux = s.getWalkerState()
    assert np.array_equal(pos, o["pos"]) and np.array_equal(aux, o["logp"])
    assert s.getAcceptedSteps() == int(o["acc"].sum())
    tr, otr = s.getTrace(W), o["trace"][:W]
    assert np.array_equal(tr["j"], otr["j"]) and np.array_equal(tr["z"], otr["z"]) and np.array_equal(tr["accept"], otr["accept"])
    s.close()


def test_diff_evo_generic_calculator_and_tuning_overrides(port, monkeypatch):
    """non-separable Calculator (dense Gaussian, reference-order evaluation by lane 0) with overrideGamma / overrideRandBounds."""
    D, W = 24, 128
    idx = np.arange(D)
    P = np.linalg.inv(0.9 ** np.abs(idx[:, None] - idx[None, :])).reshape(-1)
    s, o = run_both(port, "dense_gauss", P, W, D, 12, seed=7, gamma=0.4, rand=3.0e-3)
    assert np.array_equal(s.getChain()[1:], o["chain"]) and s.getAcceptedSteps() == int(o["acc"].sum())
    s.close()


def test_diff_evo_needs_per_walker_streams():
    import ctypes as C
    from mcmcpp_b200 import _capi
    cfg = _capi.Config(num_walkers=64, num_params=2, seed=0, rng_mode=_capi.RNG_SEQUENTIAL, alpha_num=2, alpha_denom=1, device=0,
                       rank=0, world=1, shard_mode=_capi.SHARD_NONE, trace_updates=0, max_chain_bytes=0)
    h = C.c_void_p()
    p = np.array([0.13])
    rc = _capi.lib().mcmcpp_sampler_create_builtin_mover(C.byref(cfg), _capi.MOVER_DIFF_EVOLUTION, None, 0, _capi.TARGET["skewed_gauss_2d"],
                                                         p.ctypes.data_as(C.POINTER(C.c_double)), 1, C.byref(h))
    assert rc == _capi.ERR_UNSUPPORTED


def run_walk_both(port, kind, params, W, D, S, steps, seed, trace=0):
    import mcmcpp_b200 as m
    params = np.asarray(params, dtype=np.float64)
    init = np.random.default_rng(W + D + S).standard_normal((W, D))
    lp = port.logp(kind, params if params.size else None, init)
    s = m.EnsembleSampler(seed, W, D, m.WalkMove(D, seed, CALC[kind](m, params, D), S), traceUpdates=trace)
    s.setInitialWalkerPos(init, lp)
    assert s.runMCMC(steps)
    o = port.run_walk(kind, params if params.size else None, init, lp, steps, S, seed=seed, per_walker=True, want_trace=trace > 0)
    return s, o


@pytest.mark.parametrize("kind,W,D,S,steps", [("skewed_gauss_2d", 16, 2, 3, 60), ("skewed_gauss_2d", 32, 2, 16, 30),
                                              ("rosenbrock_2d", 1024, 2, 4, 20), ("diag_gauss", 64, 5, 6, 30),
                                              ("diag_gauss", 512, 100, 8, 6), ("diag_gauss", 2100, 1001, 5, 2),
                                              ("dense_gauss", 128, 24, 7, 8)])
def test_walk_move_tolerance_tier(port, kind, W, D, S, steps):
    """Mover::WalkMove on the device against oracle_walk_run (pinned bit-exact to the unmodified reference mover), one
    engine per walker.  Partner selection is integer work: bit-exact.  The normals go through log(): the device's and
    glibc's differ in the last bit on some inputs, so normals / positions / log-probs carry the tolerance tier
    (<= 1e-12 relative, north_star); decisions must agree over these runs (a flip needs one within ~1e-13 of its threshold).
    Covers selection of ALL walkers of the other colour (S == K), odd and wide rows, separable and generic Calculators."""
    params = {"skewed_gauss_2d": [0.13], "rosenbrock_2d": [], "diag_gauss": 1.0 / (1.0 + (np.arange(D) % 10)),
              "dense_gauss": np.linalg.inv(0.9 ** np.abs(np.arange(D)[:, None] - np.arange(D)[None, :])).reshape(-1)}[kind]
    s, o = run_walk_both(port, kind, params, W, D, S, steps, seed=4, trace=W)
    tr, otr = s.getTrace(W), o["trace"][:W]
    assert np.array_equal(tr["j"], otr["j"]) and np.array_equal(tr["accept"], otr["accept"])
    assert np.allclose(tr["z"], otr["z"], rtol=1e-13, atol=0)                      # first normal of every update
    pos, aux = s.getWalkerState()
    scale = np.abs(o["pos"]) + np.abs(o["pos"]).max()
    assert np.all(np.abs(pos - o["pos"]) <= 1e-12 * scale)
    assert np.allclose(aux, o["logp"], rtol=1e-11, atol=1e-12)
    assert (s.getAcceptedSteps(), s.getTotalSteps()) == (int(o["acc"].sum()), int(o["tot"].sum()))
    s.close()


def test_walk_move_checkpoint_carries_the_normal_distribution_state(port):
    """the saved second normal of a pair is mover state: a restored sampler must continue bit for bit"""
    import mcmcpp_b200 as m
    W, D, S = 256, 3, 5                       # odd S: every update leaves or consumes a saved value
    iv = 1.0 / (1.0 + np.arange(D))
    init = np.random.default_rng(1).standard_normal((W, D))
    lp = port.logp("diag_gauss", iv, init)
    mk = lambda: m.EnsembleSampler(9, W, D, m.WalkMove(D, 9, m.DiagGaussian(iv), S))
    a = mk(); a.setInitialWalkerPos(init, lp); assert a.runMCMC(12)
    b = mk(); b.setInitialWalkerPos(init, lp); assert b.runMCMC(5)
    blob = b.saveCheckpoint(); b.close()
    c = mk(); c.loadCheckpoint(blob); assert c.runMCMC(7)
    pa, la = a.getWalkerState(); pc, lc = c.getWalkerState()
    assert np.array_equal(pa, pc) and np.array_equal(la, lc)
    a.close(); c.close()


# ---- small ensembles: the movers' whole runMCMC call in one launch (diffEvoFusedSmall, walkMoveFusedSmall) ---------------------
@pytest.mark.parametrize("kind,W,D,calls,interval", [("skewed_gauss_2d", 16, 2, (120, 7), 1), ("skewed_gauss_2d", 320, 2, (40, 3), 3),
                                                     ("rosenbrock_2d", 2048, 2, (12, 5), 1), ("diag_gauss", 64, 5, (30, 30), 2),
                                                     ("dense_gauss", 40, 8, (25, 2), 1)])
def test_diff_evo_fused_small_ensemble(port, kind, W, D, calls, interval):
    """One launch per runMCMC call (no trace requested) against the launch-per-colour path (a trace request selects it) and the
    oracle: chains, final state, counters bit-exact; the second call continues every walker's stream."""
    import mcmcpp_b200 as m
    params = np.asarray({"skewed_gauss_2d": [0.13], "rosenbrock_2d": [], "diag_gauss": 1.0 / (1.0 + (np.arange(D) % 10)),
                         "dense_gauss": np.linalg.inv(0.9 ** np.abs(np.arange(D)[:, None] - np.arange(D)[None, :])).reshape(-1)}[kind])
    init = np.random.default_rng(W + D).standard_normal((W, D))
    lp = port.logp(kind, params if params.size else None, init)
    runs = []
    for trace in (0, 1):
        s = m.EnsembleSampler(11, W, D, m.DifferentialEvolution(D, 11, CALC[kind](m, params, D)), traceUpdates=trace)
        s.setSlicingMode(interval > 1, interval)
        s.setInitialWalkerPos(init, lp)
        for n in calls:
            assert s.runMCMC(n)
        assert s.getLaunchCount() == (len(calls) if trace == 0 else 2 * sum(calls) * interval)
        runs.append((np.array(s.getChain()), s.getWalkerState(), s.getAcceptedSteps(), s.getTotalSteps()))
        s.close()
    (c0, (p0, a0), acc0, tot0), (c1, (p1, a1), acc1, tot1) = runs
    assert np.array_equal(c0, c1) and np.array_equal(p0, p1) and np.array_equal(a0, a1) and (acc0, tot0) == (acc1, tot1)
    o = port.run_de(kind, params if params.size else None, init, lp, sum(calls), interval=interval, seed=11, per_walker=True)
    assert np.array_equal(c0[1:], o["chain"]) and np.array_equal(p0, o["pos"]) and np.array_equal(a0, o["logp"])
    assert (acc0, tot0) == (int(o["acc"].sum()), int(o["tot"].sum()))


@pytest.mark.parametrize("kind,W,D,S,calls", [("skewed_gauss_2d", 16, 2, 3, (50, 9)), ("skewed_gauss_2d", 32, 2, 16, (20, 11)),
                                              ("skewed_gauss_2d", 320, 2, 5, (15, 4)), ("rosenbrock_2d", 2048, 2, 4, (6, 3)),
                                              ("diag_gauss", 64, 5, 7, (20, 5)), ("dense_gauss", 40, 8, 64 // 4, (10, 3))])
def test_walk_move_fused_small_ensemble(port, kind, W, D, S, calls):
    """The same for Mover::WalkMove (odd numbers of points: the normal distribution's saved value crosses updates and calls):
    one launch per call, bit-identical to the launch-per-colour path; against the oracle in the mover's tolerance tier."""
    import mcmcpp_b200 as m
    params = np.asarray({"skewed_gauss_2d": [0.13], "rosenbrock_2d": [], "diag_gauss": 1.0 / (1.0 + (np.arange(D) % 10)),
                         "dense_gauss": np.linalg.inv(0.9 ** np.abs(np.arange(D)[:, None] - np.arange(D)[None, :])).reshape(-1)}[kind])
    init = np.random.default_rng(W + D + S).standard_normal((W, D))
    lp = port.logp(kind, params if params.size else None, init)
    runs = []
    for trace in (0, 1):
        s = m.EnsembleSampler(6, W, D, m.WalkMove(D, 6, CALC[kind](m, params, D), S), traceUpdates=trace)
        s.setInitialWalkerPos(init, lp)
        for n in calls:
            assert s.runMCMC(n)
        assert s.getLaunchCount() == (len(calls) if trace == 0 else 2 * sum(calls))
        runs.append((np.array(s.getChain()), s.getWalkerState(), s.getAcceptedSteps(), s.getTotalSteps()))
        s.close()
    (c0, (p0, a0), acc0, tot0), (c1, (p1, a1), acc1, tot1) = runs
    assert np.array_equal(c0, c1) and np.array_equal(p0, p1) and np.array_equal(a0, a1) and (acc0, tot0) == (acc1, tot1)
    o = port.run_walk(kind, params if params.size else None, init, lp, sum(calls), S, seed=6, per_walker=True)
    scale = np.abs(o["pos"]) + np.abs(o["pos"]).max()
    assert np.all(np.abs(p0 - o["pos"]) <= 1e-12 * scale) and np.allclose(a0, o["logp"], rtol=1e-11, atol=1e-12)
    assert (acc0, tot0) == (int(o["acc"].sum()), int(o["tot"].sum()))
