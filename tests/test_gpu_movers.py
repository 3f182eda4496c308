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
