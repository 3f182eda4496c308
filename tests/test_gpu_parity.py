"""GPU parity tests (-m gpu): the CUDA path, called through the Python mirror -> C-ABI, against the
CPU oracle (oracle/oracle.c, pinned by tests/test_oracle_cpu.py) and the golden fixtures made from
the unmodified reference.  Bar: bit-exact positions, log-probs, partner indices and accept bits
(tier A); log(1-u) and the Metropolis difference within 4 ulp (device log vs glibc
log differ in the last bit on a few inputs; neither enters the stored state)."""
import numpy as np
import pytest
from conftest import load_golden, golden_names

pytestmark = pytest.mark.gpu

CALC = {
    "skewed_gauss_2d": lambda m, p, D: m.SkewedGaussianTwoDim(p[0]),
    "rosenbrock_2d": lambda m, p, D: m.RosenbrockTwoDim(),
    "diag_gauss": lambda m, p, D: m.DiagGaussian(p),
    "dense_gauss": lambda m, p, D: m.DenseGaussian(p.reshape(D, D)),
    "linreg": lambda m, p, D: m.LinearRegression(p[2:2 + int(p[0]) * D].reshape(int(p[0]), D), p[2 + int(p[0]) * D:], p[1]),
}


def make_sampler(kind, params, W, D, seed=0, **kw):
    import mcmcpp_b200 as m
    calc = CALC[kind](m, np.asarray(params, dtype=np.float64), D)
    return m.EnsembleSampler(seed, W, D, m.StretchMove(D, seed, calc), **kw)


def ulp_close(a, b, ulps=4):
    a = np.asarray(a); b = np.asarray(b)
    return np.all(np.abs(a - b) <= ulps * np.spacing(np.maximum(np.abs(a), np.abs(b))))


def assert_trace_equal(dev, ref):
    assert np.array_equal(dev["j"], ref["j"])
    assert np.array_equal(dev["accept"], ref["accept"])
    assert np.array_equal(dev["z"], ref["z"])
    assert np.array_equal(dev["newp"], ref["newp"])
    assert ulp_close(dev["e"], ref["e"])
    # diff = log(z)(D-1) + newp - logp cancels: bound it by the size of its terms, not of the result
    assert np.all(np.abs(dev["diff"] - ref["diff"]) <= 1e-14 * (1.0 + np.abs(ref["newp"]) + np.abs(ref["diff"])))


@pytest.mark.parametrize("name", golden_names())
def test_golden_fixtures_bit_exact(name):
    """Every fixture generated from the unmodified reference: chain rows, counters."""
    g = load_golden(name)
    s = make_sampler(g["kind"], g["params"], g["W"], g["D"], seed=g["seed"])
    if g["interval"] > 1:
        s.setSlicingMode(True, g["interval"])
    s.setInitialWalkerPos(g["init"], g["logp"])
    assert s.runMCMC(g["nsteps"]) is True
    chain = s.getChain()
    assert chain.shape == (g["nsteps"] + 1, g["W"], g["D"]) and s.getStoredSteps() == g["stored"]
    assert np.array_equal(chain[g["rows"]], g["chain"])
    assert (s.getAcceptedSteps(), s.getTotalSteps()) == (g["acc"], g["tot"])
    assert s.getRngRedraws() == 0
    s.close()


def test_c1_full_chain_and_trace_against_oracle(port):
    """BASELINE config 1: W=32, D=2, 10,000 stored steps; the whole chain and the first 64,000 updates."""
    g = load_golden("c1_skewed_w32")
    ntr = 64000
    s = make_sampler(g["kind"], g["params"], g["W"], g["D"], traceUpdates=ntr)
    s.setInitialWalkerPos(g["init"], g["logp"])
    assert s.runMCMC(g["nsteps"])
    o = port.run(g["kind"], g["params"], g["init"], g["logp"], g["nsteps"], want_trace=True)
    assert np.array_equal(s.getChain()[1:], o["chain"])
    pos, aux = s.getWalkerState()
    assert np.array_equal(pos, o["pos"]) and np.array_equal(aux, o["logp"])
    assert s.getAcceptedSteps() == 228883 and s.getTotalSteps() == 320032
    assert abs(s.getAcceptanceFraction() - 228883 / 320032) < 1e-15
    assert_trace_equal(s.getTrace(ntr), o["trace"][:ntr])
    assert [float.hex(v) for v in s.getChain()[-1, 0]] == ["0x1.754c17069f6f6p+0", "0x1.22e26142efda4p+0"]
    s.close()


def seeded_init(port, kind, params, W, D, seed, scale=1.0):
    rng = np.random.default_rng(seed)
    init = rng.standard_normal((W, D)) * scale
    return init, port.logp(kind, params, init)


@pytest.mark.parametrize("W,D,steps,interval", [(4096, 2, 60, 1), (1 << 16, 2, 8, 2)])
def test_rosenbrock_thread_kernel(port, W, D, steps, interval):
    init, lp = seeded_init(port, "rosenbrock_2d", None, W, D, 11)
    s = make_sampler("rosenbrock_2d", [], W, D, seed=3, traceUpdates=W * 2)
    s.setSlicingMode(interval > 1, interval)
    s.setInitialWalkerPos(init, lp)
    assert s.runMCMC(steps)
    o = port.run("rosenbrock_2d", None, init, lp, steps, interval=interval, seed=3, want_trace=True)
    assert np.array_equal(s.getChain()[1:], o["chain"])
    assert s.getAcceptedSteps() == int(o["acc"].sum()) and s.getTotalSteps() == int(o["tot"].sum())
    assert_trace_equal(s.getTrace(W * 2), o["trace"][:W * 2])
    s.close()


@pytest.mark.parametrize("W,D,steps", [(64, 3, 20), (64, 8, 20), (256, 9, 10), (256, 64, 10), (256, 65, 10), (256, 127, 6),
                                       (512, 200, 6), (2048, 500, 3), (4096, 1000, 3), (4096, 1023, 2), (4096, 1024, 2),
                                       (2100, 1025, 2), (2400, 1191, 2), (4200, 2049, 1)])
def test_diag_gauss_every_kernel_shape(port, W, D, steps):
    """thread kernel (D<=8), warp kernel with 1..16 chunks, odd D (scalar loads) and ragged tails, odd D past 1024 (rows staged in
    shared memory by plain loads):
    the separable log-posterior is reduced in the canonical order, so everything is bit-exact."""
    iv = 1.0 / (1.0 + (np.arange(D) % 10))
    init, lp = seeded_init(port, "diag_gauss", iv, W, D, 100 + D)
    s = make_sampler("diag_gauss", iv, W, D, seed=1, traceUpdates=W)
    s.setInitialWalkerPos(init, lp)
    assert s.runMCMC(steps)
    o = port.run("diag_gauss", iv, init, lp, steps, seed=1, want_trace=True)
    assert np.array_equal(s.getChain()[1:], o["chain"])
    pos, aux = s.getWalkerState()
    assert np.array_equal(pos, o["pos"]) and np.array_equal(aux, o["logp"])
    assert s.getAcceptedSteps() == int(o["acc"].sum())
    assert_trace_equal(s.getTrace(W), o["trace"][:W])
    s.close()


def test_generic_calculator_in_warp_kernel(port, monkeypatch):
    """non-separable Calculator, D > 8: proposal staged in shared memory, evaluated in the reference's order."""
    monkeypatch.setenv("MCMCPP_B200_DENSE_GENERIC", "1")       # not the tensor-core specialisation
    D, W = 24, 128
    idx = np.arange(D)
    P = np.linalg.inv(0.9 ** np.abs(idx[:, None] - idx[None, :]))
    init, lp = seeded_init(port, "dense_gauss", P.reshape(-1), W, D, 5)
    s = make_sampler("dense_gauss", P.reshape(-1), W, D, seed=2)
    s.setInitialWalkerPos(init, lp)
    assert s.runMCMC(15)
    o = port.run("dense_gauss", P.reshape(-1), init, lp, 15, seed=2)
    assert np.array_equal(s.getChain()[1:], o["chain"]) and s.getAcceptedSteps() == int(o["acc"].sum())
    s.close()


def test_per_walker_rng_mode(port):
    W, D = 2048, 2
    init, lp = seeded_init(port, "skewed_gauss_2d", [0.13], W, D, 21)
    s = make_sampler("skewed_gauss_2d", [0.13], W, D, seed=77, rngMode="per_walker", traceUpdates=W * 3)
    s.setInitialWalkerPos(init, lp)
    assert s.runMCMC(25)
    o = port.run("skewed_gauss_2d", [0.13], init, lp, 25, seed=77, per_walker=True, want_trace=True)
    assert np.array_equal(s.getChain()[1:], o["chain"]) and s.getAcceptedSteps() == int(o["acc"].sum())
    assert_trace_equal(s.getTrace(W * 3), o["trace"][:W * 3])
    s.close()


def test_alpha_other_than_two(port):
    import mcmcpp_b200 as m
    W, D = 512, 2
    init, lp = seeded_init(port, "skewed_gauss_2d", [0.13], W, D, 8)
    s = m.EnsembleSampler(4, W, D, m.StretchMove(D, 4, m.SkewedGaussianTwoDim(0.13), alphaNum=3, alphaDenom=2))
    s.setInitialWalkerPos(init, lp)
    assert s.runMCMC(30)
    o = port.run("skewed_gauss_2d", [0.13], init, lp, 30, seed=4, alpha=(3, 2))
    assert np.array_equal(s.getChain()[1:], o["chain"])
    s.close()


def test_successive_runs_continue_the_stream(port):
    g = load_golden("c1_skewed_w32")
    s = make_sampler(g["kind"], g["params"], g["W"], g["D"])
    s.setInitialWalkerPos(g["init"], g["logp"])
    for n in (7, 1, 92):
        assert s.runMCMC(n)
    o = port.run(g["kind"], g["params"], g["init"], g["logp"], 100)
    assert np.array_equal(s.getChain()[1:], o["chain"])
    s.close()


def corr_precision(D, rho=0.9):
    idx = np.arange(D)
    return np.linalg.inv(rho ** np.abs(idx[:, None] - idx[None, :]))


@pytest.mark.parametrize("W,D,steps", [(4096, 100, 12), (64, 12, 30), (256, 24, 20), (512, 101, 8), (1024, 128, 6), (130, 64, 10)])
def test_dense_gauss_tensor_core_tier_b(port, W, D, steps):
    """BASELINE config 3 path (FP64 DMMA quadratic form).  Draws, partners, proposals: bit-exact.  log-posterior:
    tensor-core summation order -> tolerance 1e-12 relative (north_star); decisions and therefore positions must
    still agree exactly over these runs (a flip needs a decision within ~1e-13 of its threshold)."""
    P = corr_precision(D).reshape(-1)
    init, lp = seeded_init(port, "dense_gauss", P, W, D, 40 + D)
    s = make_sampler("dense_gauss", P, W, D, seed=6, traceUpdates=W * 2)
    s.setInitialWalkerPos(init, lp)
    assert s.runMCMC(steps)
    o = port.run("dense_gauss", P, init, lp, steps, seed=6, want_trace=True)
    tr, otr = s.getTrace(W * 2), o["trace"][:W * 2]
    assert np.array_equal(tr["j"], otr["j"]) and np.array_equal(tr["z"], otr["z"])
    assert np.allclose(tr["newp"], otr["newp"], rtol=1e-12, atol=0)
    assert np.array_equal(tr["accept"], otr["accept"])
    pos, aux = s.getWalkerState()
    assert np.array_equal(pos, o["pos"])
    assert np.allclose(aux, o["logp"], rtol=1e-12, atol=0)
    assert np.array_equal(s.getChain()[1:], o["chain"])
    assert s.getAcceptedSteps() == int(o["acc"].sum())
    s.close()


def test_c3_full_size_dense_gauss(port):
    """BASELINE config 3 at full size: D=100, 65,536 walkers, 2 steps against the oracle."""
    W, D = 1 << 16, 100
    P = corr_precision(D).reshape(-1)
    init, lp = seeded_init(port, "dense_gauss", P, W, D, 3)
    s = make_sampler("dense_gauss", P, W, D, seed=0, maxChainSizeBytes=0)
    s.setInitialWalkerPos(init, lp)
    assert s.runMCMC(2)
    o = port.run("dense_gauss", P, init, lp, 2, seed=0, want_chain=False)
    pos, aux = s.getWalkerState()
    assert np.array_equal(pos, o["pos"]) and np.allclose(aux, o["logp"], rtol=1e-12, atol=0)
    assert s.getAcceptedSteps() == int(o["acc"].sum())
    s.close()


def linreg_problem(M, P, seed):
    rng = np.random.default_rng(seed)
    X = rng.standard_normal((M, P))
    theta = rng.standard_normal(P)
    y = X @ theta + rng.standard_normal(M)
    ols = np.linalg.lstsq(X, y, rcond=None)[0]
    return X, y, ols, np.concatenate([[float(M), 100.0], X.reshape(-1), y])


@pytest.mark.parametrize("W,P,M,steps", [(64, 3, 500, 20), (256, 50, 3000, 6), (1024, 20, 10000, 4), (600, 63, 2000, 4),
                                         (2048, 50, 20000, 2), (16, 3, 64, 30)])
def test_linreg_tensor_core_gemm_tier_b(port, W, P, M, steps):
    """BASELINE config 5 path (likelihood of all proposals of a colour as one FP64 DMMA GEMM).  Draws, partners and
    proposals bit-exact; log-posterior summed in tensor-core order: 1e-12 relative (north_star); decisions, and
    therefore positions and the chain, must still agree exactly over these runs.  Shapes cover every k-step
    instantiation (4, 8, 13, 16), ragged data tiles (M % 64 != 0) and ragged / multiple proposal tiles."""
    X, y, ols, params = linreg_problem(M, P, 7 + P)
    rng = np.random.default_rng(P)
    init = ols + 0.1 * rng.standard_normal((W, P))
    lp = port.logp("linreg", params, init)
    s = make_sampler("linreg", params, W, P, seed=9, traceUpdates=W * 2)
    s.setInitialWalkerPos(init, lp)
    assert s.runMCMC(steps)
    assert s.getLaunchCount() == 8 * steps            # propose, GEMM, reduce, accept per colour
    o = port.run("linreg", params, init, lp, steps, seed=9, want_trace=True)
    tr, otr = s.getTrace(W * 2), o["trace"][:W * 2]
    assert np.array_equal(tr["j"], otr["j"]) and np.array_equal(tr["z"], otr["z"])
    assert np.allclose(tr["newp"], otr["newp"], rtol=1e-12, atol=0)
    assert np.array_equal(tr["accept"], otr["accept"])
    pos, aux = s.getWalkerState()
    assert np.array_equal(pos, o["pos"])
    assert np.allclose(aux, o["logp"], rtol=1e-12, atol=0)
    assert np.array_equal(s.getChain()[1:], o["chain"])
    assert s.getAcceptedSteps() == int(o["acc"].sum())
    s.close()


def test_c5_full_size_linreg_against_sufficient_statistics():
    """BASELINE config 5 at full size (P=50, M=1e6, W=16,384): the oracle needs 1e8 flops per update, so the check is a
    size-independent property: after 2 steps every walker's stored log-posterior must equal the closed form
    -1/2 (y'y - 2 th'X'y + th'X'X th) - 1/2 |th|^2/100 evaluated from the sufficient statistics (a different
    algorithm; 1e-9 relative covers its own cancellation)."""
    M, P, W = 1000000, 50, 16384
    X, y, ols, params = linreg_problem(M, P, 1234)
    XtX, Xty, yty = X.T @ X, X.T @ y, float(y @ y)

    def closed(th):
        return -0.5 * (yty - 2.0 * th @ Xty + np.einsum("wi,ij,wj->w", th, XtX, th)) - 0.5 * np.sum(th * th, axis=1) / 100.0

    rng = np.random.default_rng(5)
    init = ols + 0.1 * rng.standard_normal((W, P))
    s = make_sampler("linreg", params, W, P, seed=0, maxChainSizeBytes=0)
    s.setInitialWalkerPos(init, closed(init))
    assert s.runMCMC(2)
    pos, aux = s.getWalkerState()
    moved = np.any(pos != init, axis=1)
    assert 0.05 < moved.mean() < 0.95                      # acceptance of a 50-D stretch move
    assert np.allclose(aux[moved], closed(pos[moved]), rtol=1e-9, atol=0)
    assert W + int(moved.sum()) <= s.getAcceptedSteps() <= 3 * W
    s.close()


@pytest.mark.parametrize("W,D,steps", [(256, 9, 8), (512, 100, 6), (512, 125, 6), (2048, 500, 3), (2048, 1000, 3), (2100, 1001, 2)])
def test_dim_shard_kernels_single_rank_bit_exact(port, W, D, steps):
    """The DIM-sharded half-step (draws; fused persistent kernel: partial sums -> slots of the table -> rank-ordered sum ->
    commit from shared memory) with world = 1 is the single-GPU summation order: bit-exact against the oracle, chain and trace
    included (even slices: TMA ring; odd slices: plain loads)."""
    iv = 1.0 / (1.0 + (np.arange(D) % 10))
    init, lp = seeded_init(port, "diag_gauss", iv, W, D, 200 + D)
    s = make_sampler("diag_gauss", iv, W, D, seed=1, traceUpdates=W, shardMode="dim")
    s.setInitialWalkerPos(init, lp)
    assert s.runMCMC(steps)
    assert s.getLaunchCount() == 6 * steps                            # draw, summing and finishing kernel per colour
    o = port.run("diag_gauss", iv, init, lp, steps, seed=1, want_trace=True)
    assert np.array_equal(s.getChain()[1:], o["chain"])
    pos, aux = s.getWalkerState()
    assert np.array_equal(pos, o["pos"]) and np.array_equal(aux, o["logp"])
    assert s.getAcceptedSteps() == int(o["acc"].sum())
    assert_trace_equal(s.getTrace(W), o["trace"][:W])
    s.close()


@pytest.mark.parametrize("kind,W,D,steps,interval,rng", [("skewed_gauss_2d", 32, 2, 200, 1, "sequential"),
                                                         ("rosenbrock_2d", 2048, 2, 40, 3, "per_walker"),
                                                         ("rosenbrock_2d", 2048, 2, 30, 1, "sequential"),
                                                         ("diag_gauss", 64, 8, 50, 2, "sequential"),
                                                         ("dense_gauss", 40, 5, 60, 1, "sequential")])
def test_fused_small_ensemble_kernel(port, kind, W, D, steps, interval, rng):
    """Small ensembles run the whole runMCMC loop in one launch of one CTA (mcmcpp_fused_args).  Same chain, counters and
    rng continuation as the oracle and as the launch-per-colour path (which a trace request selects)."""
    params = {"skewed_gauss_2d": [0.13], "rosenbrock_2d": [], "diag_gauss": 1.0 / (1.0 + (np.arange(D) % 10)),
              "dense_gauss": corr_precision(D).reshape(-1)}[kind]
    init, lp = seeded_init(port, kind, params if len(params) else None, W, D, 5)
    runs = []
    for trace in (0, 1):
        s = make_sampler(kind, params, W, D, seed=9, rngMode=rng, traceUpdates=trace)
        s.setSlicingMode(interval > 1, interval)
        s.setInitialWalkerPos(init, lp)
        assert s.runMCMC(steps) and s.runMCMC(3)                      # the second call continues every walker's stream
        launches = s.getLaunchCount()
        assert launches == (2 if trace == 0 else 2 * (steps + 3) * interval)
        runs.append((np.array(s.getChain()), s.getWalkerState(), s.getAcceptedSteps(), s.getTotalSteps()))   # the chain view dies with close()
        s.close()
    (c0, (p0, a0), acc0, tot0), (c1, (p1, a1), acc1, tot1) = runs
    assert np.array_equal(c0, c1) and np.array_equal(p0, p1) and np.array_equal(a0, a1) and (acc0, tot0) == (acc1, tot1)
    o = port.run(kind, params if len(params) else None, init, lp, steps + 3, interval=interval, seed=9, per_walker=(rng == "per_walker"))
    assert np.array_equal(c0[1:], o["chain"]) and np.array_equal(p0, o["pos"]) and np.array_equal(a0, o["logp"])
    assert (acc0, tot0) == (int(o["acc"].sum()), int(o["tot"].sum()))


def test_fused_small_ensemble_chain_cap(port):
    """runMCMC returns False at the byte cap (EnsembleSampler.h:293,308) at the same stored count on both paths."""
    W, D = 64, 2
    init, lp = seeded_init(port, "skewed_gauss_2d", [0.13], W, D, 6)
    out = []
    for trace in (0, 1):
        s = make_sampler("skewed_gauss_2d", [0.13], W, D, seed=1, maxChainSizeBytes=25 * W * D * 8, traceUpdates=trace)
        s.setInitialWalkerPos(init, lp)
        assert s.runMCMC(10) is True
        assert s.runMCMC(100) is False
        out.append((s.getStoredSteps(), np.array(s.getChain()), s.getTotalSteps()))
        s.close()
    assert out[0][0] == out[1][0] == 25 and np.array_equal(out[0][1], out[1][1]) and out[0][2] == out[1][2]
