"""GPU (-m gpu): device CovarianceMatrix (mcmcpp_b200/csrc/analysis.cu) through the C-ABI against the reference-made
fixtures and the oracle.  Definitions are the reference's; the summation order is per-CTA Kahan pairs, so the bar is
1e-12 relative on covariance entries measured against the scale of their terms (E[xy] and E[x]E[y]), stated below."""
import os
import numpy as np
import pytest
from conftest import GOLDEN

pytestmark = pytest.mark.gpu


def close(dev, ref, scale, rtol=1e-12):
    return np.all(np.abs(dev - ref) <= rtol * scale)


def scales(chain, interval):
    pts = chain[::interval].reshape(-1, chain.shape[2])
    m2 = (pts[:, :, None] * pts[:, None, :]).mean(axis=0)
    return np.abs(m2) + np.abs(np.outer(pts.mean(axis=0), pts.mean(axis=0))) + 1e-300


@pytest.mark.parametrize("name", ["cov_skewed_w32", "cov_diag_d5_w64_slice3"])
def test_device_covariance_against_reference_fixture(name):
    import mcmcpp_b200 as m
    z = np.load(os.path.join(GOLDEN, "analysis", name + ".npz"))
    chain, cov, corr, interval = z["chain"], z["cov"], z["corr"], int(z["cov_interval"])
    S, W, D = chain.shape
    cm = m.CovarianceMatrix(D, W)
    cm.calculateCovar(chain, interval)
    assert cm.points == len(range(0, S, interval)) * W
    assert close(cm.covarMat, cov, scales(chain, interval))
    assert np.allclose(cm.corrMat, corr, rtol=0, atol=1e-11)
    assert cm.getCovarianceMatrixElement(D - 1, 0) == cm.covarMat[D - 1, 0]
    assert abs(cm.getCorrelationMatrixElement(0, 0) - 1.0) < 1e-12


@pytest.mark.parametrize("W,D,S,interval", [(64, 1, 40, 1), (4096, 2, 30, 2), (512, 3, 25, 1), (256, 4, 33, 5), (128, 7, 50, 1),
                                            (300, 50, 12, 1), (256, 100, 9, 2), (260, 128, 5, 1)])
def test_device_covariance_every_kernel_shape_against_oracle(port, W, D, S, interval):
    """thread-per-point kernels (D <= 4), thread-per-item kernel (5..128, ragged chunks), slicing, multi-upload."""
    import mcmcpp_b200 as m
    rng = np.random.default_rng(D * 1000 + W)
    A = rng.standard_normal((D, D)) / np.sqrt(D)
    chain = rng.standard_normal((S, W, D)) @ A + rng.standard_normal(D) * 3.0
    mean, cov, corr, n = port.covariance(chain, interval)
    cm = m.CovarianceMatrix(D, W)
    cm.calculateCovar(chain, interval)
    assert cm.points == n
    assert np.allclose(cm.means, mean, rtol=1e-13, atol=1e-15)
    assert close(cm.covarMat, cov, scales(chain, interval))
    assert np.allclose(cm.corrMat, corr, rtol=0, atol=1e-10)


def test_on_the_fly_covariance_matches_chain_covariance(port):
    """covBegin / runMCMC / covResult with the chain kept: the device accumulation over the stored steps equals
    calculateCovar over the very chain the run produced (oracle restatement), slicing included."""
    import mcmcpp_b200 as m
    W, D, steps = 2048, 2, 60
    rng = np.random.default_rng(4)
    init = rng.standard_normal((W, D))
    lp = port.logp("skewed_gauss_2d", np.array([0.13]), init)
    s = m.EnsembleSampler(3, W, D, m.StretchMove(D, 3, m.SkewedGaussianTwoDim(0.13)))
    s.setInitialWalkerPos(init, lp)
    s.covBegin(sliceInterval=3, includeCurrent=True)
    assert s.runMCMC(steps)
    mean, cov, corr, n = s.covResult()
    chain = s.getChain()
    om, oc, orr, on = port.covariance(chain, 3)
    assert n == on == len(range(0, steps + 1, 3)) * W
    assert np.allclose(mean, om, rtol=1e-13, atol=1e-15) and close(cov, oc, scales(chain, 3))
    assert np.allclose(corr, orr, rtol=0, atol=1e-11)
    s.covEnd()
    s.close()


def test_on_the_fly_covariance_full_size_c2_without_chain():
    """BASELINE config 2 size (2^20 walkers, chain off): burn in, then 20 accumulated steps; the banana target's
    moments are known in closed form for x (marginal N(1, 10)): mean 1, variance 10 -- statistical check, wide bars."""
    import mcmcpp_b200 as m
    W, D = 1 << 20, 2
    rng = np.random.default_rng(9)
    x = 1.0 + np.sqrt(10.0) * rng.standard_normal(W)
    y = x * x + np.sqrt(0.1) * rng.standard_normal(W)        # exact draws from the target: -(100 (y-x^2)^2 + (1-x)^2)/20
    init = np.stack([x, y], axis=1)
    lp = -(100.0 * (y - x * x) ** 2 + (1.0 - x) ** 2) / 20.0
    s = m.EnsembleSampler(1, W, D, m.StretchMove(D, 1, m.RosenbrockTwoDim()), maxChainSizeBytes=0)
    s.setInitialWalkerPos(init, lp)
    assert s.runMCMC(30)
    s.covBegin(sliceInterval=1, includeCurrent=False)
    assert s.runMCMC(20)
    mean, cov, corr, n = s.covResult()
    assert n == 20 * W
    assert abs(mean[0] - 1.0) < 0.05 and abs(cov[0, 0] - 10.0) < 0.3
    assert abs(mean[1] - 11.0) < 0.3                           # E[y] = E[x^2] = 1 + 10
    s.close()


def test_tier_c_long_run_statistics_match_oracle_and_analytic_truth(port):
    """SURVEY s.8(c) tier C: a long run whose RNG discipline is NOT the sequential stream (one pcg64 stream per walker,
    the reference's per-thread discipline) must be statistically indistinguishable from the sequential oracle and from
    the analytic truth of the reference's 2-D target (eps = 0.13; covariance quoted in
    test/sequential/SkewedGaussian/MH/src/main.cpp:74): means, covariance (device CovarianceMatrix), acceptance fraction.
    Bars are ~6 sigma of the Monte-Carlo error of 4096 walkers x 400 kept steps at IAT ~ 3."""
    import mcmcpp_b200 as m
    W, D, burn, keep = 4096, 2, 100, 400
    eps = 0.13
    truth = np.array([[4.0 * (1 + eps) / 4.0, (1 - eps) / 2.0], [(1 - eps) / 2.0, (1 + eps) / 4.0]])   # Sigma of -(t1^2/eps + t2^2)/2
    rng = np.random.default_rng(12)
    init = rng.standard_normal((W, D))
    lp = port.logp("skewed_gauss_2d", np.array([eps]), init)
    s = m.EnsembleSampler(5, W, D, m.StretchMove(D, 5, m.SkewedGaussianTwoDim(eps)), rngMode="per_walker", maxChainSizeBytes=0)
    s.setInitialWalkerPos(init, lp)
    assert s.runMCMC(burn)
    s.reset()
    s.covBegin(1, includeCurrent=False)
    assert s.runMCMC(keep)
    mean, cov, corr, n = s.covResult()
    acc_dev = s.getAcceptanceFraction()
    s.close()
    o = port.run("skewed_gauss_2d", np.array([eps]), init, lp, burn + keep, seed=5)
    om, oc, _, _ = port.covariance(o["chain"][burn:], 1)
    o1 = port.run("skewed_gauss_2d", np.array([eps]), init, lp, burn, seed=5, want_chain=False)
    acc_orc = (o["acc"].sum() - o1["acc"].sum()) / (o["tot"].sum() - o1["tot"].sum())      # kept steps only
    n_eff = W * keep / 3.0
    sd = np.sqrt(np.diag(truth))
    assert np.all(np.abs(mean) < 6 * sd / np.sqrt(n_eff)) and np.all(np.abs(mean - om) < 9 * sd / np.sqrt(n_eff))
    tol = 6 * np.sqrt((np.outer(sd, sd) ** 2 + truth ** 2) / n_eff)
    assert np.all(np.abs(cov - truth) < tol) and np.all(np.abs(cov - oc) < 1.5 * tol)
    assert abs(acc_dev - acc_orc) < 0.004 and 0.6 < acc_dev < 0.8


@pytest.mark.parametrize("name", ["ac_diag_d5_w64", "ac_diag_d3_w48"])
def test_device_autocorr_against_reference_fixture(port, name):
    """Device AutoCorrCalc: the reference's arithmetic in the reference's order -> bit-identical to the restatement in
    both modes, and (reference mode) to the unmodified reference class on the parameters it computes reproducibly
    (>= 1; parameter 0 reads uninitialised memory there, tests/test_analysis_cpu.py)."""
    import mcmcpp_b200 as m
    z = np.load(os.path.join(GOLDEN, "analysis", name + ".npz"))
    chain, iat = z["chain"], z["iat"]
    S, W, D = chain.shape
    ac = m.AutoCorrCalc(D, W)
    ac.calcAutoCorrTimes(chain, S)
    assert np.array_equal(ac.acorrTimeList[1:], iat[1:])
    assert np.array_equal(ac.acorrTimeList, port.autocorr(chain, 4, quirk=True))
    assert ac.retrieveAutoCorrelationTime(1) == iat[1]
    clean = m.AutoCorrCalc(D, W, referenceBufferReuse=False)
    clean.calcAutoCorrTimes(chain, S)
    assert np.array_equal(clean.acorrTimeList, port.autocorr(chain, 4, quirk=False))


@pytest.mark.parametrize("S,W,D,scale", [(100, 8, 2, 4), (4096, 16, 3, 4), (5000, 4, 1, 5), (16383, 32, 2, 4)])
def test_device_autocorr_shapes_and_ar1_known_answer(port, S, W, D, scale):
    """power-of-two and ragged lengths, other window scalings, against the oracle bit for bit; the AR(1) series'
    autocorrelation time (1 + phi) / (1 - phi) is recovered on the long case (AcTime/src/main.cpp:16-22 goldens)."""
    import mcmcpp_b200 as m
    rng = np.random.default_rng(S + W)
    phis = np.array([0.8, 0.904761904762, 0.5])[:D]
    x = np.zeros((S, W, D))
    x[0] = rng.standard_normal((W, D))
    noise = rng.standard_normal((S, W, D)) * np.sqrt(1.0 - phis ** 2)
    for t in range(1, S):
        x[t] = phis * x[t - 1] + noise[t]
    for reuse in (True, False):
        ac = m.AutoCorrCalc(D, W, referenceBufferReuse=reuse)
        ac.setAutoCorrScaleFactor(scale)
        ac.calcAutoCorrTimes(x, S)
        assert np.array_equal(ac.acorrTimeList, port.autocorr(x, scale, quirk=reuse))
    if S >= 16383:
        assert np.all(np.abs(ac.acorrTimeList / ((1 + phis) / (1 - phis)) - 1.0) < 0.08)


def test_tier_c_autocorrelation_time_of_gpu_chain_matches_oracle_chain(port):
    """SURVEY s.8(c) tier C, IAT: a per-walker-stream GPU run and the sequential oracle run of the same target give
    statistically indistinguishable integrated autocorrelation times (clean definition, device AutoCorrCalc)."""
    import mcmcpp_b200 as m
    W, D, burn, keep = 256, 2, 200, 4095
    eps = 0.13
    rng = np.random.default_rng(2)
    init = rng.standard_normal((W, D))
    lp = port.logp("skewed_gauss_2d", np.array([eps]), init)
    s = m.EnsembleSampler(8, W, D, m.StretchMove(D, 8, m.SkewedGaussianTwoDim(eps)), rngMode="per_walker")
    s.setInitialWalkerPos(init, lp)
    assert s.runMCMC(burn + keep)
    dev_chain = s.getChain()[burn + 1:].copy()
    s.close()
    o = port.run("skewed_gauss_2d", np.array([eps]), init, lp, burn + keep, seed=8)
    ac = m.AutoCorrCalc(D, W, referenceBufferReuse=False)
    ac.calcAutoCorrTimes(dev_chain, keep)
    dev_iat = ac.acorrTimeList.copy()
    ac.calcAutoCorrTimes(np.ascontiguousarray(o["chain"][burn:]), keep)
    orc_iat = ac.acorrTimeList.copy()
    assert np.all(dev_iat > 1.0) and np.all(np.abs(dev_iat / orc_iat - 1.0) < 0.15)
