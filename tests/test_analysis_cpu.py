"""CPU: the CovarianceMatrix restatement (oracle/oracle.c oracle_covariance) against fixtures made by the UNMODIFIED
reference class (MCMCpp/Analysis/CovarianceMatrix.h via oracle/ref_driver.cpp, tests/golden/make_golden.py)."""
import os
import numpy as np
import pytest
from conftest import GOLDEN

CASES = ["cov_skewed_w32", "cov_diag_d5_w64_slice3"]


def load(name):
    z = np.load(os.path.join(GOLDEN, "analysis", name + ".npz"))
    return z["chain"], z["cov"], z["corr"], int(z["cov_interval"])


@pytest.mark.parametrize("name", CASES)
def test_covariance_restatement_bit_exact_against_reference_fixture(port, name):
    chain, cov, corr, interval = load(name)
    mean, c, r, n = port.covariance(chain, interval)
    assert np.array_equal(c, cov) and np.array_equal(r, corr)
    assert n == len(range(0, chain.shape[0], interval)) * chain.shape[1]
    # sanity of the definition: agrees with numpy's population covariance to rounding
    pts = chain[::interval].reshape(-1, chain.shape[2])
    assert np.allclose(c, np.cov(pts.T, bias=True), rtol=1e-10, atol=1e-13)
    assert np.allclose(mean, pts.mean(axis=0), rtol=1e-12, atol=1e-15)


def test_skewed_gaussian_covariance_near_analytic_truth(port):
    """2-D target of the reference tests: analytic covariance quoted in test/sequential/SkewedGaussian/MH/src/main.cpp:74
    (eps = 0.13): [[1.13, 0.435], [0.435, 0.2825]]; a 300-step, 32-walker chain is only a loose check."""
    chain, cov, corr, interval = load("cov_skewed_w32")
    _, c, _, _ = port.covariance(chain[100:], 1)
    assert abs(c[0, 0] - 1.13) < 0.35 and abs(c[0, 1] - 0.435) < 0.2 and abs(c[1, 1] - 0.2825) < 0.12


@pytest.mark.parametrize("name", ["ac_diag_d5_w64", "ac_diag_d3_w48"])
def test_autocorr_restatement_bit_exact_against_reference_fixture(port, name):
    """AutoCorrCalc + AutoCov restatement (oracle_autocorr, quirk mode) against the unmodified reference class.
    Parameter 0 is excluded: the reference adds the first walker's chain onto an uninitialised new[] buffer
    (AutoCorrCalc.h:228-232,316-318), so its iat[0] depends on heap contents (a NaN there poisons every later parameter; the generator retries until it
    gets a finite run); parameters >= 1 are then deterministic and
    must match bit for bit (FFT butterflies, twiddles, averaging, window rule)."""
    z = np.load(os.path.join(GOLDEN, "analysis", name + ".npz"))
    chain, iat = z["chain"], z["iat"]
    q = port.autocorr(chain, 4, quirk=True)
    assert np.array_equal(q[1:], iat[1:])
    clean = port.autocorr(chain, 4, quirk=False)
    assert np.all(clean > 1.0) and np.all(np.abs(clean / q - 1.0) < 0.25)      # the quirk perturbs, it does not dominate


def test_autocorr_recovers_ar1_autocorrelation_time(port):
    """AR(1) series x_t = phi x_{t-1} + noise has integrated autocorrelation time (1 + phi) / (1 - phi): the golden
    values quoted in test/sequential/AcTime/src/main.cpp:16-22 (9 for phi = 0.8, 20 for 0.904761904762) come from that
    formula; 64 walkers x 16383 steps recover it within a few percent (clean definition and the reference's)."""
    rng = np.random.default_rng(0)
    S, W, phis = 16383, 64, np.array([0.8, 0.904761904762])
    x = np.zeros((S, W, 2))
    x[0] = rng.standard_normal((W, 2))
    noise = rng.standard_normal((S, W, 2)) * np.sqrt(1.0 - phis ** 2)
    for t in range(1, S):
        x[t] = phis * x[t - 1] + noise[t]
    for quirk in (False, True):
        iat = port.autocorr(x, 4, quirk=quirk)
        assert np.all(np.abs(iat / np.array([9.0, 20.0]) - 1.0) < 0.06)
