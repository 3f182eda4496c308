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
