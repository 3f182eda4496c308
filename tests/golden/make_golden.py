"""Regenerates tests/golden/*.npz by running the UNMODIFIED reference (oracle/_ref, built from
/root/reference by oracle/Makefile).  Run in the build container only:  python tests/golden/make_golden.py
Every fixture holds the inputs (init positions/logp, target params) and the reference outputs
(chain rows, ensemble accepted/total), so tests need neither /root/reference nor oracle/_ref."""
import os
import sys
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle.pyoracle import Ref, build  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def corr_precision(D, rho=0.9):
    idx = np.arange(D)
    cov = rho ** np.abs(idx[:, None] - idx[None, :])
    return np.linalg.inv(cov)


def pcg32_cases(R):
    """PER_WALKER32 fixtures: vendored pcg32 + libstdc++ adapters + reference GwDistribution, one facade per walker
    (oracle/ref_driver.cpp:ref_stretch_run_pcg32)."""
    os.makedirs(os.path.join(HERE, "pcg32"), exist_ok=True)

    def case(name, kind, params, W, D, nsteps, keep_rows, interval=1, seed=0, init_seed=53):
        init = R.init_normal(init_seed, W * D).reshape(W, D)
        params = np.asarray(params if params is not None else [0.0], dtype=np.float64)
        lp = R.logp(kind, params, init)
        r = R.run_pcg32(kind, params, init, lp, nsteps, interval=interval, seed=seed)
        rows = np.asarray(keep_rows)
        np.savez_compressed(os.path.join(HERE, "pcg32", name + ".npz"), kind=kind, params=params, W=W, D=D, nsteps=nsteps,
                            interval=interval, seed=seed, init=init, logp=lp, rows=rows, chain=r["chain"][rows],
                            final_logp=r["logp"], acc=r["acc"], tot=r["tot"])
        print("pcg32/" + name, r["acc"], r["tot"])

    case("skewed_w32", "skewed_gauss_2d", [0.13], 32, 2, 2000, [0, 1, 2, 1000, 2000])
    case("rosenbrock_w512", "rosenbrock_2d", None, 512, 2, 100, [0, 1, 100], seed=3)
    case("skewed_w320_slice3", "skewed_gauss_2d", [0.13], 320, 2, 100, [0, 1, 100], interval=3, seed=11)
    case("diag_d37_w128", "diag_gauss", 1.0 / (1.0 + (np.arange(37) % 10)), 128, 37, 20, [0, 1, 20])
    case("diag_d200_w512", "diag_gauss", 1.0 / (1.0 + (np.arange(200) % 10)), 512, 200, 6, [0, 1, 6], seed=2)


def main():
    build()
    R = Ref()
    if len(sys.argv) > 1 and sys.argv[1] == "pcg32":
        return pcg32_cases(R)
    out = {}

    def case(name, kind, params, W, D, nsteps, interval=1, seed=0, init_seed=53, scale=1.0, keep_rows=None):
        init = (R.init_normal(init_seed, W * D) * scale).reshape(W, D)
        params = np.asarray(params if params is not None else [0.0], dtype=np.float64)
        lp = R.logp(kind, params, init)
        r = R.run(kind, params, init, lp, nsteps, interval=interval, seed=seed)
        chain = r["chain"]            # (nsteps+1, W, D), row 0 = initial positions
        rows = np.arange(chain.shape[0]) if keep_rows is None else np.asarray(keep_rows)
        np.savez_compressed(os.path.join(HERE, name + ".npz"), kind=kind, params=params, W=W, D=D, nsteps=nsteps,
                            interval=interval, seed=seed, init=init, logp=lp, rows=rows, chain=chain[rows],
                            acc=r["acc"], tot=r["tot"], stored=r["stored"])
        out[name] = (r["acc"], r["tot"])

    # C1: the survey KAT (BASELINE config 1): acc 228883/320032, walker 0 of the last row pinned
    case("c1_skewed_w32", "skewed_gauss_2d", [0.13], 32, 2, 10000, keep_rows=[0, 1, 2, 100, 5000, 9999, 10000])
    # the reference test's walker count, with slicing (test/sequential/SkewedGaussian/StretchMove/src/main.cpp:22-41)
    case("skewed_w320_slice3", "skewed_gauss_2d", [0.13], 320, 2, 200, interval=3, keep_rows=[0, 1, 50, 200])
    case("rosenbrock_w512", "rosenbrock_2d", None, 512, 2, 100, keep_rows=[0, 1, 100])
    iv = 1.0 / (1.0 + (np.arange(100) % 10))
    case("diag_d100_w256", "diag_gauss", iv, 256, 100, 20, keep_rows=[0, 1, 20])
    iv37 = 1.0 / (1.0 + (np.arange(37) % 10))
    case("diag_d37_w128", "diag_gauss", iv37, 128, 37, 20, keep_rows=[0, 1, 20])
    iv5 = 1.0 / (1.0 + (np.arange(5) % 10))
    case("diag_d5_w64", "diag_gauss", iv5, 64, 5, 50, keep_rows=[0, 1, 50])
    case("dense_d12_w64", "dense_gauss", corr_precision(12).reshape(-1), 64, 12, 30, keep_rows=[0, 1, 30])
    rng = np.random.default_rng(1234)
    M, P = 40, 3
    X = rng.standard_normal((M, P)); th = rng.standard_normal(P); y = X @ th + rng.standard_normal(M)
    lin = np.concatenate([[float(M), 100.0], X.reshape(-1), y])
    case("linreg_p3_w16", "linreg", lin, 16, 3, 50, scale=0.3, keep_rows=[0, 1, 50])
    # Analysis/CovarianceMatrix.h on the chains of two short runs (reference class, unmodified)
    def cov_case(name, kind, params, W, D, nsteps, cov_interval, init_seed=53):
        init = R.init_normal(init_seed, W * D).reshape(W, D)
        params = np.asarray(params, dtype=np.float64)
        lp = R.logp(kind, params, init)
        chain, cov, corr = R.covariance(kind, params, init, lp, nsteps, cov_interval=cov_interval)
        np.savez_compressed(os.path.join(HERE, name + ".npz"), chain=chain, cov=cov, corr=corr, cov_interval=cov_interval)
        print(name, cov.reshape(-1)[:3])

    os.makedirs(os.path.join(HERE, "analysis"), exist_ok=True)
    cov_case("analysis/cov_skewed_w32", "skewed_gauss_2d", [0.13], 32, 2, 300, 1)
    cov_case("analysis/cov_diag_d5_w64_slice3", "diag_gauss", iv5, 64, 5, 60, 3)
    # Analysis/AutoCorrCalc.h (reference class, unmodified, all walkers).  NB the reference reads uninitialised memory for
    # (parameter 0, first walker) -- AutoCorrCalc.h:316-318 allocates acovFuncArray with new[] and :228-232 adds the chain
    # onto it -- so iat[0] depends on heap contents (NaN has been seen); parameters >= 1 are deterministic.
    def ac_case(name, kind, params, W, D, nsteps, init_seed=53):
        init = R.init_normal(init_seed, W * D).reshape(W, D)
        params = np.asarray(params, dtype=np.float64)
        lp = R.logp(kind, params, init)
        for attempt in range(8):            # heap garbage can be NaN and then poisons every parameter: take a clean run
            chain, iat = R.autocorr(kind, params, init, lp, nsteps)
            if np.all(np.isfinite(iat[1:])):
                break
        np.savez_compressed(os.path.join(HERE, name + ".npz"), chain=chain, iat=iat)
        print(name, iat)

    ac_case("analysis/ac_diag_d5_w64", "diag_gauss", iv5, 64, 5, 300)
    ac_case("analysis/ac_diag_d3_w48", "diag_gauss", iv5[:3], 48, 3, 511)
    # Movers/DifferentialEvolution.h in the sequential sampler (reference mover, unmodified)
    def de_case(name, kind, params, W, D, nsteps, keep_rows, init_seed=53):
        init = R.init_normal(init_seed, W * D).reshape(W, D)
        params = np.asarray(params, dtype=np.float64)
        lp = R.logp(kind, params, init)
        r = R.run_de(kind, params, init, lp, nsteps)
        rows = np.asarray(keep_rows)
        np.savez_compressed(os.path.join(HERE, name + ".npz"), kind=kind, params=params, W=W, D=D, nsteps=nsteps, init=init, logp=lp,
                            rows=rows, chain=r["chain"][rows], acc=r["acc"], tot=r["tot"])
        print(name, r["acc"], r["tot"])

    os.makedirs(os.path.join(HERE, "movers"), exist_ok=True)
    de_case("movers/de_skewed_w16", "skewed_gauss_2d", [0.13], 16, 2, 300, [0, 1, 2, 150, 300])
    de_case("movers/de_diag_d5_w64", "diag_gauss", iv5, 64, 5, 100, [0, 1, 50, 100])
    # Movers/WalkMove.h in the sequential sampler (reference mover, unmodified)
    def walk_case(name, kind, params, W, D, nsteps, num_points, keep_rows, init_seed=53):
        init = R.init_normal(init_seed, W * D).reshape(W, D)
        params = np.asarray(params, dtype=np.float64)
        lp = R.logp(kind, params, init)
        r = R.run_walk(kind, params, init, lp, nsteps, num_points)
        rows = np.asarray(keep_rows)
        np.savez_compressed(os.path.join(HERE, name + ".npz"), kind=kind, params=params, W=W, D=D, nsteps=nsteps, num_points=num_points,
                            init=init, logp=lp, rows=rows, chain=r["chain"][rows], acc=r["acc"], tot=r["tot"])
        print(name, r["acc"], r["tot"])

    walk_case("movers/walk_skewed_w16_s3", "skewed_gauss_2d", [0.13], 16, 2, 300, 3, [0, 1, 2, 150, 300])
    walk_case("movers/walk_diag_d5_w64_s6", "diag_gauss", iv5, 64, 5, 100, 6, [0, 1, 50, 100])
    walk_case("movers/walk_skewed_w32_s16", "skewed_gauss_2d", [0.13], 32, 2, 100, 16, [0, 1, 100])
    pcg32_cases(R)
    for k, v in out.items():
        print(k, v)


if __name__ == "__main__":
    main()
