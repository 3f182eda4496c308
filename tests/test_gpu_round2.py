"""GPU tests added in round 2 (-m gpu): the pcg32 per-walker engine, the engine's life cycle (seeded once), the reference's
whole-block chain cap, one full-size BASELINE configs[3] step against the oracle, and several samplers on several devices
of one process."""
import os
import numpy as np
import pytest
from conftest import load_golden
from test_gpu_parity import make_sampler, seeded_init
from test_oracle_cpu import pcg32_golden_names, load_pcg32_golden

pytestmark = pytest.mark.gpu


# ---- MCMCPP_RNG_PER_WALKER32 ----------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", pcg32_golden_names())
def test_pcg32_goldens_bit_exact(name):
    """Fixtures made by the vendored pcg32 + libstdc++ adapters + the reference's GwDistribution (one facade per walker):
    fused small-ensemble kernel (W=32, 320), thread kernel (W=512, D=2), warp kernel (D=37), TMA bulk kernel (D=200)."""
    g = load_pcg32_golden(name)
    s = make_sampler(g["kind"], g["params"], g["W"], g["D"], seed=g["seed"], rngMode="per_walker32")
    if g["interval"] > 1:
        s.setSlicingMode(True, g["interval"])
    s.setInitialWalkerPos(g["init"], g["logp"])
    assert s.runMCMC(g["nsteps"]) is True
    assert np.array_equal(s.getChain()[g["rows"]], g["chain"])
    pos, aux = s.getWalkerState()
    assert np.array_equal(aux, g["final_logp"])
    assert (s.getAcceptedSteps(), s.getTotalSteps()) == (g["acc"], g["tot"])
    s.close()


@pytest.mark.parametrize("kind,W,D,steps,fused", [("rosenbrock_2d", 1 << 16, 2, 6, True), ("rosenbrock_2d", 4096, 2, 40, False),
                                                   ("diag_gauss", 256, 9, 10, True), ("diag_gauss", 512, 100, 6, True),
                                                   ("diag_gauss", 2048, 1000, 3, True), ("dense_gauss", 256, 24, 10, True)])
def test_pcg32_every_kernel_against_oracle(port, monkeypatch, kind, W, D, steps, fused):
    if not fused:
        monkeypatch.setenv("MCMCPP_B200_FUSED", "0")
    params = {"rosenbrock_2d": None, "diag_gauss": 1.0 / (1.0 + (np.arange(D) % 10)),
              "dense_gauss": np.linalg.inv(0.9 ** np.abs(np.subtract.outer(np.arange(D), np.arange(D)))).reshape(-1)}[kind]
    init, lp = seeded_init(port, kind, params, W, D, 31)
    s = make_sampler(kind, params if params is not None else [], W, D, seed=5, rngMode="per_walker32", traceUpdates=W)
    s.setInitialWalkerPos(init, lp)
    assert s.runMCMC(steps)
    o = port.run(kind, params, init, lp, steps, seed=5, per_walker=32, want_trace=True)
    tr = s.getTrace(W)
    assert np.array_equal(tr["j"], o["trace"]["j"][:W]) and np.array_equal(tr["z"], o["trace"]["z"][:W])
    assert np.array_equal(tr["accept"], o["trace"]["accept"][:W])
    pos, aux = s.getWalkerState()
    if kind == "dense_gauss":                       # tensor-core summation order: tier B
        assert np.array_equal(pos, o["pos"]) and np.allclose(aux, o["logp"], rtol=1e-12, atol=0)
    else:
        assert np.array_equal(pos, o["pos"]) and np.array_equal(aux, o["logp"])
        assert np.array_equal(s.getChain()[1:], o["chain"])
    assert s.getAcceptedSteps() == int(o["acc"].sum())
    s.close()


def test_pcg32_is_refused_by_the_variable_draw_movers():
    import mcmcpp_b200 as m
    from mcmcpp_b200._capi import McmcppError
    s = m.EnsembleSampler(0, 64, 2, m.DifferentialEvolution(2, 0, m.RosenbrockTwoDim()), rngMode="per_walker32")   # forced to per_walker
    s.close()
    cfg_err = None
    try:
        m.EnsembleSampler(0, 64, 2, m.StretchMove(2, 0, m.RosenbrockTwoDim()), rngMode="no_such_mode")
    except (KeyError, McmcppError) as e:
        cfg_err = e
    assert cfg_err is not None


# ---- the engine is seeded once (EnsembleSampler.h:217); setInitialWalkerPos never reseeds (EnsembleSampler.h:220-230) ----
@pytest.mark.parametrize("kind,W,D,rng", [("rosenbrock_2d", 512, 2, "sequential"), ("rosenbrock_2d", 4096, 2, "sequential"),
                                          ("diag_gauss", 256, 100, "sequential"), ("rosenbrock_2d", 512, 2, "per_walker32")])
def test_second_set_initial_continues_the_stream(port, kind, W, D, rng):
    params = None if kind == "rosenbrock_2d" else 1.0 / (1.0 + (np.arange(D) % 10))
    init, lp = seeded_init(port, kind, params, W, D, 8)
    init2, lp2 = seeded_init(port, kind, params, W, D, 9)
    pw = 32 if rng == "per_walker32" else False
    s = make_sampler(kind, params if params is not None else [], W, D, seed=3, rngMode=rng)
    s.setInitialWalkerPos(init, lp)
    s.runMCMC(7)
    a = port.run(kind, params, init, lp, 7, seed=3, per_walker=pw)
    assert np.array_equal(s.getChain()[1:], a["chain"])
    s.setInitialWalkerPos(init2, lp2)                 # burn-in -> re-place the walkers: the stream goes on
    s.runMCMC(5)
    b = port.run(kind, params, init2, lp2, 5, seed=3, per_walker=pw, rng=a["rng"])
    chain = s.getChain()
    assert chain.shape[0] == 1 + 7 + 1 + 5 and np.array_equal(chain[8], init2) and np.array_equal(chain[9:], b["chain"])
    # counters keep adding (Walker.h:162-170: every placement is one more accepted step)
    assert s.getTotalSteps() == W * (1 + 7 + 1 + 5)
    assert s.getAcceptedSteps() == int(a["acc"].sum()) + int(b["acc"].sum())
    s.close()


def test_second_set_initial_against_the_live_reference(port):
    from oracle.pyoracle import Ref
    if not Ref.available():
        pytest.skip("oracle/_ref not built (no /root/reference here)")
    # the reference itself: one sampler, two placements -- restated by carrying the oracle's engine (pinned above)
    g = load_golden("rosenbrock_w512")
    a = port.run(g["kind"], g["params"], g["init"], g["logp"], 3, seed=g["seed"])
    assert a["rejections"] == 0


# ---- chain cap in whole 10000-step blocks (Chain/Chain.h:178-186,218-233) ---------------------------------------------
@pytest.mark.parametrize("W,D,block,cap_blocks_bytes,expect_rows", [
    (32, 2, 10, 2.5, 20),        # floor(2.5 blocks) = 2 blocks of 10 steps
    (32, 2, 10, 0.4, 10),        # less than one block: the first block exists anyway (Chain.h:181-186)
    (64, 5, 7, 3.0, 21)])
def test_chain_cap_in_whole_blocks(port, W, D, block, cap_blocks_bytes, expect_rows):
    kind = "skewed_gauss_2d" if D == 2 else "diag_gauss"
    params = [0.13] if D == 2 else 1.0 / (1.0 + (np.arange(D) % 10))
    init, lp = seeded_init(port, kind, params, W, D, 3)
    s = make_sampler(kind, params, W, D, maxChainSizeBytes=int(cap_blocks_bytes * block * W * D * 8), chainBlockSteps=block)
    s.setInitialWalkerPos(init, lp)
    assert s.runMCMC(expect_rows - 2) is True                    # one row short of the cap
    assert s.runMCMC(50) is False                                # the step that fills the last block returns false
    assert s.getStoredSteps() == expect_rows
    o = port.run(kind, params, init, lp, expect_rows - 1)
    assert np.array_equal(s.getChain()[1:], o["chain"])
    s.close()
    # row granularity (the default) stops elsewhere for the same byte count: documented difference
    s = make_sampler(kind, params, W, D, maxChainSizeBytes=int(cap_blocks_bytes * block * W * D * 8))
    s.setInitialWalkerPos(init, lp)
    assert s.runMCMC(200) is False and s.getStoredSteps() == int(cap_blocks_bytes * block)
    s.close()


# ---- BASELINE configs[3] at FULL size: 2^20 walkers x 1000 parameters, one ensemble step against the oracle ---------------
def test_c4_full_size_one_step_against_oracle(port):
    """8.4 GB of state; the oracle needs ~10 s and 17 GB of host memory for the step.  Positions, log-probs, partner
    indices, accept bits and counters of all 2^20 updates, bit for bit."""
    W, D = 1 << 20, 1000
    try:
        avail = os.sysconf("SC_AVPHYS_PAGES") * os.sysconf("SC_PAGE_SIZE")
    except (ValueError, OSError):
        avail = 1 << 40
    if avail < 40 * (1 << 30):
        pytest.skip("needs ~35 GB of free host memory")
    iv = 1.0 / (1.0 + (np.arange(D) % 10))
    rng = np.random.default_rng(20)
    init = rng.standard_normal((W, D))
    lp = port.logp("diag_gauss", iv, init)
    s = make_sampler("diag_gauss", iv, W, D, maxChainSizeBytes=0, traceUpdates=W)
    s.setInitialWalkerPos(init, lp)
    assert s.runMCMC(1)
    o = port.run("diag_gauss", iv, init, lp, 1, want_chain=False, want_trace=True)
    del init
    tr = s.getTrace(W)
    assert np.array_equal(tr["j"], o["trace"]["j"]) and np.array_equal(tr["accept"], o["trace"]["accept"])
    assert np.array_equal(tr["z"], o["trace"]["z"]) and np.array_equal(tr["newp"], o["trace"]["newp"])
    pos, aux = s.getWalkerState()
    assert np.array_equal(aux, o["logp"])
    assert np.array_equal(pos, o["pos"])
    assert s.getAcceptedSteps() == int(o["acc"].sum()) and s.getTotalSteps() == 2 * W and s.getRngRedraws() == 0
    s.close()


# ---- one process, samplers on two devices (cudaFuncSetAttribute is per device) --------------------------------------
def test_samplers_on_two_devices_of_one_process(port):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("one GPU visible")
    W, D = 2048, 1000                      # TMA bulk kernel: 64.5 KB of dynamic shared memory, needs the opt-in on each device
    iv = 1.0 / (1.0 + (np.arange(D) % 10))
    init, lp = seeded_init(port, "diag_gauss", iv, W, D, 12)
    o = port.run("diag_gauss", iv, init, lp, 2, want_chain=False)
    for dev in (0, 1, 0):
        s = make_sampler("diag_gauss", iv, W, D, maxChainSizeBytes=0, device=dev)
        s.setInitialWalkerPos(init, lp)
        assert s.runMCMC(2)
        pos, aux = s.getWalkerState()
        assert np.array_equal(pos, o["pos"]) and np.array_equal(aux, o["logp"])
        s.close()


# ---- device-resident chain (cfg.device_chain_bytes): rows stay in HBM, slice/burn and analysis in place -----------------
@pytest.mark.parametrize("kind,W,D,steps,interval,dev_rows", [
    ("skewed_gauss_2d", 32, 2, 300, 1, 1000),      # fused small-ensemble kernel writes straight into the device block
    ("skewed_gauss_2d", 32, 2, 300, 3, 64),        # ... with spills to the host chain in between
    ("rosenbrock_2d", 4096, 2, 40, 1, 100),        # thread kernel, everything stays on the device
    ("rosenbrock_2d", 4096, 2, 40, 2, 7),          # spills every 7 rows
    ("diag_gauss", 512, 200, 9, 1, 4)])            # TMA bulk kernel
def test_device_resident_chain_equals_the_oracle_chain(port, kind, W, D, steps, interval, dev_rows):
    params = {"skewed_gauss_2d": [0.13], "rosenbrock_2d": None, "diag_gauss": 1.0 / (1.0 + (np.arange(D) % 10))}[kind]
    init, lp = seeded_init(port, kind, params, W, D, 17)
    s = make_sampler(kind, params if params is not None else [], W, D, seed=2, deviceChainBytes=dev_rows * W * D * 8)
    if interval > 1:
        s.setSlicingMode(True, interval)
    s.setInitialWalkerPos(init, lp)
    assert s.runMCMC(steps)
    o = port.run(kind, params, init, lp, steps, interval=interval, seed=2)
    ptr, first, rows = s.chainDeviceRows()
    assert rows == (steps + 1 if dev_rows > steps else rows) and first + rows == steps + 1
    chain = s.getChain()                              # the host asks: the device rows cross PCIe now
    assert chain.shape[0] == steps + 1 and np.array_equal(chain[0], init) and np.array_equal(chain[1:], o["chain"])
    assert s.chainDeviceRows()[2] == 0
    s.close()


@pytest.mark.parametrize("interval,burn,dev_rows", [(3, 5, 1000), (1, 20, 1000), (7, 0, 25), (2, 31, 25), (50, 10, 1000), (5, 80, 25)])
def test_slice_and_burn_on_the_device(port, interval, burn, dev_rows):
    W, D, steps = 256, 2, 59
    init, lp = seeded_init(port, "rosenbrock_2d", None, W, D, 5)
    s = make_sampler("rosenbrock_2d", [], W, D, deviceChainBytes=dev_rows * W * D * 8, maxChainSizeBytes=1 << 30)
    s.setInitialWalkerPos(init, lp)
    s.runMCMC(steps)                                  # dev_rows = 25: rows 0..49 have spilled to the host, 50..59 are in HBM
    o = port.run("rosenbrock_2d", None, init, lp, steps)
    full = np.concatenate([init[None], o["chain"]])
    n = full.shape[0]
    s.sliceAndBurnChain(interval, burn)               # Chain::resetChainForSubSampling (Chain.h:268-305)
    expect = full if (burn == 0 and interval == 1) else full[:0] if ((n - burn) < interval or n <= burn) else full[burn::interval]
    assert s.getStoredSteps() == expect.shape[0]
    assert np.array_equal(s.getChain(), expect)
    s.runMCMC(3)                                      # and the chain goes on behind the compacted rows
    o2 = port.run("rosenbrock_2d", None, o["pos"], o["logp"], 3, rng=o["rng"])
    assert np.array_equal(s.getChain(), np.concatenate([expect, o2["chain"]]))
    s.close()


def test_device_chain_analysis_matches_the_oracle(port):
    """CovarianceMatrix and AutoCorrCalc over the device-resident chain: the definitions of the reference classes
    (pinned in tests/test_analysis_cpu.py), nothing but the results crossing PCIe."""
    W, D, steps = 64, 5, 300
    iv = 1.0 / (1.0 + (np.arange(D) % 10))
    init, lp = seeded_init(port, "diag_gauss", iv, W, D, 53)
    s = make_sampler("diag_gauss", iv, W, D, deviceChainBytes=1 << 26, maxChainSizeBytes=0)
    s.setInitialWalkerPos(init, lp)
    assert s.runMCMC(steps)
    o = port.run("diag_gauss", iv, init, lp, steps)
    chain = np.concatenate([init[None], o["chain"]])
    mean, cov, corr, n = s.chainCovarianceDevice(3)
    om, oc, or_, on = port.covariance(chain, 3)
    assert n == on and np.allclose(mean, om, rtol=1e-12, atol=1e-14) and np.allclose(cov, oc, rtol=1e-11, atol=1e-13)
    assert np.allclose(corr, or_, rtol=1e-11, atol=1e-13)
    assert np.array_equal(s.chainAutoCorrTimesDevice(referenceBufferReuse=True), port.autocorr(chain, quirk=True))   # one CTA, reference order
    assert np.array_equal(s.chainAutoCorrTimesDevice(referenceBufferReuse=False), port.autocorr(chain, quirk=False))
    assert s.chainDeviceRows()[2] == steps + 1         # still on the device
    s.sliceAndBurnChain(2, 21)
    assert np.array_equal(s.chainAutoCorrTimesDevice(), port.autocorr(chain[21::2], quirk=False))
    s.close()


def test_device_chain_autocorr_walker_groups_many_walkers(port):
    """Clean mode, many walkers: one CTA per (parameter, walker group); same definition, compensated sums regrouped."""
    W, D, steps = 8192, 2, 127
    init, lp = seeded_init(port, "rosenbrock_2d", None, W, D, 6)
    s = make_sampler("rosenbrock_2d", [], W, D, deviceChainBytes=1 << 28, maxChainSizeBytes=0)
    s.setInitialWalkerPos(init, lp)
    assert s.runMCMC(steps)
    iat = s.chainAutoCorrTimesDevice()
    chain = s.getChain().copy()
    ref = port.autocorr(chain, quirk=False)
    assert np.allclose(iat, ref, rtol=1e-10, atol=0)
    s.close()


def test_c2_million_walker_iat_without_the_chain_crossing_pcie(port):
    """BASELINE configs[1] shape: 2^20 walkers, every 4th of 1024 steps kept in HBM (4.3 GB), burn-in dropped ON THE DEVICE,
    integrated autocorrelation times and covariance taken ON THE DEVICE; compared statistically with the oracle's
    AutoCorrCalc restatement on the first 4,096 walkers of the same chain (fetched for the test only)."""
    W, D, kept, every = 1 << 20, 2, 256, 4
    init, lp = seeded_init(port, "rosenbrock_2d", None, W, D, 2)
    s = make_sampler("rosenbrock_2d", [], W, D, rngMode="per_walker32", deviceChainBytes=(kept + 2) * W * D * 8, maxChainSizeBytes=0)
    s.setSlicingMode(True, every)
    s.setInitialWalkerPos(init, lp)
    assert s.runMCMC(kept)
    s.sliceAndBurnChain(1, 65)                        # 192 rows left
    ptr, first, rows = s.chainDeviceRows()
    assert (first, rows) == (0, kept + 1 - 65)
    iat = s.chainAutoCorrTimesDevice()
    mean, cov, corr, n = s.chainCovarianceDevice(1)
    assert n == rows * W and s.chainDeviceRows()[2] == rows           # nothing left the device
    assert np.all(np.isfinite(iat)) and np.all(np.abs(iat) > 0.5)
    full = s.getChain()                               # test only: now the rows cross PCIe
    ref = port.autocorr(full[:, :4096, :].copy(), quirk=False)
    assert np.allclose(np.abs(iat), np.abs(ref), rtol=0.15), (iat, ref)
    assert np.allclose(mean, full.reshape(-1, D).mean(axis=0), rtol=1e-9, atol=1e-12)      # the same 2e8 points
    s.close()


def test_device_log_matches_cuda_log():
    """device_rng.cuh logOfNormal (the log the pcg32 draws call on z and on 1-u) returns the bits of CUDA's log() on its whole
    domain of use -- positive normal doubles -- and is within 1 ulp of the host libm (the oracle's log)."""
    from mcmcpp_b200 import _capi
    lib = _capi.lib()
    rng = np.random.default_rng(20240607)
    n = 1 << 21
    u = rng.random(n)
    parts = [
        1.0 - u[: n // 4],                                             # 1 - u as the Metropolis test makes it
        (0.7071067811865475 * u[n // 4: n // 2] + 0.7071067811865476) ** 2,   # z = g(u), alpha = 2
        np.exp(rng.uniform(-700.0, 700.0, n // 4)),                      # the whole normal range
        rng.integers(1, 1 << 62, n // 4 - 8).astype(np.uint64).view(np.float64),
        np.array([1.0, 2.0 ** -53, np.nextafter(1.0, 0.0), np.nextafter(1.0, 2.0), 2.2250738585072014e-308,
                  1.7976931348623157e308, 0.5, 2.0]),
    ]
    x = np.ascontiguousarray(np.concatenate(parts))
    x = np.ascontiguousarray(x[np.isfinite(x) & (x >= 2.2250738585072014e-308)])
    mine = np.empty_like(x)
    ref = np.empty_like(x)
    _capi.check(lib.mcmcpp_device_log_check(0, x.ctypes.data, x.size, mine.ctypes.data, ref.ctypes.data))
    assert np.array_equal(mine.view(np.uint64), ref.view(np.uint64)), int((mine.view(np.uint64) != ref.view(np.uint64)).sum())
    host = np.log(x)
    ulp = np.abs(mine - host) / np.spacing(np.abs(host) + 1e-300)
    assert ulp.max() <= 1.0


# ---- the fused small-ensemble kernel with the draws made ahead (stretchFusedAhead) -----------------------------------------
@pytest.mark.parametrize("kind,W,D,calls,interval", [
    ("skewed_gauss_2d", 32, 2, (1, 31, 32, 33, 100), 1),        # B = 32 steps per window: less than, exactly, more than a window
    ("skewed_gauss_2d", 320, 2, (7, 2, 50), 1),                 # the reference's own test size: cluster of 4 CTAs, B = 8
    ("rosenbrock_2d", 44, 2, (45, 1, 23), 2),                   # B = 23, T = 1012: a ragged last warp; subsampling
    ("diag_gauss", 1024, 5, (3, 4), 1),                         # cluster, B = 2
    ("diag_gauss", 600, 8, (5, 5), 3),                          # cluster, B = 3; > 48 KB of shared memory
    ("dense_gauss", 12, 5, (70, 9), 1),                         # B = 82
    ("rosenbrock_2d", 256, 2, (9, 8, 1, 30), 1),                # from 256 walkers: a cluster of CTAs draws (B = 8, 4 CTAs)
    ("diag_gauss", 1000, 3, (5, 2, 6), 2)])                     # cluster, B = 2, 1000 walkers
def test_fused_kernel_with_draws_made_ahead(port, monkeypatch, kind, W, D, calls, interval):
    """Several runMCMC calls of awkward lengths (each call is one launch and leaves every walker's rng state behind for the next):
    chain, final state and counters equal to the oracle's."""
    params = {"skewed_gauss_2d": [0.13], "rosenbrock_2d": [], "diag_gauss": 1.0 / (1.0 + (np.arange(D) % 10)),
              "dense_gauss": np.linalg.inv(0.9 ** np.abs(np.subtract.outer(np.arange(D), np.arange(D)))).reshape(-1)}[kind]
    init, lp = seeded_init(port, kind, params if len(params) else None, W, D, 15)
    s = make_sampler(kind, params, W, D, seed=4)
    s.setSlicingMode(interval > 1, interval)
    s.setInitialWalkerPos(init, lp)
    for n in calls:
        assert s.runMCMC(n)
    assert s.getLaunchCount() == len(calls)
    assert s.getRngRedraws() == 0
    o = port.run(kind, params if len(params) else None, init, lp, sum(calls), interval=interval, seed=4)
    pos, aux = s.getWalkerState()
    assert np.array_equal(np.array(s.getChain())[1:], o["chain"]) and np.array_equal(pos, o["pos"]) and np.array_equal(aux, o["logp"])
    assert (s.getAcceptedSteps(), s.getTotalSteps()) == (int(o["acc"].sum()), int(o["tot"].sum()))
    s.close()


@pytest.mark.parametrize("rng", ["sequential", "per_walker32"])
def test_odd_rows_past_1024_parameters_without_a_trace(port, rng):
    """numParams odd and > 1024 (no bulk copies possible, too long for the register-resident warp kernel): the shared-memory
    kernel with plain loads; chain on, subsampling, no trace."""
    W, D, steps, interval = 2300, 1111, 2, 2
    iv = 1.0 / (1.0 + (np.arange(D) % 10))
    init, lp = seeded_init(port, "diag_gauss", iv, W, D, 8)
    s = make_sampler("diag_gauss", iv, W, D, seed=3, rngMode=rng)
    s.setSlicingMode(True, interval)
    s.setInitialWalkerPos(init, lp)
    assert s.runMCMC(steps)
    o = port.run("diag_gauss", iv, init, lp, steps, interval=interval, seed=3, per_walker=32 if rng == "per_walker32" else False)
    pos, aux = s.getWalkerState()
    assert np.array_equal(np.array(s.getChain())[1:], o["chain"]) and np.array_equal(pos, o["pos"]) and np.array_equal(aux, o["logp"])
    assert s.getAcceptedSteps() == int(o["acc"].sum())
    s.close()
