"""GPU tests of the sampler surface the reference's call sites use (EnsembleSampler.h:89-176):
chain cap / runMCMC's bool, reset, storeCurrentWalkerPositions, sliceAndBurnChain, argument checks,
and full-size runs of the BASELINE shapes checked through size-independent properties."""
import numpy as np
import pytest
from conftest import load_golden
from test_gpu_parity import make_sampler, seeded_init

pytestmark = pytest.mark.gpu


def test_argument_checks_mirror_the_reference_asserts():
    from mcmcpp_b200._capi import McmcppError, ERR_ARG
    for W, D in ((31, 2), (4, 2), (6, 3)):          # odd; W <= 2D (EnsembleSampler.h:207-208)
        with pytest.raises(McmcppError) as ei:
            make_sampler("diag_gauss", np.ones(D), W, D)
        assert ei.value.code == ERR_ARG
    s = make_sampler("skewed_gauss_2d", [0.13], 32, 2)
    with pytest.raises(McmcppError):
        s.runMCMC(5)                                   # before setInitialWalkerPos
    with pytest.raises(McmcppError):
        s.setSlicingMode(True, 0)                      # EnsembleSampler.h:327
    with pytest.raises(McmcppError):
        s.sliceAndBurnChain(0, 0)                      # EnsembleSampler.h:335
    s.close()


def test_chain_cap_returns_false_like_runMCMC(port):
    g = load_golden("c1_skewed_w32")
    rows = 11
    s = make_sampler(g["kind"], g["params"], g["W"], g["D"], maxChainSizeBytes=rows * g["W"] * g["D"] * 8)
    s.setInitialWalkerPos(g["init"], g["logp"])
    assert s.runMCMC(5) is True
    assert s.runMCMC(50) is False                      # cap reached (EnsembleSampler.h:293)
    assert s.getStoredSteps() == rows and s.getChain().shape[0] == rows
    o = port.run(g["kind"], g["params"], g["init"], g["logp"], rows - 1)
    assert np.array_equal(s.getChain()[1:], o["chain"])
    s.close()


def test_reset_store_current_and_counters(port):
    g = load_golden("c1_skewed_w32")
    W = g["W"]
    s = make_sampler(g["kind"], g["params"], W, g["D"])
    s.setInitialWalkerPos(g["init"], g["logp"])
    assert s.getAcceptedSteps() == W and s.getTotalSteps() == W        # Walker.h:162-170
    s.runMCMC(40)
    o = port.run(g["kind"], g["params"], g["init"], g["logp"], 40)
    s.reset()                                                           # EnsembleSampler.h:312-322
    assert s.getStoredSteps() == 0 and s.getTotalSteps() == 0 and s.getAcceptedSteps() == 0
    s.storeCurrentWalkerPositions()                                     # EnsembleSampler.h:232-242
    assert s.getStoredSteps() == 1 and np.array_equal(s.getChain()[0], o["pos"])
    s.runMCMC(10)
    o2 = port.run(g["kind"], g["params"], o["pos"], o["logp"], 10, rng=o["rng"], acc=np.zeros(W), tot=np.zeros(W))
    assert np.array_equal(s.getChain()[1:], o2["chain"])
    assert (s.getAcceptedSteps(), s.getTotalSteps()) == (int(o2["acc"].sum()), 10 * W)
    s.close()


@pytest.mark.parametrize("interval,burn", [(1, 0), (3, 5), (7, 0), (1, 20), (50, 10), (2, 59), (5, 80)])
def test_slice_and_burn_matches_chain_reset_semantics(interval, burn):
    g = load_golden("c1_skewed_w32")
    s = make_sampler(g["kind"], g["params"], g["W"], g["D"])
    s.setInitialWalkerPos(g["init"], g["logp"])
    s.runMCMC(59)
    full = s.getChain().copy()
    n = full.shape[0]
    s.sliceAndBurnChain(interval, burn)
    if burn == 0 and interval == 1:
        expect = full
    elif (n - burn) < interval or n <= burn:           # Chain.h:272-273 -> resetChain
        expect = full[:0]
    else:
        expect = full[burn::interval]
    assert s.getStoredSteps() == expect.shape[0]
    assert np.array_equal(s.getChain(), expect)
    s.close()


def test_c2_full_size_rosenbrock_million_walkers(port):
    """BASELINE config 2 at full size: 2^20 walkers; the oracle finishes 3 steps in seconds."""
    W, D = 1 << 20, 2
    init, lp = seeded_init(port, "rosenbrock_2d", None, W, D, 2)
    s = make_sampler("rosenbrock_2d", [], W, D, maxChainSizeBytes=0)
    s.setInitialWalkerPos(init, lp)
    assert s.runMCMC(3)
    o = port.run("rosenbrock_2d", None, init, lp, 3, want_chain=False)
    pos, aux = s.getWalkerState()
    assert np.array_equal(pos, o["pos"]) and np.array_equal(aux, o["logp"])
    assert s.getAcceptedSteps() == int(o["acc"].sum()) and s.getTotalSteps() == 4 * W
    assert s.getChain().shape[0] == 0                  # chain off
    s.close()


def test_c4_shape_properties_at_65536_walkers(port):
    """BASELINE config 4 row shape (D=1000) at a size the oracle checks directly, plus properties that hold
    at any size: logp[w] == logp(pos[w]); a rejected walker's row is untouched; counters add up."""
    W, D = 1 << 16, 1000
    iv = 1.0 / (1.0 + (np.arange(D) % 10))
    rng = np.random.default_rng(4)
    init = rng.standard_normal((W, D))
    lp = port.logp("diag_gauss", iv, init)
    s = make_sampler("diag_gauss", iv, W, D, maxChainSizeBytes=0, traceUpdates=W)
    s.setInitialWalkerPos(init, lp)
    assert s.runMCMC(1)
    pos, aux = s.getWalkerState()
    o = port.run("diag_gauss", iv, init, lp, 1, want_chain=False, want_trace=True)
    assert np.array_equal(pos, o["pos"]) and np.array_equal(aux, o["logp"])
    tr = s.getTrace(W)
    assert np.array_equal(tr["accept"], o["trace"]["accept"]) and np.array_equal(tr["j"], o["trace"]["j"])
    rejected = tr["accept"] == 0
    assert np.array_equal(pos[rejected], init[rejected])
    assert np.array_equal(aux, port.logp("diag_gauss", iv, pos))
    assert s.getAcceptedSteps() == W + int(tr["accept"].sum()) and s.getTotalSteps() == 2 * W
    s.close()


@pytest.mark.parametrize("kind,W,D,rng", [("skewed_gauss_2d", 512, 2, "sequential"), ("diag_gauss", 256, 100, "per_walker"),
                                          ("diag_gauss", 512, 200, "sequential")])
def test_checkpoint_resume_continues_bit_for_bit(port, tmp_path, kind, W, D, rng):
    """save after 7 steps, restore into a fresh sampler, run 9 more: identical to an uninterrupted 16-step run
    (positions, log-probs, counters) in both RNG disciplines and for the thread / warp / TMA kernels."""
    import mcmcpp_b200 as m
    params = np.array([0.13]) if kind == "skewed_gauss_2d" else 1.0 / (1.0 + (np.arange(D) % 10))
    calc = (lambda: m.SkewedGaussianTwoDim(0.13)) if kind == "skewed_gauss_2d" else (lambda: m.DiagGaussian(params))
    init = np.random.default_rng(D).standard_normal((W, D))
    lp = port.logp(kind, params, init)
    a = m.EnsembleSampler(11, W, D, m.StretchMove(D, 11, calc()), rngMode=rng)
    a.setInitialWalkerPos(init, lp)
    assert a.runMCMC(16)
    b = m.EnsembleSampler(11, W, D, m.StretchMove(D, 11, calc()), rngMode=rng)
    b.setInitialWalkerPos(init, lp)
    assert b.runMCMC(7)
    blob = b.saveCheckpoint()
    b.writeChain(str(tmp_path / "chain.npy"))
    assert np.array_equal(np.load(str(tmp_path / "chain.npy")), b.getChain())
    b.close()
    c = m.EnsembleSampler(11, W, D, m.StretchMove(D, 11, calc()), rngMode=rng)
    c.loadCheckpoint(blob)
    assert c.runMCMC(9)
    pa, la = a.getWalkerState()
    pc, lc = c.getWalkerState()
    assert np.array_equal(pa, pc) and np.array_equal(la, lc)
    assert (a.getAcceptedSteps(), a.getTotalSteps()) == (c.getAcceptedSteps(), c.getTotalSteps())
    assert np.array_equal(a.getChain()[8:], c.getChain())          # rows of steps 8..16
    d = m.EnsembleSampler(12, W, D, m.StretchMove(D, 12, calc()), rngMode=rng)
    from mcmcpp_b200._capi import McmcppError
    with pytest.raises(McmcppError):
        d.loadCheckpoint(blob)                                        # other seed -> refused
    for s in (a, c, d):
        s.close()


def test_device_step_action_sees_every_stored_step(port):
    """PostStepAction (EnsembleSampler.h:356-359) with a device view: called once per stored step, in order, with the
    state pointer and the sampler's stream; a copy enqueued on that stream from inside the action captures exactly the
    row the chain stores for that step."""
    import torch
    import mcmcpp_b200 as m
    W, D, steps = 256, 2, 6
    init = np.random.default_rng(3).standard_normal((W, D))
    lp = port.logp("skewed_gauss_2d", np.array([0.13]), init)
    s = m.EnsembleSampler(2, W, D, m.StretchMove(D, 2, m.SkewedGaussianTwoDim(0.13)))
    s.setSlicingMode(True, 2)                       # 2 ensemble steps per stored step: the action fires on stored ones only
    s.setInitialWalkerPos(init, lp)
    seen, snaps = [], torch.zeros((steps, W, D), dtype=torch.float64, device="cuda")
    import ctypes
    cudart = ctypes.CDLL("libcudart.so.12")          # the runtime torch already loaded
    cudart.cudaMemcpyAsync.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_void_p]

    def action(row, step, stream):
        seen.append(step)
        dst = snaps[len(seen) - 1].data_ptr()
        assert cudart.cudaMemcpyAsync(dst, row, W * D * 8, 3, stream) == 0        # 3 = cudaMemcpyDeviceToDevice

    s.setDeviceStepAction(action)
    assert s.runMCMC(steps)
    torch.cuda.synchronize()
    assert seen == list(range(1, steps + 1))
    assert np.array_equal(snaps.cpu().numpy(), s.getChain()[1:])
    s.setDeviceStepAction(None)
    assert s.runMCMC(2) and len(seen) == steps
    s.close()
