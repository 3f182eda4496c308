"""Python mirror of the reference's sampler front-ends for the ensemble-update hot path.

Same names, argument meaning and error behaviour as
  MCMC::EnsembleSampler           (MCMCpp/EnsembleSampler.h:39-197)
  MCMC::ParallelEnsembleSampler   (MCMCpp/ParallelEnsembleSampler.h:78-226)
  MCMC::Mover::StretchMove        (MCMCpp/Movers/StretchMove.h:40-131)
so the parity tests read like the reference's own test programs
(test/sequential/SkewedGaussian/StretchMove/src/main.cpp:22-80).  Everything below the method
names is a call into the C-ABI (include/mcmcpp_b200/capi.h); nothing here computes on the CPU.
The C++ drop-in with the same names is include/mcmcpp_b200/EnsembleSampler.h.
"""
import ctypes as C
import numpy as np

from . import _capi
from ._capi import check


# ---- Calculators (include/mcmcpp_b200/targets.h); params layout of the C-ABI built-ins ---------
class _Calculator:
    kind = None

    def params(self):
        return np.zeros(0)


class SkewedGaussianTwoDim(_Calculator):
    """test/sequential/SkewedGaussian/Common/SkewedGaussian.h:52-57"""
    kind = "skewed_gauss_2d"

    def __init__(self, eps):
        self.eps = float(eps)

    def params(self):
        return np.array([self.eps])


class RosenbrockTwoDim(_Calculator):
    kind = "rosenbrock_2d"


class DiagGaussian(_Calculator):
    kind = "diag_gauss"

    def __init__(self, inv_var):
        self.iv = np.ascontiguousarray(inv_var, dtype=np.float64)

    def params(self):
        return self.iv


class DenseGaussian(_Calculator):
    kind = "dense_gauss"

    def __init__(self, precision):
        self.P = np.ascontiguousarray(precision, dtype=np.float64)

    def params(self):
        return self.P.reshape(-1)


class LinearRegression(_Calculator):
    kind = "linreg"

    def __init__(self, X, y, prior_var):
        self.X = np.ascontiguousarray(X, dtype=np.float64)
        self.y = np.ascontiguousarray(y, dtype=np.float64)
        self.prior_var = float(prior_var)

    def params(self):
        M = self.X.shape[0]
        return np.concatenate([[float(M), self.prior_var], self.X.reshape(-1), self.y])


class StretchMove:
    """MCMC::Mover::StretchMove<ParamType, Calculator, GwDistribution<ParamType, alphaNum, alphaDenom>>.
    As in the reference (StretchMove.h:74-78, EnsembleSampler.h:217) `prngInit` is not what seeds the
    run: the sampler re-seeds its copy of the mover with (randSeed, stream)."""

    def __init__(self, numParams, prngInit, calc, alphaNum=2, alphaDenom=1):
        self.numParams = int(numParams)
        self.prngInit = int(prngInit)
        self.calc = calc
        self.alpha = (int(alphaNum), int(alphaDenom))


class DifferentialEvolution:
    """MCMC::Mover::DifferentialEvolution<ParamType, Calculator> (Movers/DifferentialEvolution.h:40-156): same constructor
    and tuning calls.  Its draw count per update varies, so the device runs it with one pcg64 stream per walker."""

    def __init__(self, numParams, prngInit, calc):
        self.numParams = int(numParams)
        self.prngInit = int(prngInit)
        self.calc = calc
        self.alpha = (2, 1)
        self.gamma = 0.0          # <= 0: 2.38 / sqrt(2 numParams)
        self.rand = 0.0           # <= 0: 1e-4

    def overrideGamma(self, newGamma):
        self.gamma = float(newGamma)

    def overrideRandBounds(self, newRand):
        self.rand = float(newRand)


class WalkMove:
    """MCMC::Mover::WalkMove<ParamType, Calculator>(numParams, prngInit, calc, numSamples) (Movers/WalkMove.h:40-60):
    numSamples partner walkers of the other colour per proposal.  One pcg64 stream per walker on the device."""

    def __init__(self, numParams, prngInit, calc, numSamples):
        self.numParams = int(numParams)
        self.prngInit = int(prngInit)
        self.calc = calc
        self.alpha = (2, 1)
        self.numPoints = int(numSamples)


class EnsembleSampler:
    """MCMC::EnsembleSampler<double, StretchMove<...>> on one B200 (or one shard of an ensemble
    spread over `world` GPUs, one process per GPU).

    Extra keyword arguments have no counterpart in the reference and default to its behaviour:
      rngMode      "sequential" (bit parity with the reference's single stream) | "per_walker" (pcg64(seed, w)) |
                   "per_walker32" (pcg32(seed, w), the fast engine; stretch move only)
      chainBlockSteps  0: the chain is capped at floor(maxChainSizeBytes / row bytes) rows; 10000: the reference's
                   whole-block granularity (Chain/Chain.h:178), runMCMC returns False exactly where the reference does
      deviceChainBytes > 0: stored steps stay in HBM (see capi.h); getChain() brings them to the host on demand,
                   chainCovarianceDevice()/chainAutoCorrTimesDevice() analyse them in place
      traceUpdates record the first n updates' (j, accept, z, e, newp, diff) for tests
      device/rank/world/shardMode  multi-GPU placement (see capi.h)
    """

    def __init__(self, randSeed, numWalker, numParameter, move, maxChainSizeBytes=2147483648, stepAct=None,
                 rngMode="sequential", traceUpdates=0, device=0, rank=0, world=1, shardMode="none", chainBlockSteps=0, deviceChainBytes=0):
        if not isinstance(move, (StretchMove, DifferentialEvolution, WalkMove)):
            raise TypeError("the B200 path implements Mover::StretchMove, Mover::DifferentialEvolution and Mover::WalkMove")
        if isinstance(move, (DifferentialEvolution, WalkMove)):
            rngMode = "per_walker"                      # variable draws per update: no sequential-stream indexing
        if move.numParams != numParameter:
            raise ValueError("mover and sampler disagree on numParams")
        self._lib = _capi.lib()
        self.numWalkers, self.numParams = int(numWalker), int(numParameter)
        self.stepAction = stepAct
        self.rank, self.world = int(rank), int(world)
        cfg = _capi.Config(
            num_walkers=self.numWalkers, num_params=self.numParams, seed=int(randSeed),
            rng_mode={"sequential": _capi.RNG_SEQUENTIAL, "per_walker": _capi.RNG_PER_WALKER, "per_walker32": _capi.RNG_PER_WALKER32}[rngMode],
            alpha_num=move.alpha[0], alpha_denom=move.alpha[1], device=int(device), rank=self.rank, world=self.world,
            shard_mode={"none": _capi.SHARD_NONE, "allgather": _capi.SHARD_ALLGATHER, "peer": _capi.SHARD_PEER,
                        "data": _capi.SHARD_DATA, "dim": _capi.SHARD_DIM}[shardMode],
            trace_updates=int(traceUpdates), max_chain_bytes=int(maxChainSizeBytes), chain_block_steps=int(chainBlockSteps),
            device_chain_bytes=int(deviceChainBytes))
        self.shardMode = shardMode
        params = np.ascontiguousarray(move.calc.params(), dtype=np.float64)
        self._h = C.c_void_p()
        # the reference asserts on bad walker counts (EnsembleSampler.h:207-208); here: ERR_ARG -> exception
        if isinstance(move, (DifferentialEvolution, WalkMove)):
            de = isinstance(move, DifferentialEvolution)
            mv = np.array([move.gamma, move.rand] if de else [float(move.numPoints)], dtype=np.float64)
            check(self._lib.mcmcpp_sampler_create_builtin_mover(C.byref(cfg), _capi.MOVER_DIFF_EVOLUTION if de else _capi.MOVER_WALK,
                                                                mv.ctypes.data_as(C.POINTER(C.c_double)), mv.size,
                                                                _capi.TARGET[move.calc.kind],
                                                                params.ctypes.data_as(C.POINTER(C.c_double)), params.size,
                                                                C.byref(self._h)))
        else:
            check(self._lib.mcmcpp_sampler_create_builtin(C.byref(cfg), _capi.TARGET[move.calc.kind],
                                                          params.ctypes.data_as(C.POINTER(C.c_double)), params.size,
                                                          C.byref(self._h)))
        self.lastDeviceMs = 0.0

    # -- life cycle --------------------------------------------------------------------------
    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            self._lib.mcmcpp_sampler_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- EnsembleSampler.h:220-230 -----------------------------------------------------------
    def setInitialWalkerPos(self, positions, auxValues):
        pos = np.ascontiguousarray(positions, dtype=np.float64).reshape(self.numWalkers, self.numParams)
        aux = np.ascontiguousarray(auxValues, dtype=np.float64).reshape(self.numWalkers)
        check(self._lib.mcmcpp_sampler_set_initial(self._h, pos.ctypes.data, aux.ctypes.data))

    def setInitialWalkerPosDevice(self, positionsPtr, auxPtr):
        """Device-resident inputs (raw device pointers, e.g. torch tensor .data_ptr())."""
        check(self._lib.mcmcpp_sampler_set_initial_device(self._h, positionsPtr, auxPtr))

    def setInitialWalkerPosPinned(self, positionsPtr, auxPtr):
        """Host inputs given as raw pointers (e.g. pinned torch tensors)."""
        check(self._lib.mcmcpp_sampler_set_initial(self._h, positionsPtr, auxPtr))

    def setInitialWalkerPosLocal(self, positionsPtr, auxPtr):
        """DIM sharding: this rank's columns only, contiguous [W][D_local] host buffer given as a raw pointer."""
        check(self._lib.mcmcpp_sampler_set_initial_local(self._h, positionsPtr, auxPtr))

    def getWalkerStateLocalInto(self, positionsPtr, auxPtr):
        check(self._lib.mcmcpp_sampler_get_state_local(self._h, positionsPtr, auxPtr))

    # -- EnsembleSampler.h:232-242, 312-322 --------------------------------------------------
    def storeCurrentWalkerPositions(self):
        check(self._lib.mcmcpp_sampler_store_current(self._h))

    def reset(self):
        check(self._lib.mcmcpp_sampler_reset(self._h))

    # -- EnsembleSampler.h:284-310 ------------------------------------------------------------
    def runMCMC(self, numSteps):
        """True if the sampling went to the end before the max chain size was reached."""
        ms = C.c_float(0.0)
        rc = check(self._lib.mcmcpp_sampler_run(self._h, int(numSteps), C.byref(ms)))
        self.lastDeviceMs = float(ms.value)
        if self.stepAction is not None:
            # end-of-call hook (the per-step hook of EnsembleSampler.h:356-359 stays on the host)
            self.stepAction.performAction(self.getChain())
        return rc != _capi.CHAIN_FULL

    # -- EnsembleSampler.h:244-282 ------------------------------------------------------------
    def _counts(self):
        a, t = C.c_ulonglong(), C.c_ulonglong()
        check(self._lib.mcmcpp_sampler_get_counts(self._h, C.byref(a), C.byref(t)))
        return a.value, t.value

    def getAcceptedSteps(self):
        return self._counts()[0]

    def getTotalSteps(self):
        return self._counts()[1]

    def getAcceptanceFraction(self):
        a, t = self._counts()
        return float(a) / float(t)

    # -- EnsembleSampler.h:324-339, 150 --------------------------------------------------------
    def setSlicingMode(self, useSlicing=False, slicingInterval=1):
        check(self._lib.mcmcpp_sampler_set_slicing(self._h, int(bool(useSlicing)), int(slicingInterval)))

    def sliceAndBurnChain(self, slicingInterval, burnIn):
        check(self._lib.mcmcpp_sampler_slice_and_burn(self._h, int(slicingInterval), int(burnIn)))

    def getStoredSteps(self):
        return self._lib.mcmcpp_sampler_get_stored_steps(self._h)

    # -- read side: what ChainStepIterator / ChainPsetIterator walk (ChainStepIterator.h:125) --
    def getChain(self):
        """[stored][W][D] view of the pinned host chain (valid until reset/slice/run/close)."""
        rows, n = C.POINTER(C.c_double)(), C.c_longlong()
        check(self._lib.mcmcpp_sampler_chain(self._h, C.byref(rows), C.byref(n)))
        if n.value == 0 or not rows:
            return np.zeros((0, self.numWalkers, self.numParams))
        return np.ctypeslib.as_array(rows, shape=(n.value, self.numWalkers, self.numParams))

    def getStepIttBegin(self):
        return iter(self.getChain())

    def getParamSetIttBegin(self):
        return iter(self.getChain().reshape(-1, self.numParams))

    # -- walker state (Walker::getCurrState / getCurrAuxData) ---------------------------------
    def getWalkerState(self):
        # DIM sharding: only this rank's columns are written (parallel.dim_slice); the others read 0
        pos = np.zeros((self.numWalkers, self.numParams))
        aux = np.empty(self.numWalkers)
        check(self._lib.mcmcpp_sampler_get_state(self._h, pos.ctypes.data, aux.ctypes.data))
        return pos, aux

    def getWalkerStateInto(self, positionsPtr, auxPtr):
        """Same, into caller-owned host buffers given as raw pointers (e.g. pinned torch tensors)."""
        check(self._lib.mcmcpp_sampler_get_state(self._h, positionsPtr, auxPtr))

    # -- Analysis::CovarianceMatrix accumulated on the device while sampling (chain may be off) -----
    def covBegin(self, sliceInterval=1, includeCurrent=True):
        check(self._lib.mcmcpp_sampler_cov_begin(self._h, int(sliceInterval), int(bool(includeCurrent))))

    def covResult(self):
        """(mean[D], cov[D,D], corr[D,D], points) of the steps stored since covBegin."""
        D = self.numParams
        mean, cov, corr, n = np.empty(D), np.empty((D, D)), np.empty((D, D)), C.c_longlong()
        check(self._lib.mcmcpp_sampler_cov_result(self._h, mean.ctypes.data, cov.ctypes.data, corr.ctypes.data, C.byref(n)))
        return mean, cov, corr, n.value

    def covEnd(self):
        check(self._lib.mcmcpp_sampler_cov_end(self._h))

    # -- analysis over the device-resident chain (deviceChainBytes > 0): nothing crosses PCIe but the results ---------------
    def chainDeviceRows(self):
        """(device pointer, first chain index, rows) of the stored steps currently held in HBM."""
        ptr, first, rows = C.c_void_p(), C.c_longlong(), C.c_longlong()
        check(self._lib.mcmcpp_sampler_chain_device(self._h, C.byref(ptr), C.byref(first), C.byref(rows)))
        return ptr.value, first.value, rows.value

    def chainCovarianceDevice(self, sliceInterval=1):
        """Analysis::CovarianceMatrix::calculateCovar over the device-resident chain -> (mean, cov, corr, points)."""
        D = self.numParams
        mean, cov, corr, n = np.empty(D), np.empty((D, D)), np.empty((D, D)), C.c_longlong()
        check(self._lib.mcmcpp_sampler_chain_covariance(self._h, int(sliceInterval), mean.ctypes.data, cov.ctypes.data, corr.ctypes.data, C.byref(n)))
        return mean, cov, corr, n.value

    def chainAutoCorrTimesDevice(self, windowScaling=4, referenceBufferReuse=False):
        """Analysis::AutoCorrCalc::calcAutoCorrTimes (all walkers) over the device-resident chain -> iat[D]."""
        iat = np.empty(self.numParams)
        check(self._lib.mcmcpp_sampler_chain_autocorr_times(self._h, int(windowScaling), int(bool(referenceBufferReuse)), iat.ctypes.data))
        return iat

    # -- PostStepAction with a device-resident step view (EnsembleSampler.h:356-359) --------------------------
    def setDeviceStepAction(self, fn):
        """fn(devRowPtr, stepIndex, cudaStreamPtr) is called on the host after every stored step, stream-ordered with
        the update kernels (enqueue work on the stream to look at the step in HBM); None removes it."""
        if fn is None:
            self._stepThunk = None
            check(self._lib.mcmcpp_sampler_set_step_action(self._h, None, None))
            return
        self._stepThunk = _capi.STEP_FN(lambda ctx, row, step, stream: fn(row, step, stream))
        check(self._lib.mcmcpp_sampler_set_step_action(self._h, C.cast(self._stepThunk, C.c_void_p), None))

    # -- checkpoint / resume and a minimal chain writer ------------------------------
    def saveCheckpoint(self):
        """bytes holding positions, log-probs, RNG states, step index and counters of this rank."""
        n = C.c_size_t()
        check(self._lib.mcmcpp_sampler_checkpoint_bytes(self._h, C.byref(n)))
        buf = np.empty(n.value, dtype=np.uint8)
        check(self._lib.mcmcpp_sampler_checkpoint_save(self._h, buf.ctypes.data, n.value))
        return buf.tobytes()

    def loadCheckpoint(self, blob):
        buf = np.frombuffer(blob, dtype=np.uint8)
        check(self._lib.mcmcpp_sampler_checkpoint_restore(self._h, buf.ctypes.data, buf.size))

    def writeChain(self, path):
        """Stored rows [stored][W][D] as a .npy file (float64, C order)."""
        np.save(path, self.getChain())

    # -- diagnostics -----------------------------------------------------------------------------
    def getTrace(self, maxRows):
        buf = (_capi.Trace * int(maxRows))()
        n = C.c_int()
        check(self._lib.mcmcpp_sampler_get_trace(self._h, buf, int(maxRows), C.byref(n)))
        dt = np.dtype([("j", "<i4"), ("accept", "<i4"), ("z", "<f8"), ("e", "<f8"), ("newp", "<f8"), ("diff", "<f8")])
        return np.frombuffer(buf, dtype=dt, count=n.value).copy()

    def getRngRedraws(self):
        n = C.c_int()
        check(self._lib.mcmcpp_sampler_rng_redraws(self._h, C.byref(n)))
        return n.value

    def getLaunchCount(self):
        n = C.c_ulonglong()
        check(self._lib.mcmcpp_sampler_launch_count(self._h, C.byref(n)))
        return n.value

    # -- multi-GPU plumbing (torch.distributed carries the handles; see parallel.py) -------------
    def commInit(self, uniqueId):
        check(self._lib.mcmcpp_sampler_comm_init(self._h, uniqueId))

    def peerExport(self):
        buf = C.create_string_buffer(64)
        check(self._lib.mcmcpp_sampler_peer_export(self._h, buf))
        return buf.raw

    def peerImport(self, handles):
        check(self._lib.mcmcpp_sampler_peer_import(self._h, b"".join(handles)))


class ParallelEnsembleSampler(EnsembleSampler):
    """MCMC::ParallelEnsembleSampler (ParallelEnsembleSampler.h:119-120): same surface plus a
    thread count.  On the GPU every walker of a colour is updated concurrently, so `threadCount`
    only selects the RNG discipline: 1 -> the sequential stream (the reference is bit-identical to
    the sequential sampler with one thread), >1 -> one stream per walker (the reference is
    non-deterministic there, ParallelEnsembleSampler.h:71-76)."""

    def __init__(self, randSeed, threadCount, numWalker, numParameter, move, maxChainSizeBytes=2147483648, stepAct=None,
                 **kw):
        kw.setdefault("rngMode", "sequential" if int(threadCount) <= 1 else "per_walker")
        super().__init__(randSeed, numWalker, numParameter, move, maxChainSizeBytes, stepAct, **kw)
        self.threadCount = int(threadCount)

    def setSamplingMode(self, subSamplingInt=1, burnIn=0):
        """ParallelEnsembleSampler.h:178,322-330: interval for later runs + thin/burn what is already stored.
        Later runs use the sequential sampler's subsampling (the parallel controller's own is broken)."""
        self.setSlicingMode(subSamplingInt > 1, subSamplingInt)
        self.sliceAndBurnChain(subSamplingInt, burnIn)
