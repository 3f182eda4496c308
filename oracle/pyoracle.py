"""ctypes loaders for the CPU oracle (TEST INFRASTRUCTURE, never the product path).

  Port    -- oracle/liboracle.so, the plain-C restatement (oracle/oracle.c)
  Ref     -- oracle/_ref/libmcmcpp_ref*.so, the UNMODIFIED reference headers
             (oracle/ref_driver.cpp); present only if built where /root/reference exists
"""
import ctypes as C
import os
import subprocess
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
KIND = {"skewed_gauss_2d": 0, "rosenbrock_2d": 1, "diag_gauss": 2, "dense_gauss": 3, "linreg": 4}

c_u64 = C.c_uint64
c_dp = C.POINTER(C.c_double)


def _dp(a):
    return a.ctypes.data_as(c_dp) if a is not None else None


def build():
    """(Re)build liboracle.so and, when /root/reference is present, oracle/_ref."""
    subprocess.run(["make", "-s", "-C", HERE], check=True, stdout=subprocess.DEVNULL)


class Trace(C.Structure):
    _fields_ = [("j", C.c_int32), ("accept", C.c_int32), ("z", C.c_double), ("e", C.c_double),
                ("newp", C.c_double), ("diff", C.c_double)]


TRACE_DTYPE = np.dtype([("j", "<i4"), ("accept", "<i4"), ("z", "<f8"), ("e", "<f8"), ("newp", "<f8"), ("diff", "<f8")])


class Pcg64(C.Structure):
    _fields_ = [("state_lo", c_u64), ("state_hi", c_u64), ("inc_lo", c_u64), ("inc_hi", c_u64)]


class Port:
    """oracle/oracle.c"""

    def __init__(self):
        path = os.path.join(HERE, "liboracle.so")
        if not os.path.exists(path):
            build()
        L = self.lib = C.CDLL(path)
        L.oracle_pcg64_next.restype = c_u64
        L.oracle_pcg64_bounded.restype = c_u64
        L.oracle_pcg64_bounded.argtypes = [C.c_void_p, c_u64, C.POINTER(C.c_int)]
        L.oracle_pcg64_seed.argtypes = [C.c_void_p, c_u64, c_u64]
        L.oracle_pcg64_advance_u64.argtypes = [C.c_void_p, c_u64]
        L.oracle_pcg64_backstep_u64.argtypes = [C.c_void_p, c_u64]
        L.oracle_pcg32_seed.argtypes = [C.c_void_p, c_u64]
        L.oracle_pcg32_seed_stream.argtypes = [C.c_void_p, c_u64, c_u64]
        L.oracle_pcg32_next.restype = C.c_uint32
        L.oracle_u64_to_unit.restype = C.c_double
        L.oracle_u64_to_unit.argtypes = [c_u64]
        L.oracle_gw_sample.restype = C.c_double
        L.oracle_gw_sample.argtypes = [C.c_double] * 3
        L.oracle_neg_exponential.restype = C.c_double
        L.oracle_neg_exponential.argtypes = [C.c_double]
        L.oracle_logp.restype = C.c_double
        L.oracle_logp.argtypes = [C.c_int, C.c_int, c_dp, c_dp]
        L.oracle_stretch_run.restype = C.c_int
        L.oracle_stretch_run.argtypes = [C.c_void_p, C.c_int, c_dp, C.c_int, C.c_int, C.c_int, C.c_int,
                                         c_dp, c_dp, C.c_void_p, C.c_void_p, C.c_int, C.c_int, c_dp, C.c_void_p]

        L.oracle_covariance.restype = C.c_long
        L.oracle_covariance.argtypes = [c_dp, C.c_long, C.c_int, C.c_int, C.c_int, c_dp, c_dp, c_dp]
        L.oracle_de_run.restype = None
        L.oracle_de_run.argtypes = [C.c_void_p, C.c_int, C.c_int, c_dp, C.c_int, C.c_int, C.c_double, C.c_double, C.c_double,
                                    c_dp, c_dp, C.c_void_p, C.c_void_p, C.c_int, C.c_int, c_dp, C.c_void_p]
        L.oracle_walk_run.restype = None
        L.oracle_walk_run.argtypes = [C.c_void_p, c_dp, C.c_int, C.c_int, c_dp, C.c_int, C.c_int, C.c_int,
                                      c_dp, c_dp, C.c_void_p, C.c_void_p, C.c_int, C.c_int, c_dp, C.c_void_p]
        L.oracle_autocorr.restype = None
        L.oracle_autocorr.argtypes = [c_dp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, c_dp, c_dp]
        L.oracle_stretch_run_per_walker.restype = C.c_int
        L.oracle_stretch_run_per_walker.argtypes = L.oracle_stretch_run.argtypes
        L.oracle_pcg64_seed_many.argtypes = [C.c_void_p, c_u64, C.c_int]
        L.oracle_stretch_run_per_walker32.restype = C.c_int
        L.oracle_stretch_run_per_walker32.argtypes = L.oracle_stretch_run.argtypes
        L.oracle_pcg32_seed_many.argtypes = [C.c_void_p, c_u64, C.c_int]
        L.oracle_logp_many.argtypes = [C.c_int, C.c_int, c_dp, c_dp, C.c_long, c_dp]
        L.oracle_stretch_half_slots.restype = None
        L.oracle_stretch_half_slots.argtypes = [c_u64, C.c_int, c_dp, C.c_int, C.c_int, C.c_int, C.c_int, c_dp, c_dp,
                                                C.c_void_p, C.c_long, C.c_int, C.c_int, C.c_int]

    # --- rng -----------------------------------------------------------
    def autocorr(self, chain, window_scaling=4, quirk=True, want_rho=False):
        """AutoCorrCalc::calcAutoCorrTimes over chain[S][W][D], all walkers -> iat[D] (and rho[D,S]).  quirk=True is the
        reference as written (transferWalker accumulates into the un-cleared buffer); False the clean definition."""
        chain = np.ascontiguousarray(chain, dtype=np.float64)
        S, W, D = chain.shape
        iat = np.empty(D)
        rho = np.empty((D, S)) if want_rho else None
        self.lib.oracle_autocorr(_dp(chain), S, W, D, int(window_scaling), int(bool(quirk)), _dp(iat), _dp(rho) if want_rho else None)
        return (iat, rho) if want_rho else iat

    def covariance(self, chain, interval=1):
        """CovarianceMatrix::calculateCovar over chain[S][W][D] -> (mean[D], cov[D,D], corr[D,D], points)."""
        chain = np.ascontiguousarray(chain, dtype=np.float64)
        S, W, D = chain.shape
        mean, cov, corr = np.empty(D), np.empty((D, D)), np.empty((D, D))
        n = self.lib.oracle_covariance(_dp(chain), S, W, D, int(interval), _dp(mean), _dp(cov), _dp(corr))
        return mean, cov, corr, int(n)

    def pcg64(self, seed, stream):
        g = Pcg64()
        self.lib.oracle_pcg64_seed(C.byref(g), seed & (2**64 - 1), stream & (2**64 - 1))
        return g

    def pcg64_raw(self, seed, stream, n, skip=0):
        g = self.pcg64(seed, stream)
        if skip:
            self.lib.oracle_pcg64_advance_u64(C.byref(g), skip)
        return np.array([self.lib.oracle_pcg64_next(C.byref(g)) for _ in range(n)], dtype=np.uint64)

    def pcg32_raw(self, seed, stream, n):
        g = (c_u64 * 2)()
        if stream is None:
            self.lib.oracle_pcg32_seed(g, seed)
        else:
            self.lib.oracle_pcg32_seed_stream(g, seed, stream)
        return np.array([self.lib.oracle_pcg32_next(g) for _ in range(n)], dtype=np.uint32)

    def draws(self, seed, stream, K, n, alpha=(2, 1)):
        """(j, z, log(1-u)) in StretchMove order for n updates."""
        g = self.pcg64(seed, stream)
        t1, isa = C.c_double(), C.c_double()
        self.lib.oracle_gw_constants(alpha[0], alpha[1], C.byref(t1), C.byref(isa))
        j = np.empty(n, np.int32); z = np.empty(n); e = np.empty(n)
        for i in range(n):
            j[i] = self.lib.oracle_pcg64_bounded(C.byref(g), K, None)
            z[i] = self.lib.oracle_gw_sample(self.lib.oracle_u64_to_unit(self.lib.oracle_pcg64_next(C.byref(g))), t1, isa)
            e[i] = self.lib.oracle_neg_exponential(self.lib.oracle_u64_to_unit(self.lib.oracle_pcg64_next(C.byref(g))))
        return j, z, e

    # --- targets ---------------------------------------------------------
    def logp(self, kind, params, x):
        x = np.ascontiguousarray(x, dtype=np.float64)
        params = np.ascontiguousarray(params if params is not None else [0.0], dtype=np.float64)
        D = x.shape[-1]
        flat = np.ascontiguousarray(x.reshape(-1, D))
        out = np.empty(flat.shape[0])
        self.lib.oracle_logp_many(KIND[kind], D, _dp(params), _dp(flat), flat.shape[0], _dp(out))
        return out.reshape(x.shape[:-1])

    # --- hot path ----------------------------------------------------------
    def run(self, kind, params, pos, logp, nsteps, interval=1, seed=0, stream=0, alpha=(2, 1),
            want_chain=True, want_trace=False, rng=None, acc=None, tot=None, per_walker=False):
        """Sequential-order stretch-move run.  pos [W,D] and logp [W] are NOT modified;
        returns dict(pos, logp, acc, tot, chain[(nsteps),W,D], trace, rejections, rng).
        per_walker: False = one pcg64 stream in walker order; True = pcg64(seed, w) per walker; 32 = pcg32(seed, w) per walker."""
        pos = np.array(pos, dtype=np.float64, order="C")
        logp = np.array(logp, dtype=np.float64)
        W, D = pos.shape
        params = np.ascontiguousarray(params if params is not None else [0.0], dtype=np.float64)
        if per_walker == 32:
            g = rng
            if g is None:
                g = (c_u64 * (2 * W))()                     # pcg32_t {state, inc} per walker
                self.lib.oracle_pcg32_seed_many(g, seed & (2**64 - 1), W)
        elif per_walker:
            g = rng
            if g is None:
                g = (Pcg64 * W)()
                self.lib.oracle_pcg64_seed_many(g, seed & (2**64 - 1), W)
        else:
            g = rng if rng is not None else self.pcg64(seed, stream)
        acc = np.ones(W, np.int64) if acc is None else np.array(acc, np.int64)
        tot = np.ones(W, np.int64) if tot is None else np.array(tot, np.int64)
        chain = np.empty((nsteps, W, D)) if want_chain else None
        trace = np.zeros(nsteps * interval * W, dtype=TRACE_DTYPE) if want_trace else None
        fn = (self.lib.oracle_stretch_run_per_walker32 if per_walker == 32 else
              self.lib.oracle_stretch_run_per_walker if per_walker else self.lib.oracle_stretch_run)
        rej = fn(C.byref(g), KIND[kind], _dp(params), W, D, alpha[0], alpha[1],
                                          _dp(pos), _dp(logp), acc.ctypes.data, tot.ctypes.data, nsteps, interval,
                                          _dp(chain), trace.ctypes.data if want_trace else None)
        return dict(pos=pos, logp=logp, acc=acc, tot=tot, chain=chain, trace=trace, rejections=rej, rng=g)


    def run_de(self, kind, params, pos, logp, nsteps, interval=1, seed=0, per_walker=False, gamma=None, rand=1.0e-4,
               want_chain=True, want_trace=False):
        """DifferentialEvolution mover in the red/black driver; one stream (seed, 0) in walker order, or one stream per
        walker.  Returns dict(pos, logp, acc, tot, chain[nsteps,W,D], trace)."""
        pos = np.array(pos, dtype=np.float64, order="C")
        logp = np.array(logp, dtype=np.float64)
        W, D = pos.shape
        params = np.ascontiguousarray(params if params is not None else [0.0], dtype=np.float64)
        if per_walker:
            g = (Pcg64 * W)()
            self.lib.oracle_pcg64_seed_many(g, seed & (2**64 - 1), W)
        else:
            g = self.pcg64(seed, 0)
        gamma = 2.38 / np.sqrt(2 * D) if gamma is None else gamma        # DifferentialEvolution.h:58 (integer 2*paramCount)
        acc, tot = np.ones(W, np.int64), np.ones(W, np.int64)
        chain = np.empty((nsteps, W, D)) if want_chain else None
        trace = np.zeros(nsteps * interval * W, dtype=TRACE_DTYPE) if want_trace else None
        self.lib.oracle_de_run(C.byref(g), int(per_walker), KIND[kind], _dp(params), W, D, float(gamma), -float(rand), 2.0 * float(rand),
                               _dp(pos), _dp(logp), acc.ctypes.data, tot.ctypes.data, nsteps, interval, _dp(chain),
                               trace.ctypes.data if want_trace else None)
        return dict(pos=pos, logp=logp, acc=acc, tot=tot, chain=chain, trace=trace)

    def run_walk(self, kind, params, pos, logp, nsteps, num_points, interval=1, seed=0, per_walker=False, want_chain=True,
                 want_trace=False):
        """WalkMove mover (num_points partner walkers) in the red/black driver; one engine (seed, 0) in walker order, or one
        engine per walker (stream = walker index, each with its own normal-distribution state)."""
        pos = np.array(pos, dtype=np.float64, order="C")
        logp = np.array(logp, dtype=np.float64)
        W, D = pos.shape
        params = np.ascontiguousarray(params if params is not None else [0.0], dtype=np.float64)
        if per_walker:
            g = (Pcg64 * W)()
            self.lib.oracle_pcg64_seed_many(g, seed & (2**64 - 1), W)
        else:
            g = self.pcg64(seed, 0)
        nstate = np.zeros(2 * W)
        acc, tot = np.ones(W, np.int64), np.ones(W, np.int64)
        chain = np.empty((nsteps, W, D)) if want_chain else None
        trace = np.zeros(nsteps * interval * W, dtype=TRACE_DTYPE) if want_trace else None
        self.lib.oracle_walk_run(C.byref(g), _dp(nstate), int(per_walker), KIND[kind], _dp(params), W, D, int(num_points),
                                 _dp(pos), _dp(logp), acc.ctypes.data, tot.ctypes.data, nsteps, interval, _dp(chain),
                                 trace.ctypes.data if want_trace else None)
        return dict(pos=pos, logp=logp, acc=acc, tot=tot, chain=chain, trace=trace)

    def half_slots(self, kind, params, pos, logp, acc, step, half, slot_begin, slot_count, seed=0, alpha=(2, 1)):
        """In-place update of one colour's slots [slot_begin, +slot_count) with jump-ahead draws (one rank's work)."""
        W, D = pos.shape
        params = np.ascontiguousarray(params if params is not None else [0.0], dtype=np.float64)
        assert pos.flags.c_contiguous and logp.flags.c_contiguous and acc.dtype == np.int64
        self.lib.oracle_stretch_half_slots(seed & (2**64 - 1), KIND[kind], _dp(params), W, D, alpha[0], alpha[1], _dp(pos),
                                           _dp(logp), acc.ctypes.data, step, half, slot_begin, slot_count)


class Ref:
    """oracle/_ref/libmcmcpp_ref*.so -- the unmodified reference headers (oracle/ref_driver.cpp)."""

    @staticmethod
    def available(small=False):
        return os.path.exists(os.path.join(HERE, "_ref", "libmcmcpp_ref_small.so" if small else "libmcmcpp_ref.so"))

    def __init__(self, small=False):
        path = os.path.join(HERE, "_ref", "libmcmcpp_ref_small.so" if small else "libmcmcpp_ref.so")
        L = self.lib = C.CDLL(path)
        L.ref_logp.restype = C.c_double
        L.ref_logp.argtypes = [C.c_int, C.c_int, c_dp, c_dp]
        L.ref_stretch_run.restype = C.c_int
        L.ref_stretch_run.argtypes = [C.c_int, c_dp, C.c_int, C.c_int, C.c_int, c_dp, c_dp, C.c_int, C.c_int, c_dp,
                                      C.POINTER(C.c_ulonglong), C.POINTER(C.c_ulonglong)]
        L.ref_stretch_run_pcg32.restype = C.c_int
        L.ref_stretch_run_pcg32.argtypes = L.ref_stretch_run.argtypes
        L.ref_stretch_time.restype = C.c_double
        L.ref_stretch_time.argtypes = [C.c_int, c_dp, C.c_int, C.c_int, C.c_int, c_dp, c_dp, C.c_int, C.c_int,
                                       C.POINTER(C.c_ulonglong), C.POINTER(C.c_ulonglong)]
        L.ref_init_normal.argtypes = [C.c_ulonglong, c_dp, C.c_long]
        L.ref_draws.argtypes = [C.c_longlong, C.c_longlong, C.c_int, C.c_long, C.POINTER(C.c_int), c_dp, c_dp]
        L.ref_pcg64_raw.argtypes = [C.c_ulonglong, C.c_ulonglong, C.c_ulonglong, C.POINTER(C.c_ulonglong), C.c_long]
        L.ref_pcg32_raw.argtypes = [C.c_ulonglong, C.c_ulonglong, C.POINTER(C.c_uint), C.c_long]
        L.ref_de_run.restype = C.c_int
        L.ref_de_run.argtypes = [C.c_int, c_dp, C.c_int, C.c_int, C.c_int, c_dp, c_dp, C.c_int, c_dp,
                                 C.POINTER(C.c_ulonglong), C.POINTER(C.c_ulonglong)]
        L.ref_walk_run.restype = C.c_int
        L.ref_walk_run.argtypes = [C.c_int, c_dp, C.c_int, C.c_int, C.c_int, C.c_int, c_dp, c_dp, C.c_int, c_dp,
                                   C.POINTER(C.c_ulonglong), C.POINTER(C.c_ulonglong)]
        L.ref_autocorr.restype = C.c_int
        L.ref_autocorr.argtypes = [C.c_int, c_dp, C.c_int, C.c_int, C.c_int, c_dp, c_dp, C.c_int, c_dp, c_dp]
        L.ref_covariance.restype = C.c_int
        L.ref_covariance.argtypes = [C.c_int, c_dp, C.c_int, C.c_int, C.c_int, c_dp, c_dp, C.c_int, C.c_int, c_dp, c_dp, c_dp]
        self.block_size = L.ref_block_size()
        self.hardware_concurrency = L.ref_hardware_concurrency()

    def logp(self, kind, params, x):
        x = np.ascontiguousarray(x, dtype=np.float64)
        params = np.ascontiguousarray(params if params is not None else [0.0], dtype=np.float64)
        D = x.shape[-1]
        flat = x.reshape(-1, D).copy()
        out = np.array([self.lib.ref_logp(KIND[kind], D, _dp(params), _dp(flat[i])) for i in range(flat.shape[0])])
        return out.reshape(x.shape[:-1])

    def covariance(self, kind, params, pos, logp, nsteps, cov_interval=1, seed=0):
        """Sequential reference run, then the reference's CovarianceMatrix over its chain.
        Returns (chain[(nsteps+1),W,D], cov[D,D], corr[D,D])."""
        pos = np.array(pos, dtype=np.float64, order="C")
        logp = np.array(logp, dtype=np.float64)
        W, D = pos.shape
        params = np.ascontiguousarray(params if params is not None else [0.0], dtype=np.float64)
        chain, cov, corr = np.empty((nsteps + 1, W, D)), np.empty((D, D)), np.empty((D, D))
        self.lib.ref_covariance(KIND[kind], _dp(params), W, D, seed, _dp(pos), _dp(logp), nsteps, cov_interval, _dp(chain), _dp(cov), _dp(corr))
        return chain, cov, corr

    def autocorr(self, kind, params, pos, logp, nsteps, seed=0):
        """Sequential reference run, then the reference's AutoCorrCalc (all walkers) -> (chain, iat[D])."""
        pos = np.array(pos, dtype=np.float64, order="C")
        logp = np.array(logp, dtype=np.float64)
        W, D = pos.shape
        params = np.ascontiguousarray(params if params is not None else [0.0], dtype=np.float64)
        chain, iat = np.empty((nsteps + 1, W, D)), np.empty(D)
        self.lib.ref_autocorr(KIND[kind], _dp(params), W, D, seed, _dp(pos), _dp(logp), nsteps, _dp(chain), _dp(iat))
        return chain, iat

    def init_normal(self, seed, n):
        out = np.empty(n)
        self.lib.ref_init_normal(seed, _dp(out), n)
        return out

    def draws(self, seed, stream, K, n):
        j = np.empty(n, np.int32); z = np.empty(n); e = np.empty(n)
        self.lib.ref_draws(seed, stream, K, n, j.ctypes.data_as(C.POINTER(C.c_int)), _dp(z), _dp(e))
        return j, z, e

    def pcg64_raw(self, seed, stream, n, skip=0):
        out = np.empty(n, np.uint64)
        self.lib.ref_pcg64_raw(seed, stream, skip, out.ctypes.data_as(C.POINTER(C.c_ulonglong)), n)
        return out

    def pcg32_raw(self, seed, stream, n):
        out = np.empty(n, np.uint32)
        self.lib.ref_pcg32_raw(seed, stream, out.ctypes.data_as(C.POINTER(C.c_uint)), n)
        return out

    def run(self, kind, params, pos, logp, nsteps, interval=1, seed=0, want_chain=True):
        """Sequential reference EnsembleSampler.  Returns dict(chain[(nsteps+1),W,D] incl. the
        initial row, pos (= last row), acc, tot (ensemble totals), stored)."""
        pos = np.array(pos, dtype=np.float64, order="C")
        logp = np.array(logp, dtype=np.float64)
        W, D = pos.shape
        params = np.ascontiguousarray(params if params is not None else [0.0], dtype=np.float64)
        chain = np.empty((nsteps + 1, W, D))
        acc, tot = C.c_ulonglong(), C.c_ulonglong()
        stored = self.lib.ref_stretch_run(KIND[kind], _dp(params), W, D, seed, _dp(pos), _dp(logp), nsteps, interval,
                                          _dp(chain), C.byref(acc), C.byref(tot))
        return dict(chain=chain if want_chain else None, pos=chain[-1].copy(), acc=acc.value, tot=tot.value, stored=stored)

    def run_pcg32(self, kind, params, pos, logp, nsteps, interval=1, seed=0):
        """PER_WALKER32 restatement driver (vendored pcg32 + libstdc++ adapters + reference GwDistribution, one facade per
        walker).  Returns dict(chain[(nsteps+1),W,D], pos, logp, acc, tot)."""
        pos = np.array(pos, dtype=np.float64, order="C")
        logp = np.array(logp, dtype=np.float64)
        W, D = pos.shape
        params = np.ascontiguousarray(params if params is not None else [0.0], dtype=np.float64)
        chain = np.empty((nsteps + 1, W, D))
        acc, tot = C.c_ulonglong(), C.c_ulonglong()
        self.lib.ref_stretch_run_pcg32(KIND[kind], _dp(params), W, D, seed, _dp(pos), _dp(logp), nsteps, interval, _dp(chain),
                                       C.byref(acc), C.byref(tot))
        return dict(chain=chain, pos=pos, logp=logp, acc=acc.value, tot=tot.value)

    def run_de(self, kind, params, pos, logp, nsteps, seed=0):
        """Sequential reference EnsembleSampler with the DifferentialEvolution mover -> dict(chain[(nsteps+1),W,D], acc, tot)."""
        pos = np.array(pos, dtype=np.float64, order="C")
        logp = np.array(logp, dtype=np.float64)
        W, D = pos.shape
        params = np.ascontiguousarray(params if params is not None else [0.0], dtype=np.float64)
        chain = np.empty((nsteps + 1, W, D))
        acc, tot = C.c_ulonglong(), C.c_ulonglong()
        self.lib.ref_de_run(KIND[kind], _dp(params), W, D, seed, _dp(pos), _dp(logp), nsteps, _dp(chain), C.byref(acc), C.byref(tot))
        return dict(chain=chain, acc=acc.value, tot=tot.value)

    def run_walk(self, kind, params, pos, logp, nsteps, num_points, seed=0):
        """Sequential reference EnsembleSampler with the WalkMove mover -> dict(chain[(nsteps+1),W,D], acc, tot)."""
        pos = np.array(pos, dtype=np.float64, order="C")
        logp = np.array(logp, dtype=np.float64)
        W, D = pos.shape
        params = np.ascontiguousarray(params if params is not None else [0.0], dtype=np.float64)
        chain = np.empty((nsteps + 1, W, D))
        acc, tot = C.c_ulonglong(), C.c_ulonglong()
        self.lib.ref_walk_run(KIND[kind], _dp(params), W, D, seed, int(num_points), _dp(pos), _dp(logp), nsteps, _dp(chain),
                              C.byref(acc), C.byref(tot))
        return dict(chain=chain, acc=acc.value, tot=tot.value)

    def time(self, kind, params, pos, logp, nsteps, threads=0, seed=0):
        """Seconds spent inside runMCMC for nsteps ensemble steps (threads=0: sequential sampler)."""
        pos = np.array(pos, dtype=np.float64, order="C")
        logp = np.array(logp, dtype=np.float64)
        W, D = pos.shape
        params = np.ascontiguousarray(params if params is not None else [0.0], dtype=np.float64)
        acc, tot = C.c_ulonglong(), C.c_ulonglong()
        secs = self.lib.ref_stretch_time(KIND[kind], _dp(params), W, D, seed, _dp(pos), _dp(logp), nsteps, threads,
                                         C.byref(acc), C.byref(tot))
        return secs, acc.value, tot.value
