#!/usr/bin/env python
"""bench.py -- walker-updates/s of the ensemble-update hot path on N B200s (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c4|c2|c3|c5|c1] [--shard allgather|peer]
  python bench.py --impl reference ...       the reference's own CPU implementation on the host cores

A "step" is one ensemble step (EnsembleSampler::performStep, MCMCpp/EnsembleSampler.h:341-360): every one of
the W walkers is updated once (red half then black half) = W walker-updates.  One JSON line on stdout (rank 0).
Default workload = BASELINE.json configs[3] ("1000-D Gaussian, 1M walkers"), the configuration north_star
quotes its target on and the one that is sharded over 1/2/4/8 GPUs; it fits one GPU (8 GB of state).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "walker-updates/sec"
UNIT = "walker-updates/s"

WORKLOADS = {
    # name: (description, kind, W, D, cpu sample walkers, cpu sample steps)
    "c4": ("BASELINE configs[3]: 1000-D diagonal Gaussian, 2^20 walkers, stretch move a=2", "diag_gauss", 1 << 20, 1000, 4096, 40),
    "c2": ("BASELINE configs[1]: Rosenbrock 2-D banana, 2^20 walkers, stretch move a=2", "rosenbrock_2d", 1 << 20, 2, 65536, 300),
    "c3": ("BASELINE configs[2]: 100-D correlated Gaussian (dense precision, rho=0.9), 65,536 walkers, FP64 tensor-core quadratic form",
           "dense_gauss", 1 << 16, 100, 4096, 60),
    "c5": ("BASELINE configs[4]: Bayesian linear regression, 50 params over 1e6 synthetic data points, 16,384 walkers, "
           "likelihood of each colour as one FP64 tensor-core GEMM; data rows sharded over the GPUs", "linreg", 1 << 14, 50, 128, 1),
    "c1": ("BASELINE configs[0]: 2-D skewed Gaussian, 32 walkers", "skewed_gauss_2d", 32, 2, 32, 20000),
}


C5_M = int(os.environ.get("MCMCPP_BENCH_C5_ROWS", "1000000"))     # data points of configs[4]
_LINREG = {}


def linreg_data(D):
    """X ~ N(0,1), theta* ~ N(0,1), y = X theta* + N(0,1), prior N(0, 10^2 I); seed 1234."""
    if D not in _LINREG:
        rng = np.random.default_rng(1234)
        X = rng.standard_normal((C5_M, D))
        y = X @ rng.standard_normal(D) + rng.standard_normal(C5_M)
        _LINREG[D] = (X, y, np.linalg.lstsq(X, y, rcond=None)[0])
    return _LINREG[D]


def target_params(kind, D):
    if kind == "linreg":
        X, y, _ = linreg_data(D)
        return np.concatenate([[float(X.shape[0]), 100.0], X.reshape(-1), y])
    if kind == "diag_gauss":
        return 1.0 / (1.0 + (np.arange(D) % 10))
    if kind == "skewed_gauss_2d":
        return np.array([0.13])
    if kind == "dense_gauss":
        idx = np.arange(D)
        return np.linalg.inv(0.9 ** np.abs(idx[:, None] - idx[None, :])).reshape(-1)
    return np.zeros(0)


def algorithmic_bytes_per_update(D, stored):
    """own row read + partner row read + own row write (+ chain row) + logp read/write.
    `stored` = fraction of the steps whose row goes to the chain (0, 1, or 1/N with subsampling)."""
    return 8 * D * (3 + float(stored)) + 16


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """SM clock + throttle reasons DURING the timed region (B200_PROFILING.md recipe).  NVML in-process (nvidia_ml_py), one
    sample every ~4 ms, so that a 57 ms timed region (20 steps of configs[3]) still holds >= 5 samples; falls back to
    polling nvidia-smi (0.1 s period) when NVML cannot be loaded."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    BITS = {"sw_power_cap": 0x4, "hw_slowdown": 0x8, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40}

    def __init__(self, index, bus_id=None):
        self.index, self.rows, self.stop, self.thread = index, [], threading.Event(), None
        self.nvml, self.handle, self.max_mhz = None, None, None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.handle = pynvml.nvmlDeviceGetHandleByPciBusId(bus_id.encode() if bus_id else b"") if bus_id else \
                pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM))
            self.nvml = pynvml
        except Exception:
            self.nvml = None

    def _reasons(self):
        n = self.nvml
        fn = getattr(n, "nvmlDeviceGetCurrentClocksEventReasons", None) or n.nvmlDeviceGetCurrentClocksThrottleReasons
        return int(fn(self.handle))

    def _loop(self):
        while not self.stop.is_set():
            try:
                if self.nvml is not None:
                    mhz = int(self.nvml.nvmlDeviceGetClockInfo(self.handle, self.nvml.NVML_CLOCK_SM))
                    bits = self._reasons()
                    self.rows.append([str(mhz), str(self.max_mhz)] + ["Active" if bits & self.BITS[k] else "Not Active" for k in
                                                                       ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")])
                    self.stop.wait(0.004)
                    continue
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([c.strip() for c in out.split(",")])
            except Exception:
                pass
            self.stop.wait(0.1)

    def __enter__(self):
        self.thread = threading.Thread(target=self._loop, daemon=True)
        self.thread.start()
        return self

    def __exit__(self, *a):
        self.stop.set()
        self.thread.join(timeout=6)

    def summary(self):
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows for n, v in zip(names, r[2:6]) if v.lower().startswith("active")})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(self.rows), "source": "nvml" if self.nvml is not None else "nvidia-smi"}


# ---- CPU arm: the reference's own implementation on the host cores (oracle/ is the checker; this is the one
# ---- place bench.py executes it, as the thing that is the BASELINE, never as the product) -------------------
def cpu_reference_run(workload, steps, warmup, seed=0, walkers=None):
    """Times EnsembleSampler (1 core) and ParallelEnsembleSampler (all host threads) of the unmodified reference
    (oracle/_ref, compiled from /root/reference by oracle/Makefile) on a bounded sample of the workload; falls
    back to the plain-C port (kind "port", 1 core) when oracle/_ref is absent."""
    from oracle.pyoracle import Port, Ref
    desc, kind, W, D, Wc, _ = WORKLOADS[workload]
    if walkers:
        Wc = int(walkers)
    params = target_params(kind, D)
    rng = np.random.default_rng(1)
    init = rng.standard_normal((Wc, D))
    if kind == "linreg":
        init = linreg_data(D)[2] + 0.1 * init
    port = Port()
    lp = port.logp(kind, params, init)
    cores = os.cpu_count() or 1
    updates = float(Wc) * steps
    if Ref.available(small=True):
        R = Ref(small=True)
        if warmup:
            R.time(kind, params, init, lp, warmup, threads=0, seed=seed)
        t_seq, _, _ = R.time(kind, params, init, lp, steps, threads=0, seed=seed)
        t_par, _, _ = R.time(kind, params, init, lp, steps, threads=cores, seed=seed)
        par_threads = cores
        if cores > 16:      # spin barriers can lose to fewer threads on many-core hosts: try 16 too, keep the better
            t16, _, _ = R.time(kind, params, init, lp, steps, threads=16, seed=seed)
            if t16 < t_par:
                t_par, par_threads = t16, 16
        seq, par = updates / t_seq, updates / t_par
        best, used, secs = (par, par_threads, t_par) if par >= seq else (seq, 1, t_seq)
        cores = par_threads
        return {"value": best, "unit": UNIT, "cores": used, "kind": "reference", "seconds": secs,
                "sample": "%d walkers x %d steps, D=%d (%s); EnsembleSampler 1 core %.4g/s, ParallelEnsembleSampler "
                          "%d threads %.4g/s; unmodified reference headers, chain BlockSize=%d, -O3 -march=x86-64-v3 "
                          "-ffp-contract=off" % (Wc, steps, D, kind, seq, cores, par, R.block_size)}
    t0 = time.perf_counter()
    port.run(kind, params, init, lp, steps, seed=seed, want_chain=False)
    secs = time.perf_counter() - t0
    return {"value": updates / secs, "unit": UNIT, "cores": 1, "kind": "port", "seconds": secs,
            "sample": "%d walkers x %d steps, D=%d (%s); plain-C port oracle/oracle.c, 1 core" % (Wc, steps, D, kind)}


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    desc, kind, W, D, Wc, _ = WORKLOADS[args.workload]
    r = cpu_reference_run(args.workload, args.steps, args.warmup)
    line = {"impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * r["seconds"] / args.steps, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": desc, "walkers": W, "params": D, "cpu_sample_walkers": Wc,
                       "note": "CPU throughput per update does not depend on W beyond cache effects (SURVEY.md s.8(d))"},
            "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)
    return 0


# ---- GPU arm ------------------------------------------------------------------------------------------------
def device_inputs(torch, kind, W, D, params, device):
    """Synthetic initial ensemble created in HBM: x ~ N(0, I), aux = log posterior (the caller's
    job in the reference API, EnsembleSampler.h:220).  Same seed on every rank -> identical ensembles."""
    g = torch.Generator(device=device)
    g.manual_seed(1234)
    pos = torch.randn((W, D), dtype=torch.float64, device=device, generator=g)
    aux = torch.empty(W, dtype=torch.float64, device=device)
    chunk = max(1, (1 << 27) // D)
    if kind == "diag_gauss":
        iv = torch.tensor(params, dtype=torch.float64, device=device)
        for b in range(0, W, chunk):
            x = pos[b:b + chunk]
            aux[b:b + chunk] = -0.5 * (x * x * iv).sum(dim=1)
    elif kind == "dense_gauss":
        P = torch.tensor(params, dtype=torch.float64, device=device).reshape(D, D)
        aux = -0.5 * ((pos @ P) * pos).sum(dim=1)
    elif kind == "rosenbrock_2d":
        x, y = pos[:, 0], pos[:, 1]
        aux = -(100.0 * (y - x * x) ** 2 + (1.0 - x) ** 2) / 20.0
    elif kind == "skewed_gauss_2d":
        t1, t2 = pos[:, 0] / 2 - pos[:, 1], pos[:, 0] / 2 + pos[:, 1]
        aux = -((t1 * t1) / float(params[0]) + t2 * t2) / 2.0
    elif kind == "linreg":
        X, y, ols = linreg_data(D)
        pos = torch.tensor(ols, dtype=torch.float64, device=device) + 0.1 * pos        # theta ~ OLS + 0.1 N(0, I)
        ss = torch.zeros(W, dtype=torch.float64, device=device)
        for b in range(0, X.shape[0], 1 << 17):
            Xb = torch.tensor(X[b:b + (1 << 17)], device=device)
            r = torch.tensor(y[b:b + (1 << 17)], device=device)[:, None] - Xb @ pos.T
            ss += (r * r).sum(dim=0)
        aux = -0.5 * ss - 0.5 * (pos * pos).sum(dim=1) / 100.0
    else:
        raise ValueError(kind)
    return pos.contiguous(), aux.contiguous()


def make_calc(m, kind, params, rank=0, world=1):
    if kind == "linreg":
        from mcmcpp_b200 import parallel
        D = int(round((params.size - 2) / params[0])) - 1
        X, y, _ = linreg_data(D)
        b, e = parallel.shard_data_rows(X.shape[0], rank, world)       # DATA sharding: this rank's rows
        return m.LinearRegression(X[b:e], y[b:e], 100.0)
    return {"diag_gauss": lambda: m.DiagGaussian(params), "dense_gauss": lambda: m.DenseGaussian(params.reshape(int(round(params.size ** 0.5)), -1)),
            "rosenbrock_2d": lambda: m.RosenbrockTwoDim(),
            "skewed_gauss_2d": lambda: m.SkewedGaussianTwoDim(params[0])}[kind]()


def ncu_constants():
    """Per-launch counters of the dominant kernels taken from ncu captures (profiles/traffic.json names the capture each
    comes from): DRAM bytes and executed warp instructions.  Not measured in this run -- ncu cannot run inside a timed region."""
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        with open(tpath) as f:
            return json.load(f)
    return {}


def bus_id_of(torch, local):
    try:
        p = torch.cuda.get_device_properties(local)
        return "%08x:%02x:%02x.0" % (p.pci_domain_id, p.pci_bus_id, p.pci_device_id)
    except Exception:
        return None


def measure(torch, dist, m, parallel, args, workload, steps, warm, with_e2e, with_cpu, cpu_scale=1.0, cpu_walkers=None, cpu_warm=1):
    """One workload on this process group: device-timed value, roofline, optional e2e and cpu_baseline legs."""
    rank, world, local = parallel.env_rank_world()
    device = torch.device("cuda", local)
    desc, kind, W, D, Wc, cpu_steps = WORKLOADS[workload]
    if os.environ.get("MCMCPP_BENCH_D"):          # analysis only (e.g. one rank's slice of C4 on one GPU); named in config
        D = int(os.environ["MCMCPP_BENCH_D"])
        desc += " [params overridden to %d]" % D
    if os.environ.get("MCMCPP_BENCH_W"):          # analysis only (size sweeps); named in config
        W = int(os.environ["MCMCPP_BENCH_W"])
        desc += " [walkers overridden to %d]" % W
    params = target_params(kind, D)
    # default sharding per workload: C5 shards its data rows, C4 (separable) its parameter dimension, the rest walkers
    shard = args.shard or ("data" if kind == "linreg" else "dim" if kind == "diag_gauss" else "peer")
    if kind == "linreg":
        shard = "data"
    if world == 1 and args.shard != "dim":      # explicit dim with one rank = the sharded kernels on one GPU (overhead measurement)
        shard = "none"
    rng_mode = "per_walker" if args.mover == "de" else args.rng
    if args.mover == "de":
        desc += " [mover: DifferentialEvolution instead of the stretch move]"
        if world > 1:
            shard = "peer"

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    every = args.store_every
    if every > 0 and (world > 1 or steps % every):
        raise SystemExit("--store-every N needs one GPU and --steps a multiple of N")
    # runMCMC counts STORED steps: with subsampling each is N - 1 unsaved ensemble steps + 1 saved (EnsembleSampler.h:297-308)
    calls = (lambda k: k // every) if every > 0 else (lambda k: k)
    if every > 0:
        warm = every * max(1, warm // every)
    chain_bytes = (calls(warm) + calls(steps) + 2) * W * D * 8 if every > 0 else 0

    def new_sampler():
        mover_cls = m.DifferentialEvolution if args.mover == "de" else m.StretchMove
        s = m.EnsembleSampler(0, W, D, mover_cls(D, 0, make_calc(m, kind, params, rank, world)), maxChainSizeBytes=chain_bytes,
                              rngMode=rng_mode, device=local, rank=rank, world=world, shardMode=shard)
        parallel.connect(s)
        if every > 0:
            s.setSlicingMode(True, every)
        return s

    pos, aux = device_inputs(torch, kind, W, D, params, device)

    # ---- value: inputs resident in HBM, K steps timed with CUDA events on the sampler's stream ------------
    s = new_sampler()
    s.setInitialWalkerPosDevice(pos.data_ptr(), aux.data_ptr())
    s.runMCMC(calls(warm))
    launches0 = s.getLaunchCount()
    acc0, _ = s._counts()
    barrier()
    with ClockSampler(local, bus_id_of(torch, local)) as clocks:
        wall0 = time.perf_counter()
        assert s.runMCMC(calls(steps)), "chain byte cap reached inside the timed region"
        barrier()
        wall = time.perf_counter() - wall0
    ms = torch.tensor([s.lastDeviceMs], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = float(ms.item())
    launches = s.getLaunchCount() - launches0
    acc1, _ = s._counts()
    accepted, total = parallel.reduce_counts(s)
    redraws = s.getRngRedraws()
    s.close()
    value = float(W) * steps / (ms * 1e-3)
    # acceptance of the proposals made inside the timed region (this rank's count; DIM/DATA ranks count every walker)
    owned_updates = float(W) * steps if shard in ("none", "data", "dim") else float(W) * steps / world
    acc_frac = (acc1 - acc0) / owned_updates

    # ---- e2e: the call a user makes, HOST buffers in, HOST buffers out, copies inside the timed region --------
    # DIM sharding: every process keeps its own column shard of the rows on its host side (W x D_local, contiguous)
    e2e = None
    lo, hi = parallel.dim_slice(D, rank, world) if shard == "dim" else (0, D)
    Dloc = hi - lo
    if with_e2e:
        host_pos = torch.empty((W, Dloc), dtype=torch.float64, pin_memory=True)
        host_aux = torch.empty(W, dtype=torch.float64, pin_memory=True)
        host_pos.copy_(pos[:, lo:hi])
        host_aux.copy_(aux)
        del pos, aux
        torch.cuda.empty_cache()
        s = new_sampler()
        # untimed warm-up of THIS sampler (first launches load the kernels; ALLGATHER/DATA: the first collective sets up the
        # NCCL channels, ~1 s at 8 ranks); the timed call below starts again from the host buffers
        s.setInitialWalkerPosLocal(host_pos.data_ptr(), host_aux.data_ptr())
        s.runMCMC(calls(every) if every > 0 else 2)
        if every > 0:
            s.reset()
        barrier()
        t0 = time.perf_counter()
        s.setInitialWalkerPosLocal(host_pos.data_ptr(), host_aux.data_ptr())
        ok = s.runMCMC(calls(steps))
        s.getWalkerStateLocalInto(host_pos.data_ptr(), host_aux.data_ptr())
        e2e_acc, e2e_tot = s._counts()
        barrier()
        e2e_s = time.perf_counter() - t0
        t = torch.tensor([e2e_s], dtype=torch.float64, device=device)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
        s.close()
        del host_pos, host_aux
        state_bytes = W * Dloc * 8 + W * 8
        e2e = {"value": float(W) * steps / e2e_s, "unit": UNIT, "h2d_bytes_per_step": state_bytes / steps,
               "d2h_bytes_per_step": (state_bytes + 16) / steps, "seconds": e2e_s, "ok": bool(ok),
               "what": "setInitialWalkerPos(pinned host%s) + runMCMC(%d) + walker state and counters back to pinned host, per rank"
                       % (", this rank's %d columns" % Dloc if shard == "dim" else "", steps)}
    else:
        del pos, aux
        torch.cuda.empty_cache()

    if rank != 0:
        return None

    peak, peak_src = measured_peaks()
    consts = ncu_constants()
    per_launch_updates = (W // 2) if shard in ("data", "dim") else (W // 2) // world
    launch_ms = ms / (2.0 * steps)         # one colour = one half-step
    stored_frac = 1.0 / every if every > 0 else 0.0
    alg_per_update = algorithmic_bytes_per_update(Dloc, stored_frac)            # on the slice this rank streams
    alg_bytes = per_launch_updates * alg_per_update
    achieved = alg_bytes / (launch_ms * 1e-3) / 1e9
    # bytes that MUST move: both rows read, the own row written only when the proposal is accepted (+ chain row, log-probs)
    actual_per_update = 8.0 * Dloc * (2.0 + acc_frac + stored_frac) + 16.0
    achieved_actual = per_launch_updates * actual_per_update / (launch_ms * 1e-3) / 1e9
    kernel = ("dimShardDraw + dimShardSum (TMA, partial sums over NVLink) + dimShardFinish" if shard == "dim" else "stretchDrawKernel + stretchHalfStepBulk (TMA)"
              if (D >= 130 and D % 2 == 0) else "stretchHalfStepWarp" if D > 8 else
              "stretchFusedSmall (whole runMCMC in one launch)" if (W // 2 <= 1024 and world == 1) else "stretchHalfStepThread")
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "frac_actual": achieved_actual / peak, "achieved_actual": achieved_actual,
                "traffic": consts.get(workload) if (world == 1 and shard == "none") else None,
                "kernel": kernel, "peak_source": peak_src,
                "algorithmic_bytes_per_update": alg_per_update, "actual_bytes_per_update": actual_per_update,
                "acceptance_in_timed_region": acc_frac, "updates_per_launch": per_launch_updates, "launch_ms": launch_ms,
                "note": "frac = SURVEY s.8(d) algorithmic bytes (own-row write charged on EVERY update) / launch time / peak; it can exceed 1 "
                        "because a rejected proposal writes nothing.  frac_actual charges the own-row write only for the measured "
                        "acceptance: bytes that must move = 8 D (2 + acc) + 16.  launch_ms = CUDA-event time of the timed region / "
                        "colours" + ("" if world == 1 else " (includes the per-colour exchange)")
                        + ("; traffic = ncu dram bytes of one launch (profiles/traffic.json), N=1 only" if world == 1 else "")}
    if D <= 8:
        # state is L2-resident: the kernel is bound by instruction issue, not HBM -- report the issue-slot roofline beside it
        inst = consts.get(workload + ("_per_walker32" if rng_mode == "per_walker32" else "") + "_inst_executed_per_launch")
        sm_hz = (clocks.summary().get("sm_mhz") or 1965) * 1e6
        sms = torch.cuda.get_device_properties(local).multi_processor_count
        if inst and W // 2 > 1024:
            roofline["issue"] = {"warp_instructions_per_launch": inst, "instructions_per_update": inst * 32.0 / per_launch_updates,
                                 "achieved_ginst_s": inst / (launch_ms * 1e-3) / 1e9, "peak_ginst_s": sms * 4 * sm_hz / 1e9,
                                 "frac": inst / (launch_ms * 1e-3) / (sms * 4 * sm_hz),
                                 "note": "sm__inst_executed.sum of one launch (ncu, profiles/) / live launch time over SMs x 4 schedulers x 1 "
                                         "warp-instruction/clk x SM clock"}
        if W // 2 <= 1024:
            roofline["note"] += ("; this ensemble is ONE CTA (16..1024 threads): latency-bound on the ~450-instruction dependent chain of an update, "
                                 "not a bandwidth workload -- the HBM fraction is reported only because the contract asks for it")
    if kind == "dense_gauss":
        # compute-bound: FP64 tensor cores (DMMA); flops/update = 2D^2 + 5D; peak measured on this GPU
        import ctypes
        from mcmcpp_b200 import _capi
        tf = ctypes.c_double()
        _capi.check(_capi.lib().mcmcpp_fp64_tensor_peak(local, ctypes.byref(tf)))
        flops = per_launch_updates * (2.0 * D * D + 5.0 * D)
        ach = flops / (launch_ms * 1e-3) / 1e12
        roofline = {"bound": "tensor", "achieved": ach, "peak": tf.value, "unit": "TFLOP/s", "frac": ach / tf.value,
                    "traffic": consts.get(workload) if world == 1 else None,
                    "kernel": "denseGaussHalfStep (mma.sync.m8n8k4.f64)", "peak_source": "measured here: mcmcpp_fp64_tensor_peak (register-resident DMMA chains); MEASURED_PEAKS.json has no FP64 entry",
                    "flops_per_update": 2 * D * D + 5 * D, "updates_per_launch": per_launch_updates, "launch_ms": launch_ms,
                    "hbm_view": {"achieved_gbs": achieved, "frac_of_hbm_peak": achieved / peak}}
    if kind == "linreg":
        import ctypes
        from mcmcpp_b200 import _capi
        tf = ctypes.c_double()
        _capi.check(_capi.lib().mcmcpp_fp64_tensor_peak(local, ctypes.byref(tf)))
        rows = parallel.shard_data_rows(C5_M, rank, world)
        rows = rows[1] - rows[0]
        fl = 2.0 * rows * D + 3.0 * rows                                  # 2 M P + 3 M per update
        ach = per_launch_updates * fl / (launch_ms * 1e-3) / 1e12
        roofline = {"bound": "tensor", "achieved": ach, "peak": tf.value, "unit": "TFLOP/s", "frac": ach / tf.value,
                    "traffic": consts.get(workload) if world == 1 else None,
                    "kernel": "linregSse (mma.sync.m8n8k4.f64 GEMM [M x 52] x [52 x K], square-and-sum epilogue)",
                    "peak_source": "measured here: mcmcpp_fp64_tensor_peak (register-resident DMMA chains); MEASURED_PEAKS.json has no FP64 entry",
                    "flops_per_update": fl, "updates_per_launch": per_launch_updates, "launch_ms": launch_ms, "data_rows_this_rank": rows,
                    "note": "launch_ms = half-step time (propose + GEMM + reduce + accept" + ("" if world == 1 else " + all-reduce of K partial sums") + "); the GEMM is > 99 % of it"}
    cpu = None
    if with_cpu and world == 1:
        cpu = cpu_reference_run(workload, max(1, int(round(cpu_steps * cpu_scale))), cpu_warm, walkers=cpu_walkers)
        cpu = {k: cpu[k] for k in ("value", "unit", "cores", "kind", "sample")}
    return {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": steps, "warmup": warm,
            "ms_per_step": ms / steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": {"workload": desc, "walkers": W, "params": D, "rng": rng_mode, "shard": shard,
                       "chain": "off (the CPU reference arm always stores its chain: EnsembleSampler has no chain-off mode)" if every <= 0 else
                                "every %d-th step stored: device rows -> pinned host chain on a side stream (%d rows of %.1f MB inside the timed region)"
                                % (every, calls(steps), W * D * 8 / 1e6),
                       "l2": "inputs (%.1f GB of walker rows) far exceed the 126 MB L2; no flush needed" % (W * D * 8 / 1e9)
                       if W * D * 8 > (1 << 29) else "state fits L2: L2-resident steady state is the workload itself"},
            "clocks": clocks.summary(), "e2e": e2e, "gpu_launches": int(launches) * world, "roofline": roofline,
            "cpu_baseline": cpu, "wall_s_timed_region": wall,
            "acceptance": {"accepted": accepted, "total": total, "rng_redraws": redraws}}


# short runs of the other BASELINE configs appended to the default line: (steps, warm-up, fraction of the cpu sample)
OTHERS = {"c1": (20000, 2000, 0.5, None, 1), "c2": (2000, 50, 0.3, None, 1), "c3": (400, 20, 0.3, None, 1), "c5": (2, 1, 1.0, None, 0)}


def gpu_arm(args):
    import torch
    import torch.distributed as dist
    import mcmcpp_b200 as m
    from mcmcpp_b200 import parallel
    rank, world, local = parallel.env_rank_world()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    bound = None
    if world > 1 and os.environ.get("MCMCPP_BENCH_BIND", "1") != "0":
        # one process per GPU: onto the GPU's own NUMA node before anything is pinned (rank 0's cpu legs run at N=1 only)
        bound = parallel.bind_host_to_device(bus_id_of(torch, local))
    if world > 1:
        dist.init_process_group("nccl", device_id=device)
    if world != args.gpus and rank == 0:
        print("bench.py: --gpus %d but WORLD_SIZE %d; using WORLD_SIZE" % (args.gpus, world), file=sys.stderr)
    warm = max(3, args.warmup)
    if args.shard_list:
        # analysis: several shard modes in ONE launch (one start-up on an N-GPU box), device-timed value only, one line per mode
        for mode in args.shard_list.split(","):
            args.shard = mode
            line = measure(torch, dist, m, parallel, args, args.workload, args.steps, warm, False, False)
            if rank == 0:
                print(json.dumps(line), flush=True)
        if world > 1:
            dist.destroy_process_group()
        return 0
    line = measure(torch, dist, m, parallel, args, args.workload, args.steps, warm, True, not args.no_cpu)
    if rank == 0 and world > 1:
        line["config"]["host_binding"] = ("rank processes bound to their GPU's NUMA-local cores (rank 0: %d cores)" % len(bound)) if bound else "none"
    if rank == 0 and world == 1 and args.workload == "c4" and not args.no_others and not os.environ.get("MCMCPP_BENCH_D") and not os.environ.get("MCMCPP_BENCH_W") \
            and args.shard is None and args.mover == "stretch" and args.store_every == 0:
        # the other BASELINE configs, short runs, one GPU: driver-visible numbers for C1/C2/C3/C5
        others = {}
        for name, (k_steps, k_warm, cpu_scale, cpu_walkers, cpu_warm) in OTHERS.items():
            try:
                r = measure(torch, dist, m, parallel, args, name, k_steps, k_warm, False, not args.no_cpu, cpu_scale, cpu_walkers, cpu_warm)
                others[name] = {k: r[k] for k in ("value", "unit", "steps", "warmup", "ms_per_step", "config", "clocks", "gpu_launches",
                                                 "roofline", "cpu_baseline", "acceptance")}
            except Exception as exc:                    # a failure here must not take the headline line with it
                others[name] = {"error": "%s: %s" % (type(exc).__name__, exc)}
        # C2 once more on the fast engine (one pcg32 per walker, its own pinned oracle): the device-timed value and its rooflines
        try:
            saved = args.rng
            args.rng = "per_walker32"
            k_steps, k_warm = OTHERS["c2"][0], OTHERS["c2"][1]
            r = measure(torch, dist, m, parallel, args, "c2", k_steps, k_warm, False, False)
            others["c2_per_walker32"] = {k: r[k] for k in ("value", "unit", "steps", "warmup", "ms_per_step", "config", "clocks", "gpu_launches",
                                                            "roofline", "acceptance")}
        except Exception as exc:
            others["c2_per_walker32"] = {"error": "%s: %s" % (type(exc).__name__, exc)}
        finally:
            args.rng = saved
        line["other_workloads"] = others
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c4", choices=sorted(WORKLOADS))
    ap.add_argument("--shard", default=None, choices=["allgather", "peer", "dim"])
    ap.add_argument("--shard-list", default=None, help="analysis: comma-separated shard modes measured one after the other in this launch (no e2e / cpu legs), one JSON line each")
    ap.add_argument("--rng", default="sequential", choices=["sequential", "per_walker", "per_walker32"])
    ap.add_argument("--mover", default="stretch", choices=["stretch", "de"], help="de: Mover::DifferentialEvolution (analysis runs; not the BASELINE metric)")
    ap.add_argument("--store-every", type=int, default=0, help="N > 0: chain on, every N-th step stored (SURVEY s.8(d): C2's second run stores every 100th); single GPU")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-others", action="store_true", help="default c4 line only: skip the short runs of c1/c2/c3/c5 (other_workloads)")
    args = ap.parse_args()
    if args.steps <= 0:
        raise SystemExit("--steps must be positive")
    return reference_arm(args) if args.impl == "reference" else gpu_arm(args)


if __name__ == "__main__":
    sys.exit(main())
