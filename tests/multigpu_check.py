"""Run under torchrun (one process per GPU): a walker-sharded ensemble in both shard modes must equal the
sequential oracle bit for bit on every rank's owned rows, for the thread, warp and TMA-bulk kernels; DATA and DIM sharding
must take the oracle's decisions.
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 tests/multigpu_check.py
MCMCPP_CHECK_ONE_GPU=1: every rank uses device 0 (processes sharing one GPU; torch.distributed over gloo).  Only the shard
modes that exchange through CUDA-IPC peer memory run then -- PEER and DIM; NCCL (ALLGATHER, DATA) refuses two ranks on one
device.  That is how the world > 1 path is exercised on a one-GPU box (tests/test_gpu_multiproc_one_gpu.py)."""
import os
import sys
import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import mcmcpp_b200 as m                      # noqa: E402
from mcmcpp_b200 import parallel             # noqa: E402
from oracle.pyoracle import Port             # noqa: E402


def data_sharded_linreg(port, rank, world, local):
    """DATA sharding: replicated ensemble, each rank holds its rows of the data matrix, partial sums of squares
    all-reduced per colour.  Every rank must take the oracle's decisions (log-probs within 1e-12)."""
    W, P, M, steps = 512, 50, 6000, 5
    rng = np.random.default_rng(3)
    X = rng.standard_normal((M, P))
    y = X @ rng.standard_normal(P) + rng.standard_normal(M)
    params = np.concatenate([[float(M), 100.0], X.reshape(-1), y])
    init = np.linalg.lstsq(X, y, rcond=None)[0] + 0.1 * rng.standard_normal((W, P))
    lp = port.logp("linreg", params, init)
    o = port.run("linreg", params, init, lp, steps, seed=4)
    b, e = parallel.shard_data_rows(M, rank, world)
    s = m.EnsembleSampler(4, W, P, m.StretchMove(P, 4, m.LinearRegression(X[b:e], y[b:e], 100.0)), device=local, rank=rank,
                          world=world, shardMode="data")
    parallel.connect(s)
    s.setInitialWalkerPos(init, lp)
    ok = s.runMCMC(steps)
    pos, aux = s.getWalkerState()
    acc, tot = parallel.reduce_counts(s)
    good = ok and np.array_equal(pos, o["pos"]) and np.allclose(aux, o["logp"], rtol=1e-12, atol=0)
    good = good and (acc, tot) == (int(o["acc"].sum()), int(o["tot"].sum())) and np.array_equal(s.getChain()[1:], o["chain"])
    print("rank %d linreg W=%d P=%d M=%d data-sharded: %s" % (rank, W, P, M, "ok" if good else "MISMATCH"), flush=True)
    s.close()
    dist.barrier()
    return 0 if good else 1


def dim_sharded_diag_gauss(port, rank, world, local):
    """DIM sharding: every rank holds its columns of all walkers; partial sums exchanged through peer-mapped tables.
    Decisions must be the oracle's; this rank's columns of positions and chain bit-exact; log-probs within 1e-12."""
    bad = 0
    for W, D, steps in ((512, 100, 6), (2048, 1000, 4), (256, 37, 5), (2048, 500, 3), (1024, 250, 3)):
        iv = 1.0 / (1.0 + (np.arange(D) % 10))
        rng = np.random.default_rng(23)
        init = rng.standard_normal((W, D))
        lp = port.logp("diag_gauss", iv, init)
        o = port.run("diag_gauss", iv, init, lp, steps, seed=6)
        b, e = parallel.dim_slice(D, rank, world)
        s = m.EnsembleSampler(6, W, D, m.StretchMove(D, 6, m.DiagGaussian(iv)), device=local, rank=rank, world=world, shardMode="dim")
        parallel.connect(s)
        s.setInitialWalkerPos(init, lp)
        ok = s.runMCMC(steps)
        pos, aux = s.getWalkerState()
        acc, tot = parallel.reduce_counts(s)
        chain = s.getChain()
        good = ok and np.array_equal(pos[:, b:e], o["pos"][:, b:e]) and np.allclose(aux, o["logp"], rtol=1e-12, atol=0)
        good = good and (acc, tot) == (int(o["acc"].sum()), int(o["tot"].sum()))
        good = good and np.array_equal(chain[1:, :, b:e], o["chain"][:, :, b:e]) and np.array_equal(chain[0, :, b:e], init[:, b:e])
        if world > 1:
            other = np.ones(D, dtype=bool); other[b:e] = False
            good = good and not np.any(chain[:, :, other])                  # other ranks' columns read 0
        # the rank-local entry points move only this rank's columns, contiguous [W][D_local]
        loc = np.empty((W, e - b)); laux = np.empty(W)
        s.getWalkerStateLocalInto(loc.ctypes.data, laux.ctypes.data)
        good = good and np.array_equal(loc, o["pos"][:, b:e]) and np.array_equal(laux, aux)
        s2 = m.EnsembleSampler(6, W, D, m.StretchMove(D, 6, m.DiagGaussian(iv)), device=local, rank=rank, world=world, shardMode="dim")
        parallel.connect(s2)
        mine = np.ascontiguousarray(init[:, b:e])                # keep alive across the call
        s2.setInitialWalkerPosLocal(mine.ctypes.data, lp.ctypes.data)
        s2.runMCMC(steps)
        s2.getWalkerStateLocalInto(loc.ctypes.data, laux.ctypes.data)
        good = good and np.array_equal(loc, o["pos"][:, b:e]) and np.array_equal(laux, aux)
        s2.close()
        print("rank %d diag_gauss W=%d D=%d dim-sharded cols [%d,%d): %s" % (rank, W, D, b, e, "ok" if good else "MISMATCH"), flush=True)
        bad += 0 if good else 1
        s.close()
        dist.barrier()
    return bad


def main():
    rank, world, local = parallel.env_rank_world()
    one_gpu = os.environ.get("MCMCPP_CHECK_ONE_GPU") == "1"
    if one_gpu:
        local = 0
    torch.cuda.set_device(local)
    if one_gpu:
        dist.init_process_group("gloo")
    else:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    port = Port()
    failures = 0
    # walkers per colour must divide evenly over the ranks (2, 3, 4 and 8 ranks)
    cases = [("rosenbrock_2d", 6144, 2, 12), ("diag_gauss", 768, 100, 6), ("diag_gauss", 3072, 1000, 4), ("diag_gauss", 384, 37, 5)]
    for kind, W, D, steps in cases:
        params = 1.0 / (1.0 + (np.arange(D) % 10)) if kind == "diag_gauss" else np.zeros(0)
        rng = np.random.default_rng(17)
        init = rng.standard_normal((W, D))
        lp = port.logp(kind, params if params.size else None, init)
        o = port.run(kind, params if params.size else None, init, lp, steps, seed=2, want_chain=True)
        calc = m.DiagGaussian(params) if kind == "diag_gauss" else m.RosenbrockTwoDim()
        owned = parallel.owned_walkers(W, rank, world)
        for mode in (("peer",) if one_gpu else ("allgather", "peer")):
            s = m.EnsembleSampler(2, W, D, m.StretchMove(D, 2, calc), device=local, rank=rank, world=world, shardMode=mode)
            parallel.connect(s)
            s.setInitialWalkerPos(init, lp)
            ok = s.runMCMC(steps)
            pos, aux = s.getWalkerState()
            acc, tot = parallel.reduce_counts(s)
            good = ok and np.array_equal(pos[owned], o["pos"][owned]) and np.array_equal(aux[owned], o["logp"][owned])
            if mode == "allgather":
                good = good and np.array_equal(pos, o["pos"])           # every rank holds the whole ensemble
            good = good and (acc, tot) == (int(o["acc"].sum()), int(o["tot"].sum()))
            chain = s.getChain()                                         # owned rows of every stored step
            good = good and np.array_equal(chain[1:, owned], o["chain"][:, owned])
            print("rank %d %s W=%d D=%d %s: %s" % (rank, kind, W, D, mode, "ok" if good else "MISMATCH"), flush=True)
            failures += 0 if good else 1
            s.close()
            dist.barrier()
    if not one_gpu:
        failures += data_sharded_linreg(port, rank, world, local)
    failures += dim_sharded_diag_gauss(port, rank, world, local)
    t = torch.tensor([failures], device="cpu" if one_gpu else "cuda")
    dist.all_reduce(t)
    dist.destroy_process_group()
    if int(t.item()):
        raise SystemExit(1)
    if rank == 0:
        print("MULTIGPU_OK", flush=True)


if __name__ == "__main__":
    main()
