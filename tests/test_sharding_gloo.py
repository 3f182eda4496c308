"""N>1 host-side logic on CPU (gloo, world_size 2): the walker-index shard arithmetic of mcmcpp_b200.parallel and the
jump-ahead draw indexing make a sharded ensemble reproduce the sequential reference bit for bit.  Each rank plays one
GPU: it updates its slots of a colour (oracle restatement of one rank's kernel launch), then the colour is all-gathered
in place exactly as the runtime's ncclAllGather does (runtime.cu halfStep)."""
import os
import sys
import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port_no, W, D, steps, out_dir):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port_no)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle.pyoracle import Port
    from mcmcpp_b200 import parallel
    port = Port()
    K = W // 2
    iv = 1.0 / (1.0 + (np.arange(D) % 10))
    rng = np.random.default_rng(5)
    pos = rng.standard_normal((W, D))
    logp = port.logp("diag_gauss", iv, pos)
    acc = np.zeros(W, np.int64)
    b, e = parallel.shard_slots(W, rank, world)
    n = e - b
    for t in range(steps):
        for h in (0, 1):
            port.half_slots("diag_gauss", iv, pos, logp, acc, t, h, b, n, seed=3)
            mine = torch.from_numpy(pos[h * K + b:h * K + e].copy())
            parts = [torch.empty_like(mine) for _ in range(world)]
            dist.all_gather(parts, mine)                       # the in-place all-gather of the updated colour
            pos[h * K:(h + 1) * K] = torch.cat(parts).numpy()
    # ensemble-wide counters, as parallel.reduce_counts does
    owned = parallel.owned_walkers(W, rank, world)
    tot = torch.tensor([int(acc[owned].sum())])
    dist.all_reduce(tot)
    lp_parts = [torch.empty(2 * n, dtype=torch.float64) for _ in range(world)]
    dist.all_gather(lp_parts, torch.from_numpy(logp[owned].copy()))
    if rank == 0:
        np.savez(os.path.join(out_dir, "sharded.npz"), pos=pos, acc=int(tot), logp_owned=torch.stack(lp_parts).numpy())
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("W,D,steps", [(64, 6, 5)])
def test_two_rank_sharded_ensemble_equals_sequential_reference(tmp_path, port, W, D, steps):
    world = 2
    port_no = 29500 + (os.getpid() % 400)
    mp.spawn(_worker, args=(world, port_no, W, D, steps, str(tmp_path)), nprocs=world, join=True)
    z = np.load(tmp_path / "sharded.npz")
    iv = 1.0 / (1.0 + (np.arange(D) % 10))
    rng = np.random.default_rng(5)
    init = rng.standard_normal((W, D))
    o = port.run("diag_gauss", iv, init, port.logp("diag_gauss", iv, init), steps, seed=3, acc=np.zeros(W), tot=np.zeros(W))
    assert np.array_equal(z["pos"], o["pos"])
    assert int(z["acc"]) == int(o["acc"].sum())
    from mcmcpp_b200 import parallel
    for r in range(world):
        assert np.array_equal(z["logp_owned"][r], o["logp"][parallel.owned_walkers(W, r, world)])


def test_shard_arithmetic():
    from mcmcpp_b200 import parallel
    W = 48
    for world in (1, 2, 4, 8):
        K = W // 2
        if K % world:
            with pytest.raises(ValueError):
                parallel.shard_slots(W, 0, world)
            continue
        seen = []
        for r in range(world):
            seen += parallel.owned_walkers(W, r, world)
        assert sorted(seen) == list(range(W))
    assert parallel.sequential_draw_index(0, 0, 0, W) == 0
    assert parallel.sequential_draw_index(2, 1, 5, W) == 3 * (2 * W + W // 2 + 5)


def _dim_worker(rank, world, port_no, W, D, steps, out_dir):
    """One rank of the DIM (parameter-column) sharding protocol of dim_shard_kernels.cuh, restated on the host: every rank
    holds ALL walkers x its columns, replays the same draws, builds its slice of every proposal, contributes one partial
    sum per walker to every rank's table (all_gather), adds the partials in RANK ORDER and takes the same decisions."""
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port_no)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle.pyoracle import Port
    from mcmcpp_b200 import parallel
    port = Port()
    K = W // 2
    iv = 1.0 / (1.0 + (np.arange(D) % 10))
    full = np.random.default_rng(7).standard_normal((W, D))
    logp = port.logp("diag_gauss", iv, full)
    # the draws depend on the rng alone: take (j, z, log(1-u)) of every update from the sequential restatement's trace
    tr = port.run("diag_gauss", iv, full, logp, steps, seed=11, want_chain=False, want_trace=True)["trace"]
    lo, hi = parallel.dim_slice(D, rank, world)
    mine = full[:, lo:hi].copy()                                   # this rank's columns of every walker
    accepted = 0
    for t in range(steps):
        for h in (0, 1):
            d = tr[t * W + h * K:t * W + h * K + K]
            own = mine[h * K:(h + 1) * K]
            sel = mine[(1 - h) * K + d["j"]]
            prop = sel + d["z"][:, None] * (own - sel)
            partial = torch.from_numpy((prop * prop * iv[lo:hi]).sum(axis=1))
            table = [torch.empty_like(partial) for _ in range(world)]
            dist.all_gather(table, partial)                        # "stores into every rank's table" + rendezvous
            total = table[0].numpy().copy()
            for r in range(1, world):
                total = total + table[r].numpy()                   # rank order, identical on every rank
            newp = -0.5 * total
            take = d["e"] < (np.log(d["z"]) * (D - 1) + newp) - logp[h * K:(h + 1) * K]
            mine[h * K:(h + 1) * K][take] = prop[take]
            logp[h * K:(h + 1) * K][take] = newp[take]
            accepted += int(take.sum())
    widths = [b - a for a, b in (parallel.dim_slice(D, r, world) for r in range(world))]
    padded = torch.zeros((W, max(widths)), dtype=torch.float64)      # gloo gathers equal shapes; slices may differ by two columns
    padded[:, :hi - lo] = torch.from_numpy(mine)
    cols = [torch.empty_like(padded) for _ in range(world)]
    dist.all_gather(cols, padded)
    cols = [c[:, :w] for c, w in zip(cols, widths)]
    counts = [torch.zeros(1, dtype=torch.int64) for _ in range(world)]
    dist.all_gather(counts, torch.tensor([accepted]))
    if rank == 0:
        np.savez(os.path.join(out_dir, "dim.npz"), pos=torch.cat(cols, dim=1).numpy(), logp=logp, acc=np.array([int(c) for c in counts]))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("W,D,steps", [(64, 10, 6), (48, 7, 5)])
def test_two_rank_dim_sharded_ensemble(tmp_path, port, W, D, steps):
    """Column slices from parallel.dim_slice (even and ragged), partial sums exchanged and added in rank order: every rank
    takes the decisions of the sequential reference; positions are exact, log-probs within 1e-12 (summation order)."""
    world = 2
    port_no = 29900 + (os.getpid() % 90)
    mp.spawn(_dim_worker, args=(world, port_no, W, D, steps, str(tmp_path)), nprocs=world, join=True)
    z = np.load(tmp_path / "dim.npz")
    iv = 1.0 / (1.0 + (np.arange(D) % 10))
    init = np.random.default_rng(7).standard_normal((W, D))
    o = port.run("diag_gauss", iv, init, port.logp("diag_gauss", iv, init), steps, seed=11, acc=np.zeros(W), tot=np.zeros(W))
    assert np.array_equal(z["pos"], o["pos"])
    assert np.allclose(z["logp"], o["logp"], rtol=1e-12, atol=0)
    assert z["acc"][0] == z["acc"][1] == int(o["acc"].sum())          # replicated decisions: every rank counts them all


def _data_worker(rank, world, port_no, W, P, M, steps, out_dir):
    """One rank of the DATA sharding protocol (C5, csrc/linreg.cu): the ensemble is replicated, a rank holds a contiguous
    range of the data rows, contributes its partial sum of squared residuals of every proposal, and the partials are
    all-reduced (the runtime's ncclAllReduce of K doubles); every rank then takes the same decisions."""
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port_no)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle.pyoracle import Port
    from mcmcpp_b200 import parallel
    port = Port()
    K = W // 2
    g = np.random.default_rng(3)
    X = g.standard_normal((M, P)); y = X @ g.standard_normal(P) + g.standard_normal(M)
    params = np.concatenate([[float(M), 100.0], X.reshape(-1), y])
    pos = np.linalg.lstsq(X, y, rcond=None)[0] + 0.1 * g.standard_normal((W, P))
    logp = port.logp("linreg", params, pos)
    tr = port.run("linreg", params, pos, logp, steps, seed=2, want_chain=False, want_trace=True)["trace"]
    b, e = parallel.shard_data_rows(M, rank, world)
    Xr, yr = X[b:e], y[b:e]
    accepted = 0
    for t in range(steps):
        for h in (0, 1):
            d = tr[t * W + h * K:t * W + h * K + K]
            own = pos[h * K:(h + 1) * K]
            sel = pos[(1 - h) * K + d["j"]]
            prop = sel + d["z"][:, None] * (own - sel)
            res = prop @ Xr.T - yr                                   # this rank's rows of the residual matrix [K x rows]
            sse = torch.from_numpy((res * res).sum(axis=1))
            dist.all_reduce(sse)                                     # K doubles per colour
            newp = -0.5 * sse.numpy() - 0.5 * (prop * prop).sum(axis=1) / 100.0
            take = d["e"] < (np.log(d["z"]) * (P - 1) + newp) - logp[h * K:(h + 1) * K]
            pos[h * K:(h + 1) * K][take] = prop[take]
            logp[h * K:(h + 1) * K][take] = newp[take]
            accepted += int(take.sum())
    np.savez(os.path.join(out_dir, "data%d.npz" % rank), pos=pos, logp=logp, acc=accepted)
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_data_sharded_ensemble(tmp_path, port):
    W, P, M, steps = 32, 4, 301, 6                                   # 301 rows: ragged shards (151 + 150)
    world = 2
    port_no = 29800 + (os.getpid() % 90)
    mp.spawn(_data_worker, args=(world, port_no, W, P, M, steps, str(tmp_path)), nprocs=world, join=True)
    g = np.random.default_rng(3)
    X = g.standard_normal((M, P)); y = X @ g.standard_normal(P) + g.standard_normal(M)
    params = np.concatenate([[float(M), 100.0], X.reshape(-1), y])
    init = np.linalg.lstsq(X, y, rcond=None)[0] + 0.1 * g.standard_normal((W, P))
    o = port.run("linreg", params, init, port.logp("linreg", params, init), steps, seed=2, acc=np.zeros(W), tot=np.zeros(W))
    z = [np.load(tmp_path / ("data%d.npz" % r)) for r in range(world)]
    assert np.array_equal(z[0]["pos"], z[1]["pos"]) and np.array_equal(z[0]["logp"], z[1]["logp"])   # replicas stay identical
    assert np.array_equal(z[0]["pos"], o["pos"]) and int(z[0]["acc"]) == int(o["acc"].sum())
    assert np.allclose(z[0]["logp"], o["logp"], rtol=1e-12, atol=0)


def test_bind_host_to_device_changes_nothing_without_nvml_device():
    """No GPU here: the NUMA binding helper must leave the process where it is and say so."""
    import os
    from mcmcpp_b200 import parallel
    before = os.sched_getaffinity(0)
    assert parallel.bind_host_to_device("00000000:01:00.0") is None
    assert os.sched_getaffinity(0) == before
