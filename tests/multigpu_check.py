"""Run under torchrun (one process per GPU): a walker-sharded ensemble in both shard modes must equal the
sequential oracle bit for bit on every rank's owned rows, for the thread, warp and TMA-bulk kernels.
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 tests/multigpu_check.py"""
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


def main():
    rank, world, local = parallel.env_rank_world()
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    port = Port()
    failures = 0
    cases = [("rosenbrock_2d", 4096, 2, 12), ("diag_gauss", 512, 100, 6), ("diag_gauss", 2048, 1000, 4), ("diag_gauss", 256, 37, 5)]
    for kind, W, D, steps in cases:
        params = 1.0 / (1.0 + (np.arange(D) % 10)) if kind == "diag_gauss" else np.zeros(0)
        rng = np.random.default_rng(17)
        init = rng.standard_normal((W, D))
        lp = port.logp(kind, params if params.size else None, init)
        o = port.run(kind, params if params.size else None, init, lp, steps, seed=2, want_chain=True)
        calc = m.DiagGaussian(params) if kind == "diag_gauss" else m.RosenbrockTwoDim()
        owned = parallel.owned_walkers(W, rank, world)
        for mode in ("allgather", "peer"):
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
    t = torch.tensor([failures], device="cuda")
    dist.all_reduce(t)
    dist.destroy_process_group()
    if int(t.item()):
        raise SystemExit(1)
    if rank == 0:
        print("MULTIGPU_OK", flush=True)


if __name__ == "__main__":
    main()
