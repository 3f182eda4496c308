"""World > 1 on a ONE-GPU box (-m gpu), the safe way: all ranks of a DIM-sharded ensemble live in this process and every
rank's update kernel runs inside ONE cooperative launch per colour (mcmcpp_sampler_run_group; CTA b works for rank
b % world, with exactly the device code one process per GPU runs -- dimShardFusedBody).  Ranks that wait on one another
must never be separate launches on one GPU (B200_PROFILING.md), which rules out two torchrun processes on device 0.

What it proves about the N > 1 path the SCALE benchmark runs: per-rank column slices (even: TMA ring, odd: plain loads),
partial sums handed over through the slot protocol of every rank's table, sums taken in rank order, identical decisions
on every rank.  Same assertions as tests/multigpu_check.py: decisions, counters and trace exact; this rank's columns of
positions and chain bit-exact; log-probs <= 1e-12 relative (slice-by-slice summation), against the sequential oracle."""
import numpy as np
import pytest
from test_gpu_parity import seeded_init

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("world,W,D,steps,interval", [(2, 2048, 1000, 4, 1), (2, 256, 37, 5, 1), (2, 1024, 250, 3, 2), (3, 512, 100, 6, 1),
                                                     (4, 2048, 500, 3, 1), (4, 4096, 1000, 2, 1), (3, 384, 37, 5, 1)])
def test_dim_sharded_ranks_in_one_cooperative_launch(port, world, W, D, steps, interval):
    import mcmcpp_b200 as m
    from mcmcpp_b200 import parallel
    iv = 1.0 / (1.0 + (np.arange(D) % 10))
    init, lp = seeded_init(port, "diag_gauss", iv, W, D, 23)
    o = port.run("diag_gauss", iv, init, lp, steps, interval=interval, seed=6, want_trace=True)
    ranks = [m.EnsembleSampler(6, W, D, m.StretchMove(D, 6, m.DiagGaussian(iv)), rank=r, world=world, shardMode="dim", traceUpdates=W)
             for r in range(world)]
    group = parallel.RankGroup(ranks)
    for s in ranks:
        if interval > 1:
            s.setSlicingMode(True, interval)
        s.setInitialWalkerPos(init, lp)
    assert group.runMCMC(steps)
    for r, s in enumerate(ranks):
        b, e = parallel.dim_slice(D, r, world)
        pos, aux = s.getWalkerState()
        assert np.array_equal(pos[:, b:e], o["pos"][:, b:e])                       # this rank's columns, bit for bit
        assert np.allclose(aux, o["logp"], rtol=1e-12, atol=0)
        assert np.array_equal(aux, ranks[0].getWalkerState()[1])                    # every rank holds the same log-probs
        assert s._counts() == (int(o["acc"].sum()), int(o["tot"].sum()))
        tr = s.getTrace(W)
        assert np.array_equal(tr["j"], o["trace"]["j"][:W]) and np.array_equal(tr["accept"], o["trace"]["accept"][:W])
        assert np.array_equal(tr["z"], o["trace"]["z"][:W])
        chain = s.getChain()
        assert chain.shape[0] == steps + 1
        assert np.array_equal(chain[1:, :, b:e], o["chain"][:, :, b:e]) and np.array_equal(chain[0, :, b:e], init[:, b:e])
        assert s.getRngRedraws() == 0
    for s in ranks:
        s.close()


def test_rank_group_refuses_what_it_cannot_run(port):
    import mcmcpp_b200 as m
    from mcmcpp_b200 import parallel
    from mcmcpp_b200._capi import McmcppError
    iv = np.ones(100)
    a = m.EnsembleSampler(0, 512, 100, m.StretchMove(100, 0, m.DiagGaussian(iv)), rank=0, world=2, shardMode="dim")
    b = m.EnsembleSampler(0, 512, 100, m.StretchMove(100, 0, m.DiagGaussian(iv)), rank=0, world=2, shardMode="dim")      # wrong rank
    with pytest.raises(McmcppError):
        parallel.RankGroup([a, b])
    a.close(); b.close()
