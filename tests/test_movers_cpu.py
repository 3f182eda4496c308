"""CPU: the DifferentialEvolution restatement (oracle/oracle.c oracle_de_run) against fixtures made by the UNMODIFIED
reference mover (MCMCpp/Movers/DifferentialEvolution.h in the sequential EnsembleSampler, oracle/ref_driver.cpp)."""
import os
import numpy as np
import pytest
from conftest import GOLDEN


@pytest.mark.parametrize("name", ["de_skewed_w16", "de_diag_d5_w64"])
def test_differential_evolution_restatement_bit_exact(port, name):
    z = np.load(os.path.join(GOLDEN, "movers", name + ".npz"))
    kind, params, init, lp, n = str(z["kind"]), z["params"], z["init"], z["logp"], int(z["nsteps"])
    o = port.run_de(kind, params, init, lp, n)                       # one stream (seed 0, stream 0), walker order
    full = np.concatenate([init[None], o["chain"]])
    assert np.array_equal(full[z["rows"]], z["chain"])
    assert (int(o["acc"].sum()), int(o["tot"].sum())) == (int(z["acc"]), int(z["tot"]))
    # with 8 walkers per colour the second index repeats the first every 8th update on average: the do/while is exercised
    tr = port.run_de(kind, params, init, lp, 5, want_trace=True)["trace"]
    assert np.all(tr["j"] != tr["z"].astype(int))


@pytest.mark.parametrize("name", ["walk_skewed_w16_s3", "walk_diag_d5_w64_s6", "walk_skewed_w32_s16"])
def test_walk_move_restatement_bit_exact(port, name):
    """WalkMove restatement (oracle_walk_run: selection sampling, libstdc++'s polar normal_distribution with its saved
    second value, the reference's summation order) against the unmodified reference mover; the last case selects ALL
    16 walkers of the other colour (numPoints == numWalkers, WalkMove.h:119)."""
    z = np.load(os.path.join(GOLDEN, "movers", name + ".npz"))
    kind, params, init, lp, n, S = str(z["kind"]), z["params"], z["init"], z["logp"], int(z["nsteps"]), int(z["num_points"])
    o = port.run_walk(kind, params, init, lp, n, S)
    full = np.concatenate([init[None], o["chain"]])
    assert np.array_equal(full[z["rows"]], z["chain"])
    assert (int(o["acc"].sum()), int(o["tot"].sum())) == (int(z["acc"]), int(z["tot"]))
