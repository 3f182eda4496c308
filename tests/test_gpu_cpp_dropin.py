"""The C++ drop-in headers (include/mcmcpp_b200/EnsembleSampler.h) driven like the reference's own test program
(test/sequential/SkewedGaussian/StretchMove/src/main.cpp:22-80) with a USER-defined Calculator whose kernels are
instantiated in the user's translation unit (examples/skewed_gaussian.cu, built by __graft_entry__.build())."""
import os
import re
import subprocess
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "examples", "skewed_gaussian")


def example_init(W):
    w = np.arange(W)
    init = np.stack([0.001 * ((w * 7919) % 2003) - 1.0, 0.002 * ((w * 104729) % 1009) - 1.0], axis=1)
    return init


def run_example(*args):
    assert os.path.exists(EXE), "examples/skewed_gaussian missing: run __graft_entry__.build()"
    out = subprocess.run([EXE, *map(str, args)], check=True, capture_output=True, text=True, timeout=300).stdout
    acc, tot, stored = map(int, re.search(r"accepted (\d+) total (\d+) stored (\d+)", out).groups())
    mlast = re.search(r"last (\S+) (\S+)", out)
    last = [float.fromhex(x) for x in mlast.groups()] if mlast else None
    cov = [float(x) for x in re.search(r"cov (\S+) (\S+) (\S+)", out).groups()]
    return acc, tot, stored, last, cov, out


@pytest.mark.parametrize("W,steps,slicing", [(320, 300, 3), (32, 500, 1)])
def test_cpp_dropin_sequential_is_bit_exact(port, W, steps, slicing):
    acc, tot, stored, last, cov, out = run_example("--walkers", W, "--steps", steps, "--slicing", slicing, "--dump")
    init = example_init(W)
    lp = port.logp("skewed_gauss_2d", [0.13], init)
    o = port.run("skewed_gauss_2d", [0.13], init, lp, steps, interval=slicing, seed=0)
    assert "status complete" in out
    assert (acc, tot) == (int(o["acc"].sum()), int(o["tot"].sum()))
    assert stored == steps + 1 - 20                                   # sliceAndBurnChain(1, 20)
    assert last == o["chain"][-1, 0].tolist()


def test_cpp_dropin_parallel_front_end(port):
    # one "thread" = the sequential stream (bit-identical, as in the reference); more = per-walker streams
    a1 = run_example("--walkers", 320, "--steps", 200, "--slicing", 1, "--threads", 1, "--dump")
    a0 = run_example("--walkers", 320, "--steps", 200, "--slicing", 1, "--dump")
    assert a1[:4] == a0[:4]
    init = example_init(320)
    lp = port.logp("skewed_gauss_2d", [0.13], init)
    o = port.run("skewed_gauss_2d", [0.13], init, lp, 200, seed=0, per_walker=True)
    a8 = run_example("--walkers", 320, "--steps", 200, "--slicing", 1, "--threads", 8, "--dump")
    assert (a8[0], a8[1]) == (int(o["acc"].sum()), int(o["tot"].sum())) and a8[3] == o["chain"][-1, 0].tolist()


def test_cpp_dropin_statistics():
    # long run: covariance of the target (analytic truth quoted in test/sequential/SkewedGaussian/MH/src/main.cpp:74)
    *_, cov, out = run_example("--walkers", 320, "--steps", 4000, "--slicing", 5)
    assert abs(cov[0] - 1.13) < 0.02 and abs(cov[1] - 0.435) < 0.01 and abs(cov[2] - 0.2825) < 0.006, out


def test_cpp_dropin_differential_evolution_is_bit_exact(port):
    """MCMC::Mover::DifferentialEvolution<double, UserCalc> in the sampler template (DifferentialEvolution.h:57): one
    stream per walker, every draw, decision and stored row equal to the oracle's."""
    W, steps = 320, 150
    acc, tot, stored, last, cov, out = run_example("--walkers", W, "--steps", steps, "--slicing", 1, "--mover", "de", "--dump")
    init = example_init(W)
    lp = port.logp("skewed_gauss_2d", [0.13], init)
    o = port.run_de("skewed_gauss_2d", [0.13], init, lp, steps, seed=0, per_walker=True)
    assert (acc, tot) == (int(o["acc"].sum()), int(o["tot"].sum()))
    assert stored == steps + 1 - 20 and last == o["chain"][-1, 0].tolist()


def test_cpp_dropin_walk_move(port):
    """MCMC::Mover::WalkMove<double, UserCalc>(numParams, prngInit, calc, numSamples) (WalkMove.h:59): decisions equal
    to the oracle's, positions within the tolerance tier (the normals go through log(); a last-bit difference there is
    amplified by every later proposal that uses the row -- each is a sum over S partners -- so the run is kept short
    and the bound is 1e-10)."""
    W, steps, S = 64, 30, 3
    acc, tot, stored, last, cov, out = run_example("--walkers", W, "--steps", steps, "--slicing", 1, "--mover", "walk", "--points", S, "--dump")
    init = example_init(W)
    lp = port.logp("skewed_gauss_2d", [0.13], init)
    o = port.run_walk("skewed_gauss_2d", [0.13], init, lp, steps, S, seed=0, per_walker=True)
    assert (acc, tot) == (int(o["acc"].sum()), int(o["tot"].sum()))
    assert np.allclose(last, o["chain"][-1, 0], rtol=1e-10, atol=1e-10 * np.abs(o["chain"]).max())


def test_cpp_dropin_post_step_action_call_pattern(port):
    """PostStepAction as the reference drives it (EnsembleSampler.h:289-308,356-359): performAction after every
    performStep, saved or not, with the range of steps counted so far -- the step just written is not in it yet."""
    W, steps, k = 320, 40, 3
    acc, tot, stored, last, cov, out = run_example("--walkers", W, "--steps", steps, "--slicing", k, "--action", "--dump")
    plain = run_example("--walkers", W, "--steps", steps, "--slicing", k, "--dump")
    assert (acc, tot, stored, last) == plain[:4]                       # one device call per stored step changes nothing
    calls, rng, first = re.search(r"action calls (\d+) last_range (\d+) last_first (\S+)", out).groups()
    init = example_init(W)
    o = port.run("skewed_gauss_2d", [0.13], init, port.logp("skewed_gauss_2d", [0.13], init), steps, interval=k, seed=0)
    assert int(calls) == steps * k and int(rng) == steps               # initial row + (steps - 1) counted steps
    assert float.fromhex(first) == o["chain"][steps - 2, 0, 0]
