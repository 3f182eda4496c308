"""CPU tests: pin the oracle (oracle/oracle.c) against the reference's own known-answer vectors,
the survey probes, the golden fixtures made from the unmodified reference, and -- when
oracle/_ref is present -- the compiled reference itself."""
import numpy as np
import pytest
from conftest import load_golden, golden_names

# MCMCpp/Utility/pcg-cpp/test-high/expected/check-pcg64.out, check-pcg32.out (Round 1, rng(42,54))
PCG64_42_54 = [0x86b1da1d72062b68, 0x1304aa46c9853d39, 0xa3670e9e0dd50358, 0xf9090e529a7dae00, 0xc85b9fd837996f2c,
               0x606121f8e3919196]
PCG32_42_54 = [0xa15c02b7, 0x7b47f409, 0xba1d3330, 0x83d2f293, 0xbfa4784b, 0xcbed606e]
# SURVEY.md s.8(a)/s.10 probes of the compiled reference
PCG64_0_0 = [0xd4feb4e5a4bcfe09, 0xe85a7fe071b026e6, 0x3a5b9037fe928c11, 0x7b044380d100f216, 0x1c7850a6b6d83e6a,
             0x240b82fcc04f0926]
PCG64_42_3 = [0x0807e83a2bdd2e78, 0xaaac17d0cc16f5c7, 0x9105a4f711201e29, 0x2f345982557026fa]
PCG32_53 = [0x648600c7, 0x63608e0e, 0xdb2e5e9f, 0xbb293267]


def test_pcg_known_answers(port):
    assert port.pcg64_raw(42, 54, 6).tolist() == PCG64_42_54
    assert port.pcg32_raw(42, 54, 6).tolist() == PCG32_42_54
    assert port.pcg64_raw(0, 0, 6).tolist() == PCG64_0_0
    assert port.pcg64_raw(42, 3, 4).tolist() == PCG64_42_3
    assert port.pcg32_raw(53, None, 4).tolist() == PCG32_53


def test_pcg64_advance_matches_stepping(port):
    seq = port.pcg64_raw(7, 11, 300)
    for skip in (0, 1, 2, 63, 64, 255):
        assert port.pcg64_raw(7, 11, 5, skip=skip).tolist() == seq[skip:skip + 5].tolist()


def test_stretch_draw_known_answers(port):
    # seed(0,0), K=160 -> j; and the first three (z, log(1-u)) of the stream
    j, z, e = port.draws(0, 0, 160, 3)
    assert j.tolist() == [105, 150, 38]
    assert [float.hex(v) for v in z] == ["0x1.d1cc93c2ed900p+0", "0x1.3c1b2b7d9b63cp-1", "0x1.994069f20ce4fp-1"]
    assert [float.hex(v) for v in e] == ["-0x1.08ed7dd3244b2p-2", "-0x1.36caf79505832p-3", "-0x1.fc6a19cd94141p-7"]
    j16, _, _ = port.draws(0, 0, 16, 3)
    assert j16.tolist() == [9, 6, 6]          # SURVEY.md s.8(c) design-validation trace (W=32)


def test_c1_known_answer(port):
    g = load_golden("c1_skewed_w32")
    r = port.run(g["kind"], g["params"], g["init"], g["logp"], g["nsteps"], want_trace=True)
    assert (int(r["acc"].sum()), int(r["tot"].sum())) == (228883, 320032) == (g["acc"], g["tot"])
    assert [float.hex(v) for v in r["chain"][-1, 0]] == ["0x1.754c17069f6f6p+0", "0x1.22e26142efda4p+0"]
    t = r["trace"][:3]
    assert t["j"].tolist() == [9, 6, 6] and t["accept"].tolist() == [0, 1, 1]
    assert float.hex(float(t["newp"][0])) == "-0x1.cd148759b0981p+4"
    assert float.hex(float(t["diff"][1])) == "0x1.9af42ac5319fap+1"
    assert r["rejections"] == 0


@pytest.mark.parametrize("name", golden_names())
def test_port_reproduces_reference_goldens(port, name):
    g = load_golden(name)
    r = port.run(g["kind"], g["params"], g["init"], g["logp"], g["nsteps"], interval=g["interval"], seed=g["seed"])
    full = np.concatenate([g["init"][None], r["chain"]])
    assert np.array_equal(full[g["rows"]], g["chain"])            # bit-exact positions
    assert (int(r["acc"].sum()), int(r["tot"].sum())) == (g["acc"], g["tot"])
    assert np.array_equal(g["logp"], port.logp(g["kind"], g["params"], g["init"]))


def test_port_against_live_reference(port):
    from oracle.pyoracle import Ref
    if not Ref.available():
        pytest.skip("oracle/_ref not built (no /root/reference here)")
    R = Ref()
    assert R.pcg64_raw(0, 0, 64, skip=5).tolist() == port.pcg64_raw(0, 0, 64, skip=5).tolist()
    jr, zr, er = R.draws(3, 0, 1000, 5000)
    jp, zp, ep = port.draws(3, 0, 1000, 5000)
    assert np.array_equal(jr, jp) and np.array_equal(zr, zp) and np.array_equal(er, ep)
    init = R.init_normal(7, 64 * 2).reshape(64, 2)
    lp = R.logp("rosenbrock_2d", None, init)
    a = R.run("rosenbrock_2d", None, init, lp, 300, interval=2, seed=5)
    b = port.run("rosenbrock_2d", None, init, lp, 300, interval=2, seed=5)
    assert np.array_equal(a["chain"][1:], b["chain"]) and a["acc"] == int(b["acc"].sum())


def test_per_walker_streams_are_independent_of_order(port):
    # PER_WALKER discipline: walker w's draws come from pcg64(seed, w) only
    g = load_golden("rosenbrock_w512")
    r = port.run(g["kind"], g["params"], g["init"], g["logp"], 3, per_walker=True, want_trace=True, seed=9)
    K = g["W"] // 2
    j, z, e = port.draws(9, 5, K, 3)          # walker 5 = red slot 5
    tr = r["trace"].reshape(3, g["W"])
    assert tr["j"][:, 5].tolist() == j.tolist() and np.array_equal(tr["z"][:, 5], z) and np.array_equal(tr["e"][:, 5], e)


# ---- PER_WALKER32: the fast engine (pcg32 per walker) --------------------------------------------------------------
import glob, os
from conftest import GOLDEN


def pcg32_golden_names():
    return sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, "pcg32", "*.npz")))


def load_pcg32_golden(name):
    z = np.load(os.path.join(GOLDEN, "pcg32", name + ".npz"))
    d = {k: z[k] for k in z.files}
    for k in ("W", "D", "nsteps", "interval", "seed", "acc", "tot"):
        d[k] = int(d[k])
    d["kind"] = str(d["kind"])
    return d


@pytest.mark.parametrize("name", pcg32_golden_names())
def test_port_reproduces_pcg32_goldens(port, name):
    """oracle_stretch_run_per_walker32 == vendored pcg32 + libstdc++ adapters + reference GwDistribution, bit for bit."""
    g = load_pcg32_golden(name)
    r = port.run(g["kind"], g["params"], g["init"], g["logp"], g["nsteps"], interval=g["interval"], seed=g["seed"], per_walker=32)
    full = np.concatenate([g["init"][None], r["chain"]])
    assert np.array_equal(full[g["rows"]], g["chain"])
    assert np.array_equal(r["logp"], g["final_logp"])
    assert (int(r["acc"].sum()), int(r["tot"].sum())) == (g["acc"], g["tot"])


def test_pcg32_live_reference(port):
    from oracle.pyoracle import Ref
    if not Ref.available():
        pytest.skip("oracle/_ref not built (no /root/reference here)")
    R = Ref()
    assert R.pcg32_raw(42, 54, 64).tolist() == port.pcg32_raw(42, 54, 64).tolist()
    init = R.init_normal(7, 64 * 2).reshape(64, 2)
    lp = R.logp("rosenbrock_2d", None, init)
    a = R.run_pcg32("rosenbrock_2d", None, init, lp, 200, interval=2, seed=5)
    b = port.run("rosenbrock_2d", None, init, lp, 200, interval=2, seed=5, per_walker=32)
    assert np.array_equal(a["chain"][1:], b["chain"]) and a["acc"] == int(b["acc"].sum())


def test_set_initial_twice_continues_the_stream(port):
    """The reference seeds its mover once, in the constructor (EnsembleSampler.h:217); setInitialWalkerPos never reseeds
    (EnsembleSampler.h:220-230).  The oracle mirrors it by carrying the engine between calls; checked live against the reference
    in test_gpu_api (device) -- here: two runs on one engine differ from two runs on fresh engines."""
    g = load_golden("rosenbrock_w512")
    a = port.run(g["kind"], g["params"], g["init"], g["logp"], 5, seed=3)
    b = port.run(g["kind"], g["params"], g["init"], g["logp"], 5, seed=3, rng=a["rng"])
    c = port.run(g["kind"], g["params"], g["init"], g["logp"], 5, seed=3)
    assert np.array_equal(a["pos"], c["pos"]) and not np.array_equal(a["pos"], b["pos"])
