"""The reference's OWN test program on the B200 (-m gpu): test/sequential/SkewedGaussian/StretchMove/src/main.cpp, compiled
unmodified with nvcc against the drop-in headers (tests/dropin/Makefile -> tests/_built/skewed_stretch_b200; the include path
swaps the sampler/mover headers, a sed shadow adds __host__ __device__ to the Calculator).  Everything after runMCMC in that
program is the reference's own host code -- sliceAndBurnChain (Chain::resetChainForSubSampling), ChainPsetIterator,
AutoCorrCalc, CovarianceMatrix, CornerHistograms, PercentileAndMaximumFinder -- running on the reference's own
Chain::Chain that the device sampler filled.

Expected output = the unmodified program built with g++ and run on the host (oracle/_ref/skewed_stretch_cpu, 55 s), kept as
tests/golden/ref_main_skewed_stretch.txt (stdout) and .md5 (the files it writes: 12.8 M chain points as text, histograms).
384,182,720 walker updates: acceptance counters exact, every printed statistic identical, every written file identical."""
import hashlib
import os
import subprocess
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "tests", "_built", "skewed_stretch_b200")
GOLD = os.path.join(ROOT, "tests", "golden", "ref_main_skewed_stretch")


def test_reference_main_program_on_the_device(tmp_path):
    assert os.path.exists(EXE), "tests/_built/skewed_stretch_b200 missing: run __graft_entry__.build() where /root/reference exists"
    r = subprocess.run([EXE], cwd=tmp_path, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    want = open(GOLD + ".txt").read().splitlines()
    got = r.stdout.splitlines()
    assert "Acceptance Fraction: 274789004/384182720 | 0.715256" in got          # SURVEY.md s.4 probe of the reference
    assert got == want, "\n".join(l for l in got if l not in want)
    for line in open(GOLD + ".md5"):
        digest, name = line.split()
        h = hashlib.md5()
        with open(tmp_path / name, "rb") as f:
            for block in iter(lambda: f.read(1 << 22), b""):
                h.update(block)
        assert h.hexdigest() == digest, name
