"""CPU tests of the drop-in boundary: the C-ABI library loads, exports every symbol
include/mcmcpp_b200/capi.h declares, refuses to run without a GPU (no CPU fallback), and the
__host__ form of the device RNG header reproduces the pcg-cpp known answers."""
import ctypes as C
import os
import re
import subprocess
import sys
import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "mcmcpp_b200", "capi.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(mcmcpp_[a-z0-9_]+)\s*\(", text)) - {"mcmcpp_halfstep_fn"})


def test_library_exports_every_declared_symbol():
    from mcmcpp_b200 import _capi
    L = _capi.lib()
    names = declared_symbols()
    assert len(names) >= 25
    for n in names:
        assert hasattr(L, n), n
    assert sorted(_capi.SYMBOLS) == names      # the ctypes table and the header agree


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    import mcmcpp_b200 as m
    from mcmcpp_b200._capi import McmcppError, ERR_CUDA
    with pytest.raises(McmcppError) as ei:
        m.EnsembleSampler(0, 32, 2, m.StretchMove(2, 0, m.SkewedGaussianTwoDim(0.13)))
    assert ei.value.code == ERR_CUDA


def test_product_never_imports_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "mcmcpp_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                src = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in src.replace("params layout as in oracle/oracle.c", ""), os.path.join(dirpath, f)


HOST_RNG_PROG = r"""
#include <cstdio>
#include "mcmcpp_b200/device_rng.cuh"
using namespace MCMC::Device;
int main() {
    Pcg64 g = pcg64Seed(42, 54);
    for (int i = 0; i < 6; ++i) std::printf("%016llx\n", (unsigned long long)pcg64Next(g.state, g.inc));
    /* jump-ahead == stepping */
    Pcg64 a = pcg64Seed(0, 0), b = pcg64Seed(0, 0);
    for (int i = 0; i < 3000; ++i) pcg64Next(a.state, a.inc);
    b.state = applyAffine(pcg64AdvanceMap(b.inc, 3000, 0), b.state);
    std::printf("%d\n", (a.state.lo == b.state.lo && a.state.hi == b.state.hi) ? 1 : 0);
    /* the three draws, K = 160 */
    Pcg64 d = pcg64Seed(0, 0);
    GwConstants gw = makeGwConstants(2, 1);
    unsigned long long K = 160, thr = (0ULL - K) % K, recip = (unsigned long long)((((unsigned __int128)1) << 64) / K);
    for (int i = 0; i < 3; ++i) { StretchDraws s = drawStretch(d.state, d.inc, K, thr, recip, gw); std::printf("%d %a %a\n", s.j, s.z, s.logOneMinusU); }
    /* modByRecip over awkward moduli */
    unsigned long long bad = 0, x = 0x9e3779b97f4a7c15ULL;
    unsigned long long mods[] = {2, 3, 16, 160, 524288, 524287, 1000003, 2147483647ULL};
    for (unsigned long long m : mods) { unsigned long long r = (unsigned long long)((((unsigned __int128)1) << 64) / m);
        for (int i = 0; i < 100000; ++i) { x = x * 6364136223846793005ULL + 1442695040888963407ULL; if (modByRecip(x, m, r) != x % m) ++bad; }
        if (modByRecip(~0ULL, m, r) != (~0ULL) % m) ++bad; if (modByRecip(0, m, r) != 0) ++bad; }
    std::printf("%llu\n", bad);
    return 0;
}
"""


def test_device_rng_header_host_form(tmp_path):
    src = tmp_path / "rng_host.cpp"
    src.write_text(HOST_RNG_PROG)
    exe = tmp_path / "rng_host"
    subprocess.run(["g++", "-O2", "-ffp-contract=off", "-x", "c++", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)],
                   check=True)
    out = subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.split("\n")
    assert out[:6] == ["86b1da1d72062b68", "1304aa46c9853d39", "a3670e9e0dd50358", "f9090e529a7dae00", "c85b9fd837996f2c",
                       "606121f8e3919196"]
    assert out[6] == "1"
    assert out[7].split() == ["105", "0x1.d1cc93c2ed9p+0", "-0x1.08ed7dd3244b2p-2"]
    assert out[8].split() == ["150", "0x1.3c1b2b7d9b63cp-1", "-0x1.36caf79505832p-3"]
    assert out[9].split() == ["38", "0x1.994069f20ce4fp-1", "-0x1.fc6a19cd94141p-7"]
    assert out[10] == "0"


def test_dim_slices_match_the_c_abi_and_tile_the_row():
    """DIM sharding: the Python shard arithmetic (parallel.dim_slice) is the C-ABI's (mcmcpp_dim_slice); slices tile
    [0, D), and every slice but possibly the last is even (16-byte aligned rows on the device)."""
    import ctypes as C
    from mcmcpp_b200 import _capi, parallel
    L = _capi.lib()
    for D in (2, 7, 8, 50, 100, 125, 1000, 1001, 1023):
        for world in (1, 2, 3, 4, 8):
            if D // 2 < world:
                b, c = C.c_int(), C.c_int()
                assert L.mcmcpp_dim_slice(D, 0, world, C.byref(b), C.byref(c)) == _capi.ERR_ARG
                continue
            nxt = 0
            for r in range(world):
                b, c = C.c_int(), C.c_int()
                assert L.mcmcpp_dim_slice(D, r, world, C.byref(b), C.byref(c)) == 0
                assert (b.value, b.value + c.value) == parallel.dim_slice(D, r, world)
                assert b.value == nxt and c.value > 0 and b.value % 2 == 0
                assert c.value % 2 == 0 or r == world - 1
                nxt = b.value + c.value
            assert nxt == D


def test_header_is_plain_c_and_python_mirrors_its_structs(tmp_path):
    """The boundary is a C ABI: capi.h compiles as C (no C++ or torch types in any signature), and the ctypes structures
    have the sizes and field offsets the C compiler gives the header's structs."""
    import ctypes as C
    from mcmcpp_b200 import _capi
    src = tmp_path / "abi.c"
    fields = [name for name, _ in _capi.Config._fields_]
    tfields = [name for name, _ in _capi.Trace._fields_]
    prints = "".join('    printf("config.%s %%zu\\n", offsetof(mcmcpp_config, %s));\n' % (f, f) for f in fields)
    prints += "".join('    printf("trace.%s %%zu\\n", offsetof(mcmcpp_trace, %s));\n' % (f, f) for f in tfields)
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "mcmcpp_b200/capi.h"\nint main(void) {\n'
                   '    printf("sizeof.config %zu\\nsizeof.trace %zu\\n", sizeof(mcmcpp_config), sizeof(mcmcpp_trace));\n'
                   + prints + "    return 0;\n}\n")
    exe = tmp_path / "abi"
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-I" + os.path.join(ROOT, "include"), str(src), "-o", str(exe)], check=True)
    out = dict(line.split() for line in subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.splitlines())
    assert int(out["sizeof.config"]) == C.sizeof(_capi.Config) and int(out["sizeof.trace"]) == C.sizeof(_capi.Trace)
    for f in fields:
        assert int(out["config." + f]) == getattr(_capi.Config, f).offset, f
    for f in tfields:
        assert int(out["trace." + f]) == getattr(_capi.Trace, f).offset, f
