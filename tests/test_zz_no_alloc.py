"""GPU (-m gpu), collected last on purpose: mcmcpp_covariance_device_noalloc -- the covariance kernels on a file-scope
`static __device__` scratch array instead of cudaMalloc'ed memory (include/mcmcpp_b200/capi.h; no counterpart in the reference)
-- through the C-ABI against the existing entry point mcmcpp_covariance on the same
points.  Bar: the tier the covariance tests state (tests/test_gpu_analysis.py): means 1e-13 relative, covariance entries 1e-12
of the scale of their terms, correlations 1e-11 absolute.  Several calls in a row, so that a call sees the previous call's
contents in the shared array (larger, smaller, other kernel, the maximum size)."""
import ctypes as C
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_PARAMS = 32          # MCMCPP_NOALLOC_COV_MAX_PARAMS


def existing(lib, check, pts):
    """mcmcpp_covariance over the points as a host chain of one step."""
    n, D = pts.shape
    mean, cov, corr = np.empty(D), np.empty((D, D)), np.empty((D, D))
    used = C.c_longlong(0)
    check(lib.mcmcpp_covariance(pts.ctypes.data, 1, n, D, 1, 0, mean.ctypes.data, cov.ctypes.data, corr.ctypes.data, C.byref(used)))
    assert used.value == n
    return mean, cov, corr


def no_alloc(lib, check, dev_rows, n, D, stream=None):
    mean, cov, corr = np.full(D, np.nan), np.full((D, D), np.nan), np.full((D, D), np.nan)
    check(lib.mcmcpp_covariance_device_noalloc(dev_rows.data_ptr(), n, D, 0, mean.ctypes.data, cov.ctypes.data, corr.ctypes.data, stream))
    return mean, cov, corr


def points(rng, n, D):
    mix = rng.normal(size=(D, D)) / np.sqrt(D) + np.eye(D)
    return np.ascontiguousarray(rng.normal(size=(n, D)) @ mix + rng.uniform(-3.0, 3.0, D))


def same(a, b, pts):
    m2 = (pts[:, :, None] * pts[:, None, :]).mean(axis=0) if pts.shape[1] <= 8 else np.einsum("ni,nj->ij", pts, pts) / len(pts)
    scale = np.abs(m2) + np.abs(np.outer(pts.mean(axis=0), pts.mean(axis=0))) + 1e-300
    assert np.allclose(a[0], b[0], rtol=1e-13, atol=1e-15)
    assert np.all(np.abs(a[1] - b[1]) <= 1e-12 * scale)
    assert np.allclose(a[2], b[2], rtol=0, atol=1e-11)


def test_no_alloc_covariance_equals_the_allocating_entry_point_on_calls_in_a_row():
    import torch
    from mcmcpp_b200 import _capi
    lib, check = _capi.lib(), _capi.check
    rng = np.random.default_rng(424242)
    # (points, params): item kernel, the small-D kernel on fewer pairs, the item kernel again on fewer points, the maximum size,
    # one parameter, and the first shape once more
    shapes = [(50000, 6), (30011, 2), (777, 6), (20000, MAX_PARAMS), (4096, 1), (50000, 6)]
    for n, D in shapes:
        pts = points(rng, n, D)
        dev = torch.from_numpy(pts).to("cuda:0")
        torch.cuda.synchronize()
        got = no_alloc(lib, check, dev, n, D)
        same(got, existing(lib, check, pts), pts)
        # the same call again straight away, on a stream of the caller's: the array holds this very call's pairs now
        st = torch.cuda.Stream(device="cuda:0")
        again = no_alloc(lib, check, dev, n, D, stream=st.cuda_stream)
        assert all(np.array_equal(x, y) for x, y in zip(got, again))


def test_no_alloc_covariance_refuses_what_the_array_cannot_hold():
    import torch
    from mcmcpp_b200 import _capi
    lib = _capi.lib()
    dev = torch.zeros(100 * (MAX_PARAMS + 1), dtype=torch.float64, device="cuda:0")
    out = np.empty((MAX_PARAMS + 1) ** 2)
    with pytest.raises(_capi.McmcppError) as err:
        _capi.check(lib.mcmcpp_covariance_device_noalloc(dev.data_ptr(), 100, MAX_PARAMS + 1, 0, out.ctypes.data, out.ctypes.data,
                                                         out.ctypes.data, None))
    assert "MCMCPP_NOALLOC_COV_MAX_PARAMS" in str(err.value)
    with pytest.raises(_capi.McmcppError):
        _capi.check(lib.mcmcpp_covariance_device_noalloc(None, 100, 2, 0, out.ctypes.data, out.ctypes.data, out.ctypes.data, None))
