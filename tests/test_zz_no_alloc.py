"""mf_dot_product_noalloc (include/mantacuda.h; no counterpart in the reference):
the allocation-free GridDotProduct on caller-owned device arrays, checked against the existing entry point mf_grid_dot on the
same data, two calls in a row (the static scratch keeps the first call's contents), plus the size limit.
The file name sorts last on purpose: a failure here must not hide another test."""
import ctypes as C

import numpy as np
import pytest

REL_TOL = 1e-12     # both sides sum float products in double; only the (fixed) summation orders differ


def _device_array(torch, a):
    t = torch.from_numpy(a).cuda()
    torch.cuda.synchronize()
    return t


@pytest.mark.gpu
def test_noalloc_dot_matches_grid_dot_twice():
    import torch
    from mantaflow_b200 import manta as M
    from mantaflow_b200._lib import lib, check
    L = lib()
    size = (96, 80, 72)                       # 552 960 cells: 68 blocks, the last one partial
    nx, ny, nz = size
    s = M.Solver(name="main", gridSize=M.vec3(*size), dim=3)
    ga, gb = s.create(M.RealGrid), s.create(M.RealGrid)
    out_dev = torch.zeros(1, dtype=torch.float64, device="cuda")
    results = []
    for call, seed in enumerate((1, 2)):
        r = np.random.RandomState(seed)
        a = (r.rand(nz, ny, nx).astype(np.float32) - np.float32(0.5)) * np.float32(3.0)
        b = (r.rand(nz, ny, nx).astype(np.float32) - np.float32(0.25))
        ga.from_numpy(a); gb.from_numpy(b)
        want = C.c_double()
        check(L.mf_grid_dot(ga._h, gb._h, C.byref(want)))                   # the existing entry point
        ta, tb = _device_array(torch, a.ravel()), _device_array(torch, b.ravel())
        out_dev.fill_(-1.0)
        torch.cuda.synchronize()
        check(L.mf_dot_product_noalloc(C.c_void_p(ta.data_ptr()), C.c_void_p(tb.data_ptr()), a.size, C.c_void_p(out_dev.data_ptr()), None))
        torch.cuda.synchronize()
        got = float(out_dev.cpu()[0])
        prod = (a.ravel() * b.ravel()).astype(np.float64)                    # float products, double sum
        exact, scale = float(np.sum(prod)), float(np.sum(np.abs(prod)))
        assert abs(got - want.value) <= REL_TOL * scale, (call, got, want.value)
        assert abs(got - exact) <= REL_TOL * scale, (call, got, exact)
        results.append(got)
        # same inputs again: bit-identical (fixed summation order, the counter was reset)
        check(L.mf_dot_product_noalloc(C.c_void_p(ta.data_ptr()), C.c_void_p(tb.data_ptr()), a.size, C.c_void_p(out_dev.data_ptr()), None))
        torch.cuda.synchronize()
        assert float(out_dev.cpu()[0]) == got
    assert results[0] != results[1]


@pytest.mark.gpu
def test_noalloc_dot_edges_and_limit():
    import torch
    from mantaflow_b200._lib import lib, check, MantaError
    L = lib()
    out_dev = torch.full((1,), 7.0, dtype=torch.float64, device="cuda")
    a = torch.arange(1, 8193 + 1, dtype=torch.float32, device="cuda")          # one full block + one element
    ones = torch.ones_like(a)
    for n in (0, 1, 8192, 8193):
        check(L.mf_dot_product_noalloc(C.c_void_p(a.data_ptr()), C.c_void_p(ones.data_ptr()), n, C.c_void_p(out_dev.data_ptr()), None))
        torch.cuda.synchronize()
        assert float(out_dev.cpu()[0]) == n * (n + 1) / 2.0
    limit = 33554432
    with pytest.raises(MantaError, match="exceeds the limit"):
        check(L.mf_dot_product_noalloc(C.c_void_p(a.data_ptr()), C.c_void_p(ones.data_ptr()), limit + 1, C.c_void_p(out_dev.data_ptr()), None))
    with pytest.raises(MantaError):
        check(L.mf_dot_product_noalloc(None, C.c_void_p(ones.data_ptr()), 4, C.c_void_p(out_dev.data_ptr()), None))
    # at the limit itself, on a non-default stream (4096 blocks: every partial slot in use)
    big = torch.full((limit,), 0.5, dtype=torch.float32, device="cuda")
    st = torch.cuda.Stream()
    st.wait_stream(torch.cuda.current_stream())
    check(L.mf_dot_product_noalloc(C.c_void_p(big.data_ptr()), C.c_void_p(big.data_ptr()), limit, C.c_void_p(out_dev.data_ptr()), C.c_void_p(st.cuda_stream)))
    st.synchronize()
    assert float(out_dev.cpu()[0]) == 0.25 * limit
