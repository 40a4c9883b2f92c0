"""Kernel-level parity of the pressure-matrix pieces (C-ABI entries mf_make_laplace_matrix / mf_apply_matrix /
mf_apply_laplace_rowcode) against a numpy restatement of MakeLaplaceMatrix (conjugategrad.h:154-187) and ApplyMatrix
(conjugategrad.h:118-151, :205-222).  All comparisons are bit-exact: coefficients are small integers and the row sums are
evaluated in the reference's order with separate multiplies and adds.
"""
import ctypes as Ct

import numpy as np
import pytest

FLUID, OBSTACLE, EMPTY = 1, 2, 4


def _flags(rs, size, dim, border_fluid=False):
    nx, ny, nz = size
    f = rs.choice([FLUID, OBSTACLE, EMPTY], size=(nz, ny, nx), p=[0.7, 0.15, 0.15]).astype(np.int32)
    if not border_fluid:
        f[:, :, 0] = f[:, :, -1] = OBSTACLE
        f[:, 0, :] = f[:, -1, :] = OBSTACLE
        if dim == 3:
            f[0] = f[-1] = OBSTACLE
    return f


def _np_laplace(f, dim):
    """MakeLaplaceMatrix without fractions: rows of interior (bnd 1) fluid cells only"""
    nz, ny, nx = f.shape
    A0 = np.zeros(f.shape, np.float32); Ai = np.zeros_like(A0); Aj = np.zeros_like(A0); Ak = np.zeros_like(A0)
    inter = np.zeros(f.shape, bool)
    if dim == 3:
        inter[1:-1, 1:-1, 1:-1] = True
    else:
        inter[:, 1:-1, 1:-1] = True
    row = inter & ((f & FLUID) != 0)
    def sh(a, axis, d):   # value of the neighbour in direction d along axis (wrap is irrelevant: only interior rows are used)
        return np.roll(a, -d, axis=axis)
    axes = [(2, Ai), (1, Aj)] + ([(0, Ak)] if dim == 3 else [])
    for ax, Ap in axes:
        for d in (-1, 1):
            nb = sh(f, ax, d)
            A0 += (row & ((nb & OBSTACLE) == 0)).astype(np.float32)
        Ap[row & ((sh(f, ax, 1) & FLUID) != 0)] = -1.0
    return A0, Ai, Aj, Ak


def _np_apply(f, src, A0, Ai, Aj, Ak, dim):
    """ApplyMatrix: linear-index stencil, neighbours outside the array contribute nothing"""
    nz, ny, nx = f.shape
    N = f.size
    s = src.ravel(); a0 = A0.ravel(); ai = Ai.ravel(); aj = Aj.ravel(); ak = Ak.ravel()
    X, Y, Z = 1, nx, nx * ny
    def lo(v, o):   # v[idx - o] (0 where idx - o < 0)
        r = np.zeros_like(v); r[o:] = v[:N - o]; return r
    def hi(v, o):   # v[idx + o] (0 where idx + o >= N)
        r = np.zeros_like(v); r[:N - o] = v[o:]; return r
    r = s * a0
    r = r + lo(s, X) * lo(ai, X)
    r = r + hi(s, X) * ai
    r = r + lo(s, Y) * lo(aj, Y)
    r = r + hi(s, Y) * aj
    if dim == 3:
        r = r + lo(s, Z) * lo(ak, Z)
        r = r + hi(s, Z) * ak
    out = np.where((f.ravel() & FLUID) != 0, r, s).astype(np.float32)
    return out.reshape(f.shape)


SIZES = [((64, 48, 40), 3), ((8, 8, 8), 3), ((4, 3, 3), 3), ((128, 20, 1), 2), ((36, 5, 7), 3)]


@pytest.mark.gpu
@pytest.mark.parametrize("size,dim", SIZES)
@pytest.mark.parametrize("border_fluid", [False, True])
def test_laplace_matrix_and_apply(size, dim, border_fluid):
    from mantaflow_b200 import manta
    from mantaflow_b200._lib import lib, check
    rs = np.random.RandomState(11)
    s = manta.Solver(name="s", gridSize=manta.vec3(*size), dim=dim)
    f = _flags(rs, size, dim, border_fluid)
    src = (rs.rand(*f.shape).astype(np.float32) - np.float32(0.5))
    flags = s.create(manta.FlagGrid); flags.from_numpy(f)
    g = [s.create(manta.RealGrid) for _ in range(7)]
    A0, Ai, Aj, Ak, S, D, D2 = g
    S.from_numpy(src)
    check(lib().mf_make_laplace_matrix(flags._h, A0._h, Ai._h, Aj._h, Ak._h, None))
    ref = _np_laplace(f, dim)
    for ours, r, nm in zip((A0, Ai, Aj, Ak), ref, "A0 Ai Aj Ak".split()):
        assert np.array_equal(ours.to_numpy(), r), nm
    check(lib().mf_apply_matrix(flags._h, D._h, S._h, A0._h, Ai._h, Aj._h, Ak._h))
    want = _np_apply(f, src, *ref, dim)
    got = D.to_numpy()
    assert np.array_equal(got, want)
    if size[0] % 4 == 0:
        check(lib().mf_apply_laplace_rowcode(flags._h, D2._h, S._h))
        got2 = D2.to_numpy()
        assert np.array_equal(got2.view(np.int32), got.view(np.int32)), "row-code SpMV differs from the plane SpMV"
    else:
        assert lib().mf_apply_laplace_rowcode(flags._h, D2._h, S._h) != 0


@pytest.mark.gpu
@pytest.mark.parametrize("size", [(64, 48, 40), (8, 8, 8), (132, 21, 70), (256, 9, 33), (4, 3, 3)])
@pytest.mark.parametrize("zc", [32, 5])
def test_rowcode_spmv_zmarch_matches_flat(size, zc, monkeypatch):
    """the z-marching register-pipeline SpMV (k_cg_spmv_dot_zm, MF_SPMV_ZM=1) against the default 1-D kernel and numpy:
    bit-identical products, on rows shorter / longer than one warp segment and z chunks that do not divide the grid"""
    from mantaflow_b200 import manta
    from mantaflow_b200._lib import lib, check
    rs = np.random.RandomState(13)
    f = _flags(rs, size, 3, True)
    src = (rs.rand(*f.shape).astype(np.float32) - np.float32(0.5))
    outs = []
    for flat in (False, True):
        monkeypatch.setenv("MF_SPMV_ZC", str(zc))
        if flat:
            monkeypatch.delenv("MF_SPMV_ZM", raising=False)
        else:
            monkeypatch.setenv("MF_SPMV_ZM", "1")
        s = manta.Solver(name="s", gridSize=manta.vec3(*size), dim=3)
        flags = s.create(manta.FlagGrid); flags.from_numpy(f)
        S = s.create(manta.RealGrid); D = s.create(manta.RealGrid); S.from_numpy(src)
        check(lib().mf_apply_laplace_rowcode(flags._h, D._h, S._h))
        outs.append(D.to_numpy())
    want = _np_apply(f, src, *_np_laplace(f, 3), 3)
    assert np.array_equal(outs[0].view(np.int32), outs[1].view(np.int32))
    assert np.array_equal(outs[0], want)


@pytest.mark.gpu
@pytest.mark.parametrize("size", [(32, 24, 16), (64, 48, 40)])
def test_half_level1_with_float_rows_matches_float(size, monkeypatch):
    """the level-1 operator in binary16 with the rows binary16 cannot hold (around the pinned cell of zeroPressureFixing) kept in
    float -- NaN-diagonal marker + per-colour vertex lists swept by one extra CTA row of each launch -- against the all-float
    level (MF_MG_HALF_EXACT_ONLY=1): same pressure bit for bit, same iteration count, with the wavefront legs and with the
    per-colour sweeps, on one rank and on emulated z-slab ranks"""
    from mantaflow_b200 import manta
    res = []
    for env in ({"MF_MG_HALF_EXACT_ONLY": "1"}, {}, {"MF_MG_NOWAVE": "1"}, {"MF_MG_UNFUSED": "1"}):
        for k in ("MF_MG_HALF_EXACT_ONLY", "MF_MG_NOWAVE", "MF_MG_UNFUSED"):
            monkeypatch.delenv(k, raising=False)
        for k, v in env.items():
            monkeypatch.setenv(k, v)
        s = manta.Solver(name="s", gridSize=manta.vec3(*size), dim=3)
        flags = s.create(manta.FlagGrid); flags.initDomain(); flags.fillGrid()
        vel = s.create(manta.MACGrid); p = s.create(manta.RealGrid)
        rs2 = np.random.RandomState(9)
        vel.from_numpy(rs2.rand(size[2], size[1], size[0], 3).astype(np.float32) - np.float32(0.5))
        manta.solvePressure(flags=flags, vel=vel, pressure=p, cgAccuracy=1e-6, preconditioner=manta.PcMGStatic, zeroPressureFixing=True)
        res.append((p.to_numpy(), manta.getSolveInfo()["iterations"]))
        fmt = Ct.c_int(0)
        from mantaflow_b200._lib import lib, check
        check(lib().mf_mg_level_format(s._h, 1, Ct.byref(fmt)))
        assert fmt.value == (4 if "MF_MG_HALF_EXACT_ONLY" in env else 2)
        manta.releaseMG(s)
    for r in res[1:]:
        assert r[1] == res[0][1]
        assert np.array_equal(r[0].view(np.int32), res[0][0].view(np.int32))


@pytest.mark.gpu
@pytest.mark.parametrize("nranks", [2, 3])
def test_half_level1_lists_partitioned(nranks, monkeypatch):
    """the same on a z-slab partitioned level 1 (>= 2^20 vertices, emulated ranks): every rank sweeps the listed float-row
    vertices of its own z pairs; against the all-float level on one rank"""
    from mantaflow_b200 import manta
    size = (256, 128, 256)
    rs2 = np.random.RandomState(9)
    v0 = rs2.rand(size[2], size[1], size[0], 3).astype(np.float32) - np.float32(0.5)

    def run(comm):
        s = manta.Solver(name="s", gridSize=manta.vec3(*size), dim=3, comm=comm)
        flags = s.create(manta.FlagGrid); flags.initDomain(); flags.fillGrid()
        vel = s.create(manta.MACGrid); p = s.create(manta.RealGrid)
        vel.from_numpy(v0)
        manta.solvePressure(flags=flags, vel=vel, pressure=p, cgAccuracy=1e-4, preconditioner=manta.PcMGStatic, zeroPressureFixing=True)
        out = (p.to_numpy(), manta.getSolveInfo()["iterations"])
        manta.releaseMG(s)
        return out

    monkeypatch.setenv("MF_MG_HALF_EXACT_ONLY", "1")
    ref = run(None)
    monkeypatch.delenv("MF_MG_HALF_EXACT_ONLY")
    comm = manta.Comm.emulated(nranks)
    try:
        got = run(comm)
    finally:
        comm.destroy()
    assert got[1] == ref[1]
    assert np.array_equal(got[0].view(np.int32), ref[0].view(np.int32))


@pytest.mark.gpu
@pytest.mark.parametrize("size,nranks", [((64, 48, 40), 1), ((128, 96, 80), 1), ((256, 128, 256), 1), ((256, 128, 256), 2)])
def test_fused_dot_beta_matches_separate(size, nranks, monkeypatch):
    """sigmaNew = z . r reduced by the level-0 up leg of the V-cycle while it stores z (k_mg_up0w<true>) against the separate
    k_cg_dot_beta pass (MF_MG_NODOT=1): same iteration count and pressure, one launch less per PCG iteration"""
    from mantaflow_b200 import manta
    from mantaflow_b200._lib import lib
    rs2 = np.random.RandomState(11)
    v0 = rs2.rand(size[2], size[1], size[0], 3).astype(np.float32) - np.float32(0.5)
    lib().mf_solver_launch_count.restype = Ct.c_longlong
    res = []
    for env in ({"MF_MG_NODOT": "1"}, {}):
        monkeypatch.delenv("MF_MG_NODOT", raising=False)
        for k, v in env.items():
            monkeypatch.setenv(k, v)
        comm = manta.Comm.emulated(nranks) if nranks > 1 else None
        try:
            s = manta.Solver(name="s", gridSize=manta.vec3(*size), dim=3, comm=comm)
            flags = s.create(manta.FlagGrid); flags.initDomain(); flags.fillGrid()
            vel = s.create(manta.MACGrid); p = s.create(manta.RealGrid)
            vel.from_numpy(v0)
            n0 = lib().mf_solver_launch_count(s._h)
            manta.solvePressure(flags=flags, vel=vel, pressure=p, cgAccuracy=1e-5, preconditioner=manta.PcMGStatic, zeroPressureFixing=True)
            res.append((p.to_numpy(), manta.getSolveInfo()["iterations"], lib().mf_solver_launch_count(s._h) - n0))
            manta.releaseMG(s)
            del flags, vel, p, s
        finally:
            if comm is not None:
                comm.destroy()
    assert res[0][1] == res[1][1] and res[0][1] >= 3
    assert np.array_equal(res[0][0].view(np.int32), res[1][0].view(np.int32))
    assert res[1][2] < res[0][2], "the fused reduction must save launches"


@pytest.mark.gpu
@pytest.mark.parametrize("nranks", [2, 3])
def test_rowcode_spmv_partitioned(nranks):
    """z-slab partition (emulated ranks): the row-code SpMV reads the neighbour planes of other slabs"""
    from mantaflow_b200 import manta
    from mantaflow_b200._lib import lib, check
    rs = np.random.RandomState(5)
    size, dim = (32, 12, 14), 3
    comm = manta.Comm.emulated(nranks)
    try:
        s = manta.Solver(name="s", gridSize=manta.vec3(*size), dim=dim, comm=comm)
        f = _flags(rs, size, dim, True)
        src = rs.rand(*f.shape).astype(np.float32)
        flags = s.create(manta.FlagGrid); flags.from_numpy(f)
        S = s.create(manta.RealGrid); D = s.create(manta.RealGrid); S.from_numpy(src)
        check(lib().mf_apply_laplace_rowcode(flags._h, D._h, S._h))
        want = _np_apply(f, src, *_np_laplace(f, dim), dim)
        assert np.array_equal(D.to_numpy(), want)
        del flags, S, D, s
    finally:
        comm.destroy()


@pytest.mark.gpu
def test_pressure_solve_rowcode_vs_planes(monkeypatch):
    """the CG with row codes and with coefficient planes converges to the same pressure (same iteration count)"""
    from mantaflow_b200 import manta
    rs = np.random.RandomState(2)
    size = (32, 24, 16)
    res = []
    for planes in (False, True):
        if planes:
            monkeypatch.setenv("MF_CG_PLANES", "1")
        s = manta.Solver(name="s", gridSize=manta.vec3(*size), dim=3)
        flags = s.create(manta.FlagGrid); flags.initDomain(); flags.fillGrid()
        vel = s.create(manta.MACGrid); p = s.create(manta.RealGrid)
        rs2 = np.random.RandomState(9)
        vel.from_numpy(rs2.rand(size[2], size[1], size[0], 3).astype(np.float32) - np.float32(0.5))
        manta.solvePressure(flags=flags, vel=vel, pressure=p, cgAccuracy=1e-6, preconditioner=manta.PcMGStatic, zeroPressureFixing=True)
        res.append((p.to_numpy(), vel.to_numpy()))
        manta.releaseMG(s)
    assert np.allclose(res[0][0], res[1][0], rtol=0, atol=2e-5)
    assert np.allclose(res[0][1], res[1][1], rtol=0, atol=2e-5)


VC_SIZES = [(64, 50, 37), (8, 8, 8), (60, 29, 33), (120, 31, 18), (56, 26, 64), (116, 60, 41), (4, 70, 9), (128, 96, 80)]


@pytest.mark.gpu
@pytest.mark.parametrize("size", VC_SIZES)
def test_vcycle_wavefront_matches_unfused(size, monkeypatch):
    """one V(1,1)-cycle: the z-marching wavefront legs (mg_wave0.cuh), the tiled fused legs (mg_fused0.cuh) and the plain
    per-colour sweeps give bit-identical results; so does the binary16 copy of the level-1 operator (exact by construction)
    and so do the persistent small-level kernels (mg_tail.cuh, MF_MG_TAIL=1: grid barrier, MF_MG_TAIL=cluster: one cluster)"""
    from mantaflow_b200 import manta
    from mantaflow_b200._lib import lib, check
    rs = np.random.RandomState(21)
    f = _flags(rs, size, 3)
    f[f == EMPTY] = FLUID                 # closed box with obstacles
    nz, ny, nx = f.shape
    f[nz // 2, ny // 2, nx // 2] = EMPTY  # one Dirichlet neighbour keeps the operator regular
    rhs = (rs.rand(*f.shape).astype(np.float32) - np.float32(0.5))
    outs = []
    fmt = []
    for env in ({}, {"MF_MG_NOWAVE": "1"}, {"MF_MG_UNFUSED": "1"}, {"MF_MG_NOHALF": "1"}, {"MF_MG_NOFOLD": "1"}, {"MF_MG_TAIL": "1"}, {"MF_MG_TAIL": "cluster"}):
        for k in ("MF_MG_NOWAVE", "MF_MG_UNFUSED", "MF_MG_NOHALF", "MF_MG_NOFOLD", "MF_MG_TAIL"):
            monkeypatch.delenv(k, raising=False)
        for k, v in env.items():
            monkeypatch.setenv(k, v)
        s = manta.Solver(name="s", gridSize=manta.vec3(*size), dim=3)
        flags = s.create(manta.FlagGrid); flags.from_numpy(f)
        A0, Ai, Aj, Ak, R, D = [s.create(manta.RealGrid) for _ in range(6)]
        R.from_numpy(rhs)
        check(lib().mf_make_laplace_matrix(flags._h, A0._h, Ai._h, Aj._h, Ak._h, None))
        check(lib().mf_mg_vcycle(s._h, A0._h, Ai._h, Aj._h, Ak._h, R._h, D._h, 1e-7))
        outs.append(D.to_numpy())
        b = Ct.c_int(0); nl = Ct.c_int(0); fb = Ct.c_int(0)
        check(lib().mf_mg_info(s._h, Ct.byref(nl), Ct.byref(fb)))
        if nl.value > 1:
            check(lib().mf_mg_level_format(s._h, 1, Ct.byref(b)))
        fmt.append(b.value)
        manta.releaseMG(s)
    assert fmt in ([2, 2, 2, 4, 2, 2, 2], [0, 0, 0, 0, 0, 0, 0]), "level-1 operator of a flags-only Laplacian must be stored in binary16"
    assert np.array_equal(outs[3].view(np.int32), outs[2].view(np.int32)), "binary16 level-1 operator changes the V-cycle"
    assert np.array_equal(outs[4].view(np.int32), outs[0].view(np.int32)), "folding the first colour-0 sweep into the restriction changes the V-cycle"
    assert np.array_equal(outs[5].view(np.int32), outs[0].view(np.int32)), "persistent small-level kernels change the V-cycle"
    assert np.array_equal(outs[6].view(np.int32), outs[0].view(np.int32)), "small-level kernels as one thread-block cluster change the V-cycle"
    assert np.isfinite(outs[0]).all() and np.abs(outs[2]).max() > 0
    assert np.array_equal(outs[1].view(np.int32), outs[2].view(np.int32)), "tiled fused legs differ from the plain sweeps"
    assert np.array_equal(outs[0].view(np.int32), outs[2].view(np.int32)), "wavefront legs differ from the plain sweeps"


@pytest.mark.gpu
@pytest.mark.parametrize("size", [(64, 50, 37), (128, 96, 80), (120, 31, 18), (256, 256, 256)])
def test_vcycle_bottom_matches_unfused(size, monkeypatch):
    """the last swept level + the coarsest-level CG in one CTA out of shared memory (mg_bottom.cuh, opt-in MF_MG_BOTTOM=1) against
    the per-phase launches: bit-identical V-cycle, 18 launches fewer (7 + 8 colour sweeps, residual, restriction, interpolation, CG -> 1)"""
    from mantaflow_b200 import manta
    from mantaflow_b200._lib import lib, check
    rs = np.random.RandomState(5)
    f = _flags(rs, size, 3)
    f[f == EMPTY] = FLUID
    nz, ny, nx = f.shape
    f[nz // 2, ny // 2, nx // 2] = EMPTY
    rhs = (rs.rand(*f.shape).astype(np.float32) - np.float32(0.5))
    outs, launches = [], []
    lib().mf_solver_launch_count.restype = Ct.c_longlong
    for env in ({"MF_MG_BOTTOM": "1"}, {}):
        monkeypatch.delenv("MF_MG_BOTTOM", raising=False)
        for k, v in env.items():
            monkeypatch.setenv(k, v)
        s = manta.Solver(name="s", gridSize=manta.vec3(*size), dim=3)
        flags = s.create(manta.FlagGrid); flags.from_numpy(f)
        A0, Ai, Aj, Ak, R, D = [s.create(manta.RealGrid) for _ in range(6)]
        R.from_numpy(rhs)
        check(lib().mf_make_laplace_matrix(flags._h, A0._h, Ai._h, Aj._h, Ak._h, None))
        check(lib().mf_mg_vcycle(s._h, A0._h, Ai._h, Aj._h, Ak._h, R._h, D._h, 1e-7))     # builds the hierarchy
        n0 = lib().mf_solver_launch_count(s._h)
        check(lib().mf_mg_vcycle(s._h, None, None, None, None, R._h, D._h, 1e-7))          # re-applies it: launches of one V-cycle
        launches.append(lib().mf_solver_launch_count(s._h) - n0)
        outs.append(D.to_numpy())
        manta.releaseMG(s)
    assert np.isfinite(outs[0]).all() and np.abs(outs[0]).max() > 0
    assert np.array_equal(outs[0].view(np.int32), outs[1].view(np.int32)), "one-CTA bottom of the V-cycle differs from the per-phase launches"
    assert launches[1] - launches[0] == 18, launches


@pytest.mark.gpu
@pytest.mark.parametrize("nranks", [1, 2, 3])
def test_vcycle_wavefront_partitioned(nranks, monkeypatch):
    """a grid large enough for a partitioned level 1 (>= 2^20 vertices): wavefront legs on z-slabs (emulated ranks) against the
    plain sweeps on one rank"""
    from mantaflow_b200 import manta
    from mantaflow_b200._lib import lib, check
    size = (256, 128, 256)
    rs = np.random.RandomState(4)
    f = _flags(rs, size, 3)
    f[f == EMPTY] = FLUID
    f[128, 64, 128] = EMPTY
    rhs = (rs.rand(*f.shape).astype(np.float32) - np.float32(0.5))

    def run(comm):
        s = manta.Solver(name="s", gridSize=manta.vec3(*size), dim=3, comm=comm)
        flags = s.create(manta.FlagGrid); flags.from_numpy(f)
        A0, Ai, Aj, Ak, R, D = [s.create(manta.RealGrid) for _ in range(6)]
        R.from_numpy(rhs)
        check(lib().mf_make_laplace_matrix(flags._h, A0._h, Ai._h, Aj._h, Ak._h, None))
        check(lib().mf_mg_vcycle(s._h, A0._h, Ai._h, Aj._h, Ak._h, R._h, D._h, 1e-7))
        out = D.to_numpy()
        manta.releaseMG(s)
        return out

    monkeypatch.setenv("MF_MG_UNFUSED", "1")
    ref = run(None)
    monkeypatch.delenv("MF_MG_UNFUSED")
    comm = manta.Comm.emulated(nranks) if nranks > 1 else None
    try:
        got = run(comm)
    finally:
        if comm is not None:
            comm.destroy()
    assert np.array_equal(got.view(np.int32), ref.view(np.int32))


@pytest.mark.gpu
@pytest.mark.parametrize("gridtype", ["RealGrid", "MACGrid"])
@pytest.mark.parametrize("order,clampMode,strength", [(1, 2, 1.0), (2, 2, 1.0), (2, 1, 1.0), (2, 2, 0.8), (2, 1, 2.0), (2, 2, 0.25)])
def test_advect_fast_matches_generic(gridtype, order, clampMode, strength, monkeypatch):
    """the 32-bit-index advection kernels (advect_fast.cuh: 3-D, linear interpolation, first-order trace) against the generic
    ones (MF_ADVECT_GENERIC=1) on a grid with obstacles, empty cells and velocities of up to ~6 cells per step (so that traces
    leave the grid and hit every edge case of BUILD_INDEX and of the clamp lookups): bit-identical.  strength 1 / 2 / 0.25 take the
    float form of the MacCormack correction (mcCorrect: strength * 0.5 is a power of two), 0.8 the reference's double expression;
    the data carries zeros, denormals and values 2^-30 .. 2^-120 apart in magnitude to hit the roundings that could tell them apart"""
    from mantaflow_b200 import manta
    size = (70, 45, 38)
    rs = np.random.RandomState(31)
    f = _flags(rs, size, 3)
    vel0 = ((rs.rand(size[2], size[1], size[0], 3).astype(np.float32) - np.float32(0.5)) * np.float32(12.0))
    vel0[rs.rand(size[2], size[1], size[0]) < 0.5] *= np.float32(0.05)
    d0 = rs.rand(size[2], size[1], size[0]).astype(np.float32)
    # magnitudes spread over the float range, exact zeros and denormals
    sel = rs.rand(size[2], size[1], size[0])
    scale = np.where(sel < 0.1, 0.0, np.where(sel < 0.2, 1e-40, np.where(sel < 0.3, 1e-33, np.where(sel < 0.4, 1e-12, np.where(sel < 0.45, 3e-38, 1.0))))).astype(np.float32)
    d0 = d0 * scale
    vsel = rs.rand(size[2], size[1], size[0], 3)
    vel0 = vel0 * np.where(vsel < 0.05, 0.0, np.where(vsel < 0.1, 1e-40, np.where(vsel < 0.15, 1e-20, 1.0))).astype(np.float32)
    outs = []
    for generic in (False, True):
        monkeypatch.delenv("MF_ADVECT_GENERIC", raising=False)
        if generic:
            monkeypatch.setenv("MF_ADVECT_GENERIC", "1")
        s = manta.Solver(name="s", gridSize=manta.vec3(*size), dim=3)
        flags = s.create(manta.FlagGrid); flags.from_numpy(f)
        vel = s.create(manta.MACGrid); vel.from_numpy(vel0)
        if gridtype == "RealGrid":
            g = s.create(manta.RealGrid); g.from_numpy(d0)
        else:
            g = vel
        manta.advectSemiLagrange(flags=flags, vel=vel, grid=g, order=order, clampMode=clampMode, strength=strength)
        outs.append(g.to_numpy())
    assert np.isfinite(outs[0]).all()
    assert np.array_equal(outs[0].view(np.int32), outs[1].view(np.int32))


@pytest.mark.gpu
@pytest.mark.parametrize("size,amp,zc", [((64, 48, 40), 1.5, 32), ((72, 45, 38), 3.0, 16), ((36, 20, 9), 12.0, 4), ((128, 96, 70), 2.0, 32)])
@pytest.mark.parametrize("order,clampMode", [(1, 2), (2, 2), (2, 1)])
def test_advect_tma_matches_ldg(size, amp, zc, order, clampMode, monkeypatch):
    """the TMA-staged MAC advection (advect_tma.cuh: cp.async.bulk.tensor.3d boxes with halos in a shared-memory ring, stencils and
    in-window gathers from shared memory, global fallback for traces that leave the window) against the per-thread-LDG kernels of
    advect_fast.cuh: bit-identical, on ragged grids (tiles and z chunks that hang over the grid), with obstacles and empty cells,
    with velocities inside the window (|v| dt < 2) and far outside it (fallback gathers, traces that leave the grid)"""
    from mantaflow_b200 import manta
    rs = np.random.RandomState(47)
    f = _flags(rs, size, 3)
    vel0 = ((rs.rand(size[2], size[1], size[0], 3).astype(np.float32) - np.float32(0.5)) * np.float32(2.0 * amp))
    vel0[rs.rand(size[2], size[1], size[0]) < 0.3] *= np.float32(0.05)
    outs = []
    for tma in (False, True):
        monkeypatch.setenv("MF_ADVECT_TMA", "1" if tma else "0")
        monkeypatch.setenv("MF_TMA_ZC", str(zc))
        s = manta.Solver(name="s", gridSize=manta.vec3(*size), dim=3)
        flags = s.create(manta.FlagGrid); flags.from_numpy(f)
        vel = s.create(manta.MACGrid); vel.from_numpy(vel0)
        l0 = s.launchCount()
        manta.advectSemiLagrange(flags=flags, vel=vel, grid=vel, order=order, clampMode=clampMode)
        outs.append(vel.to_numpy())
    assert np.isfinite(outs[0]).all()
    assert np.array_equal(outs[0].view(np.int32), outs[1].view(np.int32))


@pytest.mark.gpu
@pytest.mark.parametrize("pc", ["none", "mic", "mg"])
@pytest.mark.parametrize("dim", [3, 2])
def test_gridcg_object_matches_solve_pressure(pc, dim, monkeypatch):
    """GridCg driven as an object, the way the reference's C++ plugins do (solvePressureSystem, pressure.cpp:372-452: matrix,
    GridCg(dst, rhs, residual, search, flags, tmp, A0..Ak), setAccuracy, setUseL2Norm(false), set*Preconditioner, iterate until
    it returns false), must produce the pressure of the solvePressureSystem plugin bit for bit, with the same iteration count;
    then the accessors, forceReinit and solve(maxIter)."""
    from mantaflow_b200 import manta
    from mantaflow_b200._lib import lib, check
    monkeypatch.setattr(manta, "allowPcNone", True)    # the plugin refuses precondition=False like the reference (pressure.cpp:417)
    monkeypatch.setenv("MF_CG_PLANES", "1")      # the plugin's CG reads the float coefficient planes too (same kernels, same reduction grids)
    size = (28, 22, 18) if dim == 3 else (40, 30, 1)
    rs = np.random.RandomState(17)
    f = _flags(rs, size, dim)
    f[f == EMPTY] = FLUID
    nz, ny, nx = f.shape
    f[nz // 2, ny // 2, nx // 2] = EMPTY
    vel0 = (rs.rand(nz, ny, nx, 3).astype(np.float32) - np.float32(0.5))
    if dim == 2:
        vel0[..., 2] = 0
    s = manta.Solver(name="s", gridSize=manta.vec3(*size), dim=dim)
    flags = s.create(manta.FlagGrid); flags.from_numpy(f)
    vel = s.create(manta.MACGrid); vel.from_numpy(vel0)
    p_ref = s.create(manta.RealGrid); rhs = s.create(manta.RealGrid)
    pcs = {"none": dict(precondition=False), "mic": dict(preconditioner=manta.PcMIC), "mg": dict(preconditioner=manta.PcMGDynamic)}[pc]
    manta.computePressureRhs(rhs=rhs, vel=vel, pressure=p_ref, flags=flags)
    manta.solvePressureSystem(rhs=rhs, vel=vel, pressure=p_ref, flags=flags, cgAccuracy=1e-5, **pcs)
    info = dict(manta.getSolveInfo())
    manta.releaseMG(s)
    # the same through the object
    dst, residual, search, tmp, A0, Ai, Aj, Ak = [s.create(manta.RealGrid) for _ in range(8)]
    check(lib().mf_make_laplace_matrix(flags._h, A0._h, Ai._h, Aj._h, Ak._h, None))
    cg = manta.GridCg(dst, rhs, residual, search, flags, tmp, A0, Ai, Aj, Ak)
    assert cg.getAccuracy() == np.float32(1e-6) and cg.getIterations() == 0
    cg.setAccuracy(1e-5)
    cg.setUseL2Norm(False)
    if pc == "mic":
        cg.setICPreconditioner(manta.GridCg.PC_mICP, A0, Ai, Aj, Ak)
    elif pc == "mg":
        cg.setMGPreconditioner(manta.GridCg.PC_MGP, None, rebuild=True)
    it = 0
    while it < info["maxIter"]:
        it += 1
        if not cg.iterate():
            break
    assert cg.getIterations() == it == info["iterations"]
    assert np.array_equal(dst.to_numpy().view(np.int32), p_ref.to_numpy().view(np.int32))
    assert abs(cg.getResNorm() - info["resNorm"]) <= 1e-12
    if cg.getResNorm() < 1e-5:                         # (the unpreconditioned solve may stop at maxIter instead)
        assert cg.getSigma() == cg.getResNorm()        # mSigma = mResNorm on convergence (conjugategrad.cpp:276)
    else:
        assert it == info["maxIter"]
    if pc == "mg":
        assert cg.getResNorm() < 1e-5
    # a second solve with the same object: forceReinit + solve(maxIter) reproduces it
    first = dst.to_numpy()
    cg.forceReinit()
    cg.solve(info["maxIter"])
    assert cg.getIterations() == info["iterations"]
    assert np.array_equal(dst.to_numpy().view(np.int32), first.view(np.int32))
    with pytest.raises(manta.MantaError):
        cg.setICPreconditioner(manta.GridCg.PC_MGP)
    del cg
    manta.releaseMG(s)


@pytest.mark.gpu
def test_rowcode_cache_follows_flag_changes(monkeypatch):
    """solvePressureSystem keeps the CG's Laplace row codes (and the pinned cell) while the flag grid is unchanged; every way of
    writing flags must drop them: the solves with the cache are bit-identical to solves with MF_CG_NOCODECACHE=1"""
    from mantaflow_b200 import manta
    size = (32, 24, 20)
    rs = np.random.RandomState(3)
    vel0 = (rs.rand(size[2], size[1], size[0], 3).astype(np.float32) - np.float32(0.5))

    def run():
        s = manta.Solver(name="s", gridSize=manta.vec3(*size), dim=3)
        flags = s.create(manta.FlagGrid); vel = s.create(manta.MACGrid); p = s.create(manta.RealGrid); other = s.create(manta.FlagGrid)
        flags.initDomain(); flags.fillGrid()
        out = []

        def solve():
            vel.from_numpy(vel0)
            manta.setWallBcs(flags=flags, vel=vel)
            manta.solvePressure(flags=flags, vel=vel, pressure=p, preconditioner=manta.PcMGStatic, zeroPressureFixing=True, cgAccuracy=1e-5)
            out.append((p.to_numpy(), vel.to_numpy(), manta.getSolveInfo()["iterations"]))
        solve(); solve()                                   # second solve: cached codes
        manta.Box(parent=s, p0=manta.vec3(8, 6, 5), p1=manta.vec3(14, 12, 11)).applyToGrid(grid=flags, value=manta.FlagObstacle)
        manta.releaseMG(s); solve(); solve()               # shape written into the flags
        f = flags.to_numpy(); f[10:14, 14:18, 20:26] = manta.FlagObstacle; flags.from_numpy(f)
        manta.releaseMG(s); solve()                        # upload
        other.initDomain(); other.fillGrid(); flags.swap(other)
        manta.releaseMG(s); solve()                        # swap
        flags.copyFrom(other)
        manta.releaseMG(s); solve(); solve()               # copy
        return out

    a = run()
    monkeypatch.setenv("MF_CG_NOCODECACHE", "1")
    b = run()
    assert len(a) == len(b) == 8
    for (pa, va, ia), (pb, vb, ib) in zip(a, b):
        assert ia == ib and np.array_equal(pa.view(np.int32), pb.view(np.int32)) and np.array_equal(va.view(np.int32), vb.view(np.int32))
    assert not np.array_equal(a[1][0], a[2][0]) and not np.array_equal(a[3][0], a[4][0]) and not np.array_equal(a[4][0], a[5][0])


@pytest.mark.gpu
@pytest.mark.parametrize("gravity", [(0.0, -6e-3, 0.0), (0.0, 0.0, 0.0), (-0.0, 2e-3, 0.0), (1e-3, -2e-3, 3e-3)])
def test_add_buoyancy_bits(gravity):
    """KnAddBuoyancy (extforces.cpp:73-81) bit for bit against the reference's expression evaluated in numpy (double product and
    sum, rounded once), with the inputs that tell the kernel's shortcuts apart: negative zeros in the velocity, an infinite and a
    NaN factor, zero strength components of either sign (float path) and non-zero ones (double path)"""
    from mantaflow_b200 import manta
    rs = np.random.RandomState(5)
    size = (24, 10, 12)
    f = _flags(rs, size, 3)
    nz, ny, nx = f.shape
    dens = rs.rand(nz, ny, nx).astype(np.float32)
    dens[3, 4, 5] = np.inf; dens[6, 5, 9] = np.nan; dens[7, 3, 11] = -dens[7, 3, 11]
    vel0 = (rs.rand(nz, ny, nx, 3).astype(np.float32) - np.float32(0.5))
    vel0[rs.rand(nz, ny, nx, 3) < 0.3] = np.float32(-0.0)
    s = manta.Solver(name="s", gridSize=manta.vec3(*size), dim=3)
    flags = s.create(manta.FlagGrid); flags.from_numpy(f)
    density = s.create(manta.RealGrid); density.from_numpy(dens)
    vel = s.create(manta.MACGrid); vel.from_numpy(vel0)
    manta.addBuoyancy(flags=flags, density=density, vel=vel, gravity=manta.vec3(*gravity))
    got = vel.to_numpy()
    dt, dx = np.float32(s.timestep), np.float32(1.0 / max(size))
    want = vel0.copy()
    fluid = (f & FLUID) != 0
    for c, sh in ((0, (0, 0, 1)), (1, (0, 1, 0)), (2, (1, 0, 0))):
        strength = np.float32(np.float32(np.float32(-np.float32(gravity[c])) * dt) / dx) * np.float32(1.0)   # -gravity * (dt / dx) * coefficient, extforces.cpp:86-88
        lo = tuple(slice(1 - d, (n - 1) - d) for d, n in zip(sh, (nz, ny, nx)))
        hi = (slice(1, nz - 1), slice(1, ny - 1), slice(1, nx - 1))
        m = fluid[hi] & fluid[lo]
        with np.errstate(invalid="ignore", over="ignore"):
            fs = (dens[hi] + dens[lo]).astype(np.float32)
            upd = (want[hi + (c,)].astype(np.float64) + (0.5 * np.float64(strength)) * fs.astype(np.float64)).astype(np.float32)
        w = want[hi + (c,)]
        w[m] = upd[m]
    nan = np.isnan(want)
    assert np.array_equal(np.isnan(got), nan)
    assert np.array_equal(got.view(np.int32)[~nan], want.view(np.int32)[~nan])
