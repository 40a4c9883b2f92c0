"""CPU-side checks of the host logic that mirrors the reference's Python surface (no GPU needed):
vec3 arithmetic in float32, module constants of source/python/defines.py, keyword checking (PbArgs::check),
FluidSolver::step / adaptTimestep bookkeeping, and consistency of the golden fixtures with tests/cases.py."""
import os

import numpy as np
import pytest

import cases as C
import scene_runner as sr
from mantaflow_b200 import manta as M
from mantaflow_b200 import plugins as P

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_vec3_is_float32():
    v = M.vec3(0.1, 0.2, 0.3) * 3 + M.vec3(1)
    assert isinstance(v.x, np.float32)
    assert v.x == np.float32(np.float32(0.1) * np.float32(3) + np.float32(1))
    assert M.vec3(2, 3).z == 0 and M.vec3(5).y == 5
    assert (M.vec3(64, 96, 64) * M.vec3(0.5, 0.1, 0.5)).y == np.float32(96) * np.float32(0.1)
    assert M.vec3([1, 2, 3]) == M.vec3(1, 2, 3) and -M.vec3(1, 2, 3) == M.vec3(-1, -2, -3)


def test_module_constants():
    """source/python/defines.py:24-50, source/grid.h:306-320, source/plugin/pressure.cpp:27"""
    assert (M.FlagFluid, M.FlagObstacle, M.FlagEmpty, M.FlagInflow, M.FlagOutflow, M.FlagOpen, M.FlagStick) == (1, 2, 4, 8, 16, 32, 64)
    assert (M.PcNone, M.PcMIC, M.PcMGDynamic, M.PcMGStatic) == (0, 1, 2, 3)
    assert (M.IntEuler, M.IntRK2, M.IntRK4) == (0, 1, 2)
    assert M.GUI is False and M.CUDA is True and M.DOUBLEPRECISION is False


def test_unknown_keyword_is_an_error():
    """PbArgs::check (source/pwrapper/pconvert.cpp:460-474)"""
    with pytest.raises(M.MantaError, match="Argument 'foo' unknown"):
        M._checkArgs("advectSemiLagrange", {"foo": 1}, ())
    M._checkArgs("advectSemiLagrange", {"notiming": True}, ())
    with pytest.raises(M.MantaError, match="Argument 'bogus' unknown"):
        P._params("solvePressure", {"bogus": 1})
    with pytest.raises(M.MantaError, match="Argument 'obvel' unknown"):
        P._params("solvePressureSystem", {"obvel": None}, allowObvel=False)
    with pytest.raises(M.MantaError, match="Argument 'retRhs' unknown"):
        P._params("correctVelocity", {"retRhs": None})


def test_missing_and_mistyped_arguments():
    with pytest.raises(M.MantaError, match="Argument 'flags' is not defined"):
        M.advectSemiLagrange(vel=None, grid=None)
    with pytest.raises(M.MantaError, match="can't convert argument 'flags'"):
        M.setWallBcs(flags=3, vel=None)
    with pytest.raises(M.MantaError, match="Argument 'gridSize' is not defined"):
        M.Solver(name="x")
    with pytest.raises(M.MantaError, match="Only 2D and 3D"):
        M.Solver(name="x", gridSize=M.vec3(4, 4, 4), dim=4)
    with pytest.raises(M.MantaError, match="size.z != 1"):
        M.Solver(name="x", gridSize=M.vec3(4, 4, 4), dim=2)


def test_pcnone_throws_like_the_reference():
    """solvePressure(preconditioner=PcNone) asserts in setICPreconditioner (conjugategrad.cpp:312 via pressure.cpp:417)"""
    with pytest.raises(M.MantaError, match="setICPreconditioner"):
        P._params("solvePressure", {"preconditioner": M.PcNone})
    with pytest.raises(M.MantaError, match="setICPreconditioner"):
        P._params("solvePressure", {"precondition": False})
    p, _ = P._params("solvePressure", {"preconditioner": M.PcMGStatic, "cgAccuracy": 1e-6, "zeroPressureFixing": True})
    assert p.preconditioner == 3 and p.zeroPressureFixing == 1 and np.float32(p.cgAccuracy) == np.float32(1e-6)


class _FakeSolver(M.FluidSolver):
    """time bookkeeping only: bypass the device handle"""
    def __init__(self):
        self._h = None
        self._dt = np.float32(1.0); self.timeTotal = np.float32(0); self.frame = 0; self._count = 0
        self.cfl = np.float32(1000); self.timestepMin = np.float32(1); self.timestepMax = np.float32(1)
        self.frameLength = np.float32(1); self.timePerFrame = np.float32(0); self._lockDt = False

    timestep = property(lambda self: self._dt, lambda self, v: setattr(self, "_dt", np.float32(v)))


def test_solver_step_counts_frames():
    """FluidSolver::step, source/fluidsolver.cpp:143-166"""
    s = _FakeSolver()
    for _ in range(3):
        s.step()
    assert s.frame == 3 and s.timeTotal == 3.0 and s.timePerFrame == 0.0
    s.timestep = 0.25; s.timestepMin = np.float32(0.1)
    s.step(); s.step()
    assert s.frame == 3 and abs(float(s.timePerFrame) - 0.5) < 1e-6
    s.step(); s.step()
    assert s.frame == 4 and s.timeTotal == 4.0
    s.step(frame=17)
    assert s.frame == 17


def test_adapt_timestep():
    """FluidSolver::adaptTimestep, source/fluidsolver.cpp:184-205"""
    s = _FakeSolver()
    s.cfl = np.float32(2.0); s.timestepMin = np.float32(0.2); s.timestepMax = np.float32(2.0); s.frameLength = np.float32(4.0)
    s.adaptTimestep(4.0)        # dt * cfl / (maxVel*dt) = 0.5
    assert abs(float(s.timestep) - 0.5) < 1e-4
    s.adaptTimestep(0.001)      # clamped to timestepMax
    assert float(s.timestep) == 2.0


@pytest.mark.parametrize("name", list(C.CASES))
def test_golden_fixture_matches_case(name):
    """every case has a golden file made from exactly the inputs the case regenerates, with the declared shapes"""
    case = C.CASES[name]
    path = os.path.join(GOLDEN, name + ".npz")
    assert os.path.exists(path), "missing golden fixture (tests/golden/make_golden.py)"
    gold = np.load(path)
    assert int(gold["in_crc"][0]) == C.checksum(C.inputs_for(case))
    for g in case["save"]:
        a = gold["out_" + g]
        if sr.is_particle_type(case["grids"][g]):
            assert a.ndim == 2 and a.shape[1] == 3 and a.shape[0] > 0 and a.dtype == np.float32 and np.isfinite(a).all()
            continue
        assert a.shape == sr.grid_shape(case, case["grids"][g])
        assert a.dtype == (np.int32 if case["grids"][g] in ("FlagGrid", "IntGrid") else np.float32)
        if a.dtype != np.int32:
            assert np.isfinite(a).all()


def test_flag_fixture_semantics():
    """initDomain(boundaryWidth=1)+fillGrid+Box obstacle from the reference: bit patterns of source/grid.h:306-320"""
    f = np.load(os.path.join(GOLDEN, "flags_domain.npz"))["out_flags"]
    assert set(np.unique(f)) <= {1, 2}
    assert (f[0] == 2).all() and (f[:, :2, :] == 2).all() and (f[:, :, -2:] == 2).all()
    assert f[6, 8, 7] == 2          # inside Box p0=(5,6,4) p1=(9,11,8)
    assert f[3, 3, 3] == 1
    fo = np.load(os.path.join(GOLDEN, "flags_open.npz"))["out_flags"]
    assert {int(v) for v in np.unique(fo)} >= {1, 20}   # fluid, outflow|empty
