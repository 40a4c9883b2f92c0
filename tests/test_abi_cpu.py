"""CPU-side checks (no GPU needed): the C-ABI library loads and exports every symbol include/mantacuda.h declares,
the ctypes table matches the header, and the product path fails loudly without a CUDA device (no CPU fallback)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "mantacuda.h")


def _declared():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(mf_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_the_path():
    names = _declared()
    for must in ("mf_advect_semi_lagrange", "mf_set_wall_bcs", "mf_add_buoyancy", "mf_compute_pressure_rhs",
                 "mf_solve_pressure_system", "mf_correct_velocity", "mf_solve_pressure", "mf_release_mg",
                 "mf_grid_upload", "mf_grid_download", "mf_solver_create"):
        assert must in names


def test_library_exports_every_declared_symbol():
    from mantaflow_b200 import _lib
    assert os.path.exists(_lib.LIB_PATH), "libmantacuda.so not built: run __graft_entry__.build()"
    l = C.CDLL(_lib.LIB_PATH)
    missing = [n for n in _declared() if not hasattr(l, n)]
    assert not missing, "declared in mantacuda.h but not exported: %s" % missing


def test_ctypes_table_matches_header():
    from mantaflow_b200 import _lib
    declared = set(_declared())
    bound = set(_lib.SIGNATURES)
    assert declared == bound, (sorted(declared - bound), sorted(bound - declared))
    _lib.lib()   # resolves every symbol with its signature


def test_struct_layouts_match_header():
    """field order of mf_pressure_params / mf_solve_info in the header == ctypes Structures"""
    from mantaflow_b200 import _lib
    src = open(HEADER).read()
    body = re.search(r"typedef struct mf_pressure_params \{(.*?)\} mf_pressure_params;", src, re.S).group(1)
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    fields = [re.findall(r"(\w+)\s*;", ln)[0] for ln in body.splitlines() if ";" in ln]
    assert fields == [f[0] for f in _lib.PressureParams._fields_]
    body = re.search(r"typedef struct mf_solve_info \{(.*?)\} mf_solve_info;", src, re.S).group(1)
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    fields = [re.findall(r"(\w+)\s*;", ln)[0] for ln in body.splitlines() if ";" in ln]
    assert fields == [f[0] for f in _lib.SolveInfo._fields_]


def test_pressure_defaults_are_the_reference_defaults():
    """solvePressure keyword defaults, source/plugin/pressure.cpp:482-497"""
    from mantaflow_b200 import _lib
    p = _lib.PressureParams()
    _lib.lib().mf_pressure_params_default(C.byref(p))
    assert np.float32(p.cgAccuracy) == np.float32(1e-3) and np.float32(p.gfClamp) == np.float32(1e-4)
    assert np.float32(p.cgMaxIterFac) == np.float32(1.5) and p.precondition == 1 and p.preconditioner == 1
    assert (p.enforceCompatibility, p.useL2Norm, p.zeroPressureFixing) == (0, 0, 0)
    assert not p.phi and not p.fractions and not p.retRhs and p.surfTens == 0.0


def test_no_cpu_fallback():
    """without a CUDA device the solver must refuse to exist (on a GPU box this test checks the device count instead)"""
    from mantaflow_b200 import manta as M
    from mantaflow_b200 import _lib
    if _lib.lib().mf_device_count() > 0:
        pytest.skip("CUDA device present")
    with pytest.raises(M.MantaError, match="no CUDA device"):
        M.Solver(name="main", gridSize=M.vec3(8, 8, 8), dim=3)


def test_null_handles_return_errors_not_crashes():
    from mantaflow_b200 import _lib
    l = _lib.lib()
    assert l.mf_grid_clear(None) != 0
    assert b"null" in l.mf_last_error()
    assert l.mf_advect_semi_lagrange(None, None, None, 1, 1.0, 1, 2, 1) != 0
    assert l.mf_solve_pressure(None, None, None, None, None) != 0
    assert l.mf_solver_sync(None) != 0
    l.mf_solver_destroy(None); l.mf_grid_destroy(None)   # no-ops
    assert l.mf_version() >= 100
    assert b"sm_100a" in l.mf_build_info()
