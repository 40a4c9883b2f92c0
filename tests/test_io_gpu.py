"""`.uni` interoperability with the reference (fileio/iogrids.cpp:38-44, 385-505): a file the unmodified reference
wrote loads bit for bit into a device grid, and a file written from a device grid is read back by the reference; plus the numpy
bridge plugin names of source/plugin/numpyconvert.cpp:145-222."""
import gzip
import os
import subprocess
import tempfile

import numpy as np
import pytest

import cases as C
import scene_runner as sr

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
MANTA_REF = os.path.join(ROOT, "oracle", "_ref", "manta")
GRIDS = {"density": "RealGrid", "vel": "MACGrid", "flags": "FlagGrid", "phi": "LevelsetGrid", "vec": "VecGrid", "cnt": "IntGrid"}


def _ref(d, size, env_extra=None):
    env = dict(os.environ, MF_CASE_DIR=d, MF_UNI_SIZE="%d,%d,%d" % tuple(size), MANTA_DISABLE_UI="1", OMP_NUM_THREADS="1",
               MF_UNI_GRIDS=",".join("%s:%s" % kv for kv in GRIDS.items()))
    p = subprocess.run([MANTA_REF, os.path.join(HERE, "golden", "uni_driver.py")], env=env, cwd=d, capture_output=True, text=True, timeout=600)
    assert p.returncode == 0 and "Error" not in p.stdout and "Traceback" not in p.stderr, p.stdout[-2000:] + p.stderr[-2000:]


def _inputs(size):
    nx, ny, nz = size
    r = np.random.RandomState(9)
    return {"density": C.gen_real(size, 1, -2.0, 2.0), "vel": C.gen_vel(size, 2, 1.0, 3), "flags": C.gen_flags_liquid(size, 0.5, 3 if nz > 1 else 2),
            "phi": C.gen_phi_sphere(size, nx * 0.4, ny * 0.5, nz * 0.5, 3.0), "vec": C.gen_vec(size, 3),
            "cnt": r.randint(-5, 1000, size=(nz, ny, nx)).astype(np.int32)}


@pytest.mark.gpu
@pytest.mark.parametrize("size", [C.S3, C.S2])
def test_uni_files_interoperate_with_the_reference(size):
    if not os.path.exists(MANTA_REF):
        pytest.skip("oracle/_ref/manta not built")
    from mantaflow_b200 import manta as M
    ins = _inputs(size)
    s = M.Solver(name="main", gridSize=M.vec3(*size), dim=3 if size[2] > 1 else 2)
    with tempfile.TemporaryDirectory() as d:
        grids = {}
        for name, gt in GRIDS.items():
            sr.write_raw(os.path.join(d, "in_%s.raw" % name), ins[name])
            grids[name] = s.create(getattr(M, gt))
            grids[name].from_numpy(ins[name])
            grids[name].save(os.path.join(d, "our_%s.uni" % name))
        _ref(d, size)
        for name, gt in GRIDS.items():
            # (1) reference-written file -> device grid
            g = s.create(getattr(M, gt))
            g.load(os.path.join(d, "ref_%s.uni" % name))
            assert np.array_equal(g.to_numpy(), ins[name]), name
            # (2) our file, as read by the reference
            back = sr.read_raw(os.path.join(d, "back_%s.raw" % name), ins[name].dtype, ins[name].shape)
            assert np.array_equal(back, ins[name]), name
            # (3) same header layout: magic + 288-byte UniHeader, fields identical except info string and timestamp
            a = gzip.open(os.path.join(d, "ref_%s.uni" % name), "rb").read(4 + 288)
            b = gzip.open(os.path.join(d, "our_%s.uni" % name), "rb").read(4 + 288)
            assert a[:28] == b[:28] and a[280:284] == b[280:284], name
        # a grid of the wrong type is refused like readGridUni does (iogrids.cpp:494)
        with pytest.raises(M.MantaError):
            s.create(M.RealGrid).load(os.path.join(d, "ref_vel.uni"))


@pytest.mark.gpu
def test_numpy_bridge_plugins():
    """copyArrayToGrid* / copyGridToArray* (numpyconvert.cpp:145-222): C-order arrays of shape (z, y, x[, 3]), sizes checked"""
    from mantaflow_b200 import manta as M
    size = C.S3
    ins = _inputs(size)
    s = M.Solver(name="main", gridSize=M.vec3(*size), dim=3)
    for name, gt, up, down in [("density", "RealGrid", "copyArrayToGridReal", "copyGridToArrayReal"),
                               ("phi", "LevelsetGrid", "copyArrayToGridLevelset", "copyGridToArrayLevelset"),
                               ("vec", "VecGrid", "copyArrayToGridVec3", "copyGridToArrayVec3"),
                               ("vel", "MACGrid", "copyArrayToGridMAC", "copyGridToArrayMAC"),
                               ("flags", "FlagGrid", "copyArrayToGridFlag", "copyGridToArrayFlag")]:
        g = s.create(getattr(M, gt))
        getattr(M, up)(source=ins[name], target=g)
        assert np.array_equal(g.to_numpy(), ins[name])
        out = np.zeros_like(ins[name])
        getattr(M, down)(source=g, target=out)
        assert np.array_equal(out, ins[name])
        with pytest.raises(M.MantaError):
            getattr(M, up)(source=ins[name].ravel()[:-3], target=g)
        if gt in ("RealGrid", "VecGrid"):     # float64 arrays are converted element by element (N_DOUBLE branch)
            getattr(M, up)(source=ins[name].astype(np.float64), target=g)
            assert np.array_equal(g.to_numpy(), ins[name])
            out64 = np.zeros(ins[name].shape, np.float64)
            getattr(M, down)(source=g, target=out64)
            assert np.array_equal(out64, ins[name].astype(np.float64))
