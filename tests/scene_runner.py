"""Runs one test 'case' (a small scene described as data) against a manta-style API.

The SAME case description is executed
  * by the unmodified reference (`oracle/_ref/manta tests/golden/ref_driver.py`, in generate mode), and
  * by mantaflow_b200.manta on the GPU (tests/test_*_gpu.py),
so the parity tests are literally the same scene calls on identical inputs.

case = {
  "size": [nx, ny, nz], "dim": 3, "dt": 1.0,
  "grids": {"flags": "FlagGrid", "vel": "MACGrid", ...},
  "load":  ["vel", "density"],          # grids whose initial content comes from inputs (in_<name>)
  "ops":   [["call", "advectSemiLagrange", {"flags": "@flags", "vel": "@vel", "grid": "@density", "order": 2}],
            ["method", "flags", "initDomain", {"boundaryWidth": 0}],
            ["repeat", 5, [ ...ops... ]]],
  "save":  ["density", "vel"],
}
Values: "@name" -> grid object, ["vec3", x, y, z] -> vec3, everything else literal.
"""
import gzip
import os

# numpy is imported lazily: the reference's embedded interpreter may not have it
DTYPES = {"FlagGrid": "int32", "IntGrid": "int32"}
NCOMP = {"MACGrid": 3, "VecGrid": 3, "Vec3Grid": 3}


def grid_shape(case, gtype):
    nx, ny, nz = case["size"]
    return (nz, ny, nx, 3) if NCOMP.get(gtype, 1) == 3 else (nz, ny, nx)


PART_HEADER = 4 + 6 * 4 + 256 + 8      # "PD01" + UniPartHeader (fileio/ioparticles.cpp:37-43)


def read_pdata_uni(path, ncomp=3):
    """particle data written by ParticleDataImpl::save (writePdataUni, fileio/ioparticles.cpp:226): float32 values after the header"""
    import numpy as np
    import struct
    with gzip.open(path, "rb") as f:
        raw = f.read()
    assert raw[:4] == b"PD01", raw[:4]
    n, = struct.unpack("<i", raw[4:8])
    a = np.frombuffer(raw[PART_HEADER:], dtype=np.float32)
    assert a.size == n * ncomp, (a.size, n, ncomp)
    return a.reshape(n, ncomp).copy() if ncomp > 1 else a.copy()


def is_particle_type(gtype):
    return gtype.split(":")[0] in ("BasicParticleSystem", "PdataVec3", "PdataReal")


def write_raw(path, arr):
    import numpy as np
    with gzip.open(path, "wb", compresslevel=1) as f:
        f.write(np.ascontiguousarray(arr).tobytes())


def read_raw(path, dtype, shape):
    import numpy as np
    with gzip.open(path, "rb") as f:
        return np.frombuffer(f.read(), dtype=dtype).reshape(shape).copy()


def _resolve(v, grids, api):
    if isinstance(v, str) and v.startswith("@"):
        return grids[v[1:]]
    if isinstance(v, list) and len(v) == 4 and v[0] == "vec3":
        return api["vec3"](v[1], v[2], v[3])
    if isinstance(v, str) and v.startswith("$"):
        return api[v[1:]]          # module constant, e.g. "$PcMGStatic"
    return v


def _run_ops(ops, grids, api, solver, log):
    for op in ops:
        kind = op[0]
        if kind == "call":
            kw = {k: _resolve(v, grids, api) for k, v in op[2].items()}
            api[op[1]](**kw)
        elif kind == "method":
            obj = solver if op[1] == "solver" else grids[op[1]]
            kw = {k: _resolve(v, grids, api) for k, v in op[3].items()}
            r = getattr(obj, op[2])(**kw)
            if r is not None:
                log.append([op[1], op[2], float(r) if not isinstance(r, (str, bytes)) else r])
        elif kind == "shape":   # ["shape", "Box", {ctor kwargs}, "applyToGrid", {"grid": "@flags", "value": 2}]
            ckw = {k: _resolve(v, grids, api) for k, v in op[2].items()}
            shp = api[op[1]](parent=solver, **ckw)
            kw = {k: _resolve(v, grids, api) for k, v in op[4].items()}
            getattr(shp, op[3])(**kw)
        elif kind == "shape_ls":   # ["shape_ls", "Sphere", {ctor kwargs}, "phi"]: phi.copyFrom(shape.computeLevelset())
            ckw = {k: _resolve(v, grids, api) for k, v in op[2].items()}
            shp = api[op[1]](parent=solver, **ckw)
            grids[op[3]].copyFrom(shp.computeLevelset())
        elif kind == "object":   # ["object", "src", "Cylinder", {ctor kwargs}, {attributes}]: grids["src"] = solver.create(Cylinder, ...)
            ckw = {k: _resolve(v, grids, api) for k, v in op[3].items()}
            obj = solver.create(api[op[2]], **ckw)
            for k, v in (op[4] if len(op) > 4 else {}).items():
                setattr(obj, k, _resolve(v, grids, api))
            grids[op[1]] = obj
        elif kind == "repeat":
            for _ in range(op[1]):
                _run_ops(op[2], grids, api, solver, log)
                solver.step()
        else:
            raise ValueError("unknown op kind %r" % (kind,))


def run_case(case, api, loader, saver):
    """loader(name, grid, gtype) fills a grid from the case inputs; saver(name, grid, gtype) collects an output."""
    nx, ny, nz = case["size"]
    s = api["Solver"](name="main", gridSize=api["vec3"](nx, ny, nz), dim=case.get("dim", 3))
    s.timestep = case.get("dt", 1.0)
    grids = {}
    for name, gtype in case["grids"].items():
        if ":" in gtype:     # "PdataVec3:pp": particle data created by its particle system
            t, owner = gtype.split(":")
            grids[name] = grids[owner].create(api[t])
        else:
            grids[name] = s.create(api[gtype])
    for name in case.get("load", []):
        loader(name, grids[name], case["grids"][name])
    log = []
    _run_ops(case["ops"], grids, api, s, log)
    out = {}
    for name in case.get("save", []):
        out[name] = saver(name, grids[name], case["grids"][name])
    return out, log
