"""Case descriptions shared by tests/golden/make_golden.py (reference side) and the GPU parity tests.

Inputs are regenerated from seeds (numpy RandomState / MT19937 is stable across numpy versions); the golden
npz stores only the reference's OUTPUTS plus a checksum of the inputs they were produced from.
"""
import zlib

import numpy as np

S3 = [18, 22, 14]      # small odd-ish 3-D grid
S2 = [26, 20, 1]       # 2-D grid


def _coords(size):
    nx, ny, nz = size
    z, y, x = np.meshgrid(np.arange(nz), np.arange(ny), np.arange(nx), indexing="ij")
    return x.astype(np.float64), y.astype(np.float64), z.astype(np.float64)


def gen_real(size, seed, lo=0.0, hi=1.0):
    r = np.random.RandomState(seed)
    nx, ny, nz = size
    return (lo + (hi - lo) * r.rand(nz, ny, nx)).astype(np.float32)


def gen_smooth(size, seed, amp=1.0):
    r = np.random.RandomState(seed)
    x, y, z = _coords(size)
    nx, ny, nz = size
    a = r.rand(6) * 2 * np.pi
    f = np.sin(2 * np.pi * x / nx * 1.5 + a[0]) * np.cos(2 * np.pi * y / ny + a[1]) * np.cos(2 * np.pi * z / max(nz, 2) + a[2])
    f += 0.5 * np.sin(2 * np.pi * y / ny * 2 + a[3]) * np.sin(2 * np.pi * x / nx + a[4])
    f += 0.25 * (r.rand(nz, ny, nx) - 0.5)
    return (amp * f).astype(np.float32)


def gen_vel(size, seed, amp=1.7, dim=3):
    v = np.stack([gen_smooth(size, seed + 11 * c, amp) for c in range(3)], axis=-1)
    if dim == 2:
        v[..., 2] = 0
    return np.ascontiguousarray(v.astype(np.float32))


def gen_vec(size, seed, amp=1.0):
    return gen_vel(size, seed + 1000, amp)


def gen_phi_sphere(size, cx, cy, cz, rad):
    x, y, z = _coords(size)
    return (np.sqrt((x + 0.5 - cx) ** 2 + (y + 0.5 - cy) ** 2 + (z + 0.5 - cz) ** 2) - rad).astype(np.float32)


def gen_fractions(size, seed):
    r = np.random.RandomState(seed)
    nx, ny, nz = size
    f = 0.2 + 0.8 * r.rand(nz, ny, nx, 3)
    f[r.rand(nz, ny, nx, 3) > 0.6] = 1.0
    return f.astype(np.float32)


def gen_weight(size, seed):
    """Vec3 weights as left by mapPartsToMAC: positive in a blob, exactly 0 elsewhere"""
    r = np.random.RandomState(seed)
    nx, ny, nz = size
    x, y, z = _coords(size)
    w = np.zeros((nz, ny, nx, 3), np.float32)
    for c in range(3):
        blob = ((x - nx * (0.4 + 0.05 * c)) ** 2 + (y - ny * 0.45) ** 2 + (z - nz * 0.5) ** 2) < (0.27 * min(nx, ny, max(nz, ny))) ** 2
        w[..., c] = np.where(blob, 0.1 + 3 * r.rand(nz, ny, nx), 0.0)
    return w


def gen_flags_liquid(size, fill=0.55, dim=3):
    """walls on the border, fluid below a height, empty above, one obstacle block"""
    nx, ny, nz = size
    f = np.full((nz, ny, nx), 4, np.int32)
    f[:, : int(ny * fill), :] = 1
    f[:, 2:5, 3:7] = 2
    f[:, 0, :] = 2; f[:, -1, :] = 2; f[:, :, 0] = 2; f[:, :, -1] = 2
    if dim == 3:
        f[0] = 2; f[-1] = 2
    return f


def gen_flags_stick(size):
    nx, ny, nz = size
    f = np.full((nz, ny, nx), 1, np.int32)
    f[:, 0, :] = 2; f[:, -1, :] = 2; f[:, :, 0] = 2; f[:, :, -1] = 2; f[0] = 2; f[-1] = 2
    f[3:6, 5:9, 4:8] = 2 | 64
    f[:, 0, :] |= 64
    return f


def gen_flags_thin(size):
    """a fluid block with 1-cell-wide channels at odd coordinates (their even neighbours are obstacles): the channel
    vertices keep >= 2 free interpolation vertices, which forces the order-dependent (serial) branch of
    GridMg::genCoarseGrid (multigrid.cpp:543-575).  Everything is 6-connected to the block."""
    nx, ny, nz = size
    f = np.full((nz, ny, nx), 2, np.int32)
    f[2:-2, 2:8, 2:7] = 1
    for k in range(3, nz - 2, 2):
        f[k, 8:ny - 2, 3] = 1                  # risers along y at odd (i, k)
        for j in range(9, ny - 2, 4):
            f[k, j, 3:nx - 2] = 1              # branches along x at odd (j, k)
    return f


INPUTS = {}   # name -> callable(case) -> array


def inputs_for(case):
    """arrays for case['load'] from case['inputs'] = {grid: [generator, args...]}"""
    out = {}
    for name in case.get("load", []):
        spec = case["inputs"][name]
        fn = globals()[spec[0]]
        out[name] = fn(case["size"], *spec[1:])
    return out


def checksum(arrs):
    c = 0
    for k in sorted(arrs):
        c = zlib.crc32(np.ascontiguousarray(arrs[k]).tobytes(), c)
    return c


V = lambda x, y, z: ["vec3", x, y, z]  # noqa: E731

DOMAIN = [["method", "flags", "initDomain", {}], ["method", "flags", "fillGrid", {}]]
OBST3 = [["shape", "Box", {"p0": V(5, 6, 4), "p1": V(9, 11, 8)}, "applyToGrid", {"grid": "@flags", "value": 2}]]
OBST2 = [["shape", "Box", {"p0": V(8, 6, 0), "p1": V(13, 10, 1)}, "applyToGrid", {"grid": "@flags", "value": 2}]]
BASEGRIDS = {"flags": "FlagGrid", "vel": "MACGrid", "density": "RealGrid", "pressure": "RealGrid"}


def _adv_case(size, dim, gridtype, extra, seed, pre=None, dt=1.2, amp=1.7):
    grids = dict(BASEGRIDS)
    gname = {"RealGrid": "density", "MACGrid": "vel", "VecGrid": "vec"}[gridtype]
    if gridtype == "VecGrid":
        grids["vec"] = "VecGrid"
    load = ["vel"] + ([] if gname == "vel" else [gname])
    inputs = {"vel": ["gen_vel", seed, amp, dim], "density": ["gen_real", seed + 1], "vec": ["gen_vec", seed + 2]}
    ops = DOMAIN + (OBST3 if dim == 3 else OBST2) + (pre or [])
    kw = {"flags": "@flags", "vel": "@vel", "grid": "@" + gname}
    kw.update(extra)
    ops = ops + [["call", "advectSemiLagrange", kw]]
    return {"size": size, "dim": dim, "dt": dt, "grids": grids, "load": load, "inputs": {k: inputs[k] for k in load},
            "ops": ops, "save": [gname, "flags"], "exact": True}


CASES = {}

# ---- flags -------------------------------------------------------------------------------------------------
CASES["flags_domain"] = {"size": S3, "dim": 3, "grids": {"flags": "FlagGrid", "phiWalls": "RealGrid"}, "ops": [
    ["method", "flags", "initDomain", {"boundaryWidth": 1, "phiWalls": "@phiWalls"}], ["method", "flags", "fillGrid", {}]] + OBST3,
    "save": ["flags", "phiWalls"], "exact": True}
CASES["flags_open"] = {"size": S3, "dim": 3, "grids": {"flags": "FlagGrid"}, "ops": [
    ["method", "flags", "initDomain", {"boundaryWidth": 0, "open": "xY", "inflow": "z", "outflow": "X"}],
    ["method", "flags", "fillGrid", {}],
    ["call", "setOpenBound", {"flags": "@flags", "bWidth": 0, "openBound": "yZ", "type": 20}]],
    "save": ["flags"], "exact": True}
CASES["flags_open2d"] = {"size": S2, "dim": 2, "grids": {"flags": "FlagGrid"}, "ops": [
    ["method", "flags", "initDomain", {"boundaryWidth": 1}], ["method", "flags", "fillGrid", {}],
    ["call", "setOpenBound", {"flags": "@flags", "bWidth": 1, "openBound": "xY", "type": 20}]],
    "save": ["flags"], "exact": True}

# ---- advection (bit-exact) -----------------------------------------------------------------------------------
CASES["adv_real_o1"] = _adv_case(S3, 3, "RealGrid", {"order": 1}, 10)
CASES["adv_real_o2_c2"] = _adv_case(S3, 3, "RealGrid", {"order": 2}, 11)
CASES["adv_real_o2_c1"] = _adv_case(S3, 3, "RealGrid", {"order": 2, "clampMode": 1, "strength": 0.8}, 12, amp=2.6)
CASES["adv_real_o2_cubic"] = _adv_case(S3, 3, "RealGrid", {"order": 2, "orderSpace": 2}, 13)
CASES["adv_real_o1_trace2"] = _adv_case(S3, 3, "RealGrid", {"order": 1, "orderTrace": 2}, 14)
CASES["adv_real_o2_c1_cubic_trace2"] = _adv_case(S3, 3, "RealGrid", {"order": 2, "clampMode": 1, "orderSpace": 2, "orderTrace": 2}, 15)
CASES["adv_mac_o1"] = _adv_case(S3, 3, "MACGrid", {"order": 1}, 20)
CASES["adv_mac_o2_c2"] = _adv_case(S3, 3, "MACGrid", {"order": 2}, 21)
CASES["adv_mac_o2_c1"] = _adv_case(S3, 3, "MACGrid", {"order": 2, "clampMode": 1, "strength": 0.7}, 22, amp=2.6)
CASES["adv_mac_o2_trace2"] = _adv_case(S3, 3, "MACGrid", {"order": 2, "orderTrace": 2}, 23)
CASES["adv_mac_o2_cubic"] = _adv_case(S3, 3, "MACGrid", {"order": 2, "orderSpace": 2}, 24)
CASES["adv_mac_o1_cubic_trace2"] = _adv_case(S3, 3, "MACGrid", {"order": 1, "orderSpace": 2, "orderTrace": 2}, 25)
CASES["adv_vec_o1"] = _adv_case(S3, 3, "VecGrid", {"order": 1}, 30)
CASES["adv_vec_o2_c2"] = _adv_case(S3, 3, "VecGrid", {"order": 2}, 31)
CASES["adv_vec_o2_c1"] = _adv_case(S3, 3, "VecGrid", {"order": 2, "clampMode": 1}, 32, amp=2.6)
CASES["adv_vec_o2_cubic"] = _adv_case(S3, 3, "VecGrid", {"order": 2, "orderSpace": 2, "strength": 0.9}, 33)
CASES["adv2d_real_o2_c2"] = _adv_case(S2, 2, "RealGrid", {"order": 2}, 40)
CASES["adv2d_real_o2_c1_cubic"] = _adv_case(S2, 2, "RealGrid", {"order": 2, "clampMode": 1, "orderSpace": 2}, 41)
CASES["adv2d_mac_o2_c2"] = _adv_case(S2, 2, "MACGrid", {"order": 2}, 42)
CASES["adv2d_mac_o2_c1"] = _adv_case(S2, 2, "MACGrid", {"order": 2, "clampMode": 1}, 43)
CASES["adv2d_mac_o1_trace2"] = _adv_case(S2, 2, "MACGrid", {"order": 1, "orderTrace": 2}, 44)
_OPEN3 = [["call", "setOpenBound", {"flags": "@flags", "bWidth": 0, "openBound": "xYz", "type": 20}]]
CASES["adv_mac_o2_outflow"] = _adv_case(S3, 3, "MACGrid", {"order": 2}, 50, pre=_OPEN3)
CASES["adv_mac_o1_outflow"] = _adv_case(S3, 3, "MACGrid", {"order": 1}, 51, pre=_OPEN3)
CASES["adv2d_mac_o2_outflow"] = _adv_case(S2, 2, "MACGrid", {"order": 2}, 52,
                                          pre=[["call", "setOpenBound", {"flags": "@flags", "bWidth": 0, "openBound": "XY", "type": 20}]])
CASES["adv_real_o2_bigvel"] = _adv_case(S3, 3, "RealGrid", {"order": 2, "clampMode": 1}, 53, amp=9.0)
CASES["adv_mac_o2_bigvel"] = _adv_case(S3, 3, "MACGrid", {"order": 2, "clampMode": 2}, 54, amp=9.0)


# ---- boundary conditions / forces (bit-exact) ------------------------------------------------------------------
def _bc_case(size, dim, ops, load, inputs, save=("vel",), grids=None):
    g = dict(BASEGRIDS)
    g.update(grids or {})
    return {"size": size, "dim": dim, "dt": 0.7, "grids": g, "load": load, "inputs": inputs,
            "ops": DOMAIN + (OBST3 if dim == 3 else OBST2) + ops, "save": list(save), "exact": True}


CASES["wallbcs"] = _bc_case(S3, 3, [["call", "setWallBcs", {"flags": "@flags", "vel": "@vel"}]], ["vel"], {"vel": ["gen_vel", 60, 1.0, 3]})
CASES["wallbcs_2d"] = _bc_case(S2, 2, [["call", "setWallBcs", {"flags": "@flags", "vel": "@vel"}]], ["vel"], {"vel": ["gen_vel", 61, 1.0, 3]})
CASES["wallbcs_obvel"] = _bc_case(S3, 3, [["call", "setWallBcs", {"flags": "@flags", "vel": "@vel", "obvel": "@obvel"}]], ["vel", "obvel"],
                                  {"vel": ["gen_vel", 62, 1.0, 3], "obvel": ["gen_vel", 63, 0.5, 3]}, grids={"obvel": "MACGrid"})
CASES["wallbcs_stick"] = {"size": S3, "dim": 3, "dt": 1.0, "grids": dict(BASEGRIDS), "load": ["vel", "flags"],
                          "inputs": {"vel": ["gen_vel", 64, 1.0, 3], "flags": ["gen_flags_stick"]},
                          "ops": [["call", "setWallBcs", {"flags": "@flags", "vel": "@vel"}]], "save": ["vel"], "exact": True}
CASES["wallbcs_frac"] = _bc_case(S3, 3, [["call", "setWallBcs", {"flags": "@flags", "vel": "@vel", "fractions": "@fractions", "phiObs": "@phiObs"}]],
                                 ["vel", "fractions", "phiObs"],
                                 {"vel": ["gen_vel", 65, 1.0, 3], "fractions": ["gen_fractions", 66], "phiObs": ["gen_phi_sphere", 7.0, 8.5, 6.0, 3.2]},
                                 grids={"fractions": "MACGrid", "phiObs": "RealGrid"})
CASES["wallbcs_frac_2d"] = _bc_case(S2, 2, [["call", "setWallBcs", {"flags": "@flags", "vel": "@vel", "fractions": "@fractions", "phiObs": "@phiObs"}]],
                                    ["vel", "fractions", "phiObs"],
                                    {"vel": ["gen_vel", 67, 1.0, 2], "fractions": ["gen_fractions", 68], "phiObs": ["gen_phi_sphere", 10.5, 8.0, 0.5, 3.2]},
                                    grids={"fractions": "MACGrid", "phiObs": "RealGrid"})
CASES["buoyancy"] = _bc_case(S3, 3, [["call", "addBuoyancy", {"flags": "@flags", "density": "@density", "vel": "@vel", "gravity": V(0.01, -0.6, 0.02), "coefficient": 0.9}]],
                             ["vel", "density"], {"vel": ["gen_vel", 70, 1.0, 3], "density": ["gen_real", 71]})
CASES["buoyancy_2d_noscale"] = _bc_case(S2, 2, [["call", "addBuoyancy", {"flags": "@flags", "density": "@density", "vel": "@vel", "gravity": V(0, -0.006, 0), "scale": False}]],
                                        ["vel", "density"], {"vel": ["gen_vel", 72, 1.0, 2], "density": ["gen_real", 73]})
CASES["gravity"] = {"size": S3, "dim": 3, "dt": 0.7, "grids": dict(BASEGRIDS), "load": ["vel", "flags", "density"],
                    "inputs": {"vel": ["gen_vel", 74, 1.0, 3], "flags": ["gen_flags_liquid"], "density": ["gen_real", 75, -1.0, 1.0]},
                    "ops": [["call", "addGravity", {"flags": "@flags", "vel": "@vel", "gravity": V(0.0, -0.3, 0.01), "exclude": "@density"}]],
                    "save": ["vel"], "exact": True}
CASES["inflow"] = _bc_case(S3, 3, [["call", "setInflowBcs", {"vel": "@vel", "dir": "xY", "value": V(0.5, 0.25, -0.1)}]], ["vel"], {"vel": ["gen_vel", 76, 1.0, 3]})


# ---- pressure: RHS and velocity correction (bit-exact), solves (tolerance) ------------------------------------
def _p_case(size, dim, ops, load, inputs, save, grids=None, exact=False, tol=None, setup=None, dt=1.0):
    g = dict(BASEGRIDS)
    g.update(grids or {})
    pre = setup if setup is not None else (DOMAIN + (OBST3 if dim == 3 else OBST2))
    c = {"size": size, "dim": dim, "dt": dt, "grids": g, "load": load, "inputs": inputs, "ops": pre + ops, "save": list(save), "exact": exact}
    if tol:
        c["tol"] = tol
    return c


_WB = [["call", "setWallBcs", {"flags": "@flags", "vel": "@vel"}]]
CASES["rhs_plain"] = _p_case(S3, 3, _WB + [["call", "computePressureRhs", {"rhs": "@rhs", "vel": "@vel", "pressure": "@pressure", "flags": "@flags"}]],
                             ["vel"], {"vel": ["gen_vel", 80, 1.0, 3]}, ["rhs"], grids={"rhs": "RealGrid"}, exact=True)
CASES["rhs_2d"] = _p_case(S2, 2, _WB + [["call", "computePressureRhs", {"rhs": "@rhs", "vel": "@vel", "pressure": "@pressure", "flags": "@flags"}]],
                          ["vel"], {"vel": ["gen_vel", 81, 1.0, 2]}, ["rhs"], grids={"rhs": "RealGrid"}, exact=True)
CASES["rhs_fractions_obvel_corr"] = _p_case(S3, 3, [["call", "computePressureRhs", {"rhs": "@rhs", "vel": "@vel", "pressure": "@pressure", "flags": "@flags",
                                            "fractions": "@fractions", "obvel": "@obvel", "perCellCorr": "@density"}]],
                                            ["vel", "fractions", "obvel", "density"],
                                            {"vel": ["gen_vel", 82, 1.0, 3], "fractions": ["gen_fractions", 83], "obvel": ["gen_vel", 84, 0.3, 3], "density": ["gen_real", 85, -0.1, 0.1]},
                                            ["rhs"], grids={"rhs": "RealGrid", "fractions": "MACGrid", "obvel": "MACGrid"}, exact=True)
CASES["rhs_surftens"] = _p_case(S3, 3, [["call", "computePressureRhs", {"rhs": "@rhs", "vel": "@vel", "pressure": "@pressure", "flags": "@flags",
                                "phi": "@phi", "curv": "@curv", "surfTens": 0.3, "gfClamp": 0.01}]],
                                ["vel", "flags", "phi", "curv"],
                                {"vel": ["gen_vel", 86, 1.0, 3], "flags": ["gen_flags_liquid"], "phi": ["gen_smooth", 87, 1.5], "curv": ["gen_smooth", 88, 0.4]},
                                ["rhs"], grids={"rhs": "RealGrid", "phi": "RealGrid", "curv": "RealGrid"}, exact=True, setup=[])
CASES["rhs_compat"] = _p_case(S3, 3, _WB + [["call", "computePressureRhs", {"rhs": "@rhs", "vel": "@vel", "pressure": "@pressure", "flags": "@flags", "enforceCompatibility": True}]],
                              ["vel"], {"vel": ["gen_vel", 89, 1.0, 3]}, ["rhs"], grids={"rhs": "RealGrid"}, tol={"rhs": 2e-6})
CASES["correct_vel"] = _p_case(S3, 3, [["call", "correctVelocity", {"vel": "@vel", "pressure": "@pressure", "flags": "@flags"}]],
                               ["vel", "flags", "pressure"], {"vel": ["gen_vel", 90, 1.0, 3], "flags": ["gen_flags_liquid"], "pressure": ["gen_smooth", 91, 2.0]},
                               ["vel"], exact=True, setup=[])
CASES["correct_vel_2d"] = _p_case(S2, 2, [["call", "correctVelocity", {"vel": "@vel", "pressure": "@pressure", "flags": "@flags"}]],
                                  ["vel", "flags", "pressure"], {"vel": ["gen_vel", 92, 1.0, 2], "flags": ["gen_flags_liquid", 0.55, 2], "pressure": ["gen_smooth", 93, 2.0]},
                                  ["vel"], exact=True, setup=[])
CASES["correct_vel_gf"] = _p_case(S3, 3, [["call", "correctVelocity", {"vel": "@vel", "pressure": "@pressure", "flags": "@flags", "phi": "@phi", "curv": "@curv", "surfTens": 0.2, "gfClamp": 0.05}]],
                                  ["vel", "flags", "pressure", "phi", "curv"],
                                  {"vel": ["gen_vel", 94, 1.0, 3], "flags": ["gen_flags_liquid"], "pressure": ["gen_smooth", 95, 2.0], "phi": ["gen_smooth", 96, 1.5], "curv": ["gen_smooth", 97, 0.4]},
                                  ["vel"], grids={"phi": "RealGrid", "curv": "RealGrid"}, exact=True, setup=[])

# solves: same stopping rule as the reference; tolerances are relative L-inf (see tests/test_parity_gpu.py)
_TIGHT = {"cgAccuracy": 1e-6, "cgMaxIterFac": 99, "zeroPressureFixing": True}


def _solve(extra, seed=100, size=S3, dim=3, setup=None, load=("vel",), inputs=None, grids=None, tol=None, wb=True):
    kw = {"flags": "@flags", "vel": "@vel", "pressure": "@pressure"}
    kw.update(extra)
    inp = {"vel": ["gen_vel", seed, 0.6, dim]}
    inp.update(inputs or {})
    return _p_case(size, dim, (_WB if wb else []) + [["call", "solvePressure", kw]], list(load), inp, ["pressure", "vel"],
                   grids=grids, tol=tol or {"pressure": 1e-4, "vel": 1e-4}, setup=setup)


CASES["solve_mic_tight"] = _solve(dict(_TIGHT))
CASES["solve_mgstatic_tight"] = _solve(dict(_TIGHT, preconditioner="$PcMGStatic"), seed=101)
CASES["solve_mgdynamic_tight"] = _solve(dict(_TIGHT, preconditioner="$PcMGDynamic"), seed=102)
# closed box without a pinned cell: the pressure is defined up to a constant -> compared after removing the mean over the fluid cells
CASES["solve_mic_default"] = _solve({}, seed=103, tol={"pressure": ["meanfree", 3e-2], "vel": 3e-2})
CASES["solve_mg_default"] = _solve({"preconditioner": "$PcMGStatic", "zeroPressureFixing": True}, seed=104, tol={"pressure": 3e-2, "vel": 3e-2})
CASES["solve_2d_tight"] = _solve(dict(_TIGHT), seed=105, size=S2, dim=2)
CASES["solve_2d_mg_tight"] = _solve(dict(_TIGHT, preconditioner="$PcMGDynamic"), seed=106, size=S2, dim=2)
CASES["solve_l2norm"] = _solve({"cgAccuracy": 1e-10, "useL2Norm": True, "zeroPressureFixing": True, "cgMaxIterFac": 99}, seed=107)
CASES["solve_liquid_gf"] = _solve({"cgAccuracy": 1e-6, "cgMaxIterFac": 99, "phi": "@phi", "gfClamp": 0.01}, seed=108, setup=[], wb=True,
                                  load=("vel", "flags", "phi"), inputs={"flags": ["gen_flags_liquid"], "phi": ["gen_smooth", 109, 1.5]}, grids={"phi": "RealGrid"})
CASES["solve_liquid_mg"] = _solve({"cgAccuracy": 1e-6, "cgMaxIterFac": 99, "preconditioner": "$PcMGDynamic"}, seed=110, setup=[],
                                  load=("vel", "flags"), inputs={"flags": ["gen_flags_liquid"]})
CASES["solve_fractions"] = _solve(dict(_TIGHT, fractions="@fractions", obvel="@obvel"), seed=111, load=("vel", "fractions", "obvel"),
                                  inputs={"fractions": ["gen_fractions", 112], "obvel": ["gen_vel", 113, 0.1, 3]}, grids={"fractions": "MACGrid", "obvel": "MACGrid"})
CASES["solve_thin_mg"] = _solve(dict(_TIGHT, preconditioner="$PcMGDynamic"), seed=114, setup=[], load=("vel", "flags"), inputs={"flags": ["gen_flags_thin"]})
CASES["solve_retrhs"] = _p_case(S3, 3, _WB + [["call", "solvePressure", dict({"flags": "@flags", "vel": "@vel", "pressure": "@pressure", "retRhs": "@rhs"}, **_TIGHT)]],
                                ["vel"], {"vel": ["gen_vel", 115, 0.6, 3]}, ["rhs"], grids={"rhs": "RealGrid"}, exact=True)
CASES["diffusion_real"] = _p_case(S3, 3, [["call", "cgSolveDiffusion", {"flags": "@flags", "grid": "@density", "alpha": 0.4, "cgAccuracy": 1e-8}]],
                                  ["density"], {"density": ["gen_real", 116]}, ["density"], tol={"density": 1e-5})
CASES["diffusion_mac_2d"] = _p_case(S2, 2, [["call", "cgSolveDiffusion", {"flags": "@flags", "grid": "@vel", "alpha": 0.3, "cgAccuracy": 1e-8}]],
                                    ["vel"], {"vel": ["gen_vel", 117, 1.0, 2]}, ["vel"], tol={"vel": 1e-5})

# ---- multi-step plume (the reference's test_2010_plume3d.py loop, smaller grid) -----------------------------------
_PLUME_STEP = [
    ["shape", "Cylinder", {"center": V(12, 3.2, 10), "radius": 3.4, "z": V(0, 0.64, 0)}, "applyToGrid", {"grid": "@density", "value": 1}],
    ["call", "advectSemiLagrange", {"flags": "@flags", "vel": "@vel", "grid": "@density", "order": 2, "clampMode": 1}],
    ["call", "advectSemiLagrange", {"flags": "@flags", "vel": "@vel", "grid": "@vel", "order": 2, "clampMode": 1}],
    ["call", "setWallBcs", {"flags": "@flags", "vel": "@vel"}],
    ["call", "addBuoyancy", {"density": "@density", "vel": "@vel", "gravity": V(0, -6e-3, 0), "flags": "@flags"}],
    ["call", "solvePressure", {"flags": "@flags", "vel": "@vel", "pressure": "@pressure"}],
    ["call", "setWallBcs", {"flags": "@flags", "vel": "@vel"}],
]
CASES["plume3d_mic"] = {"size": [24, 32, 20], "dim": 3, "dt": 1.0, "grids": dict(BASEGRIDS), "load": [], "inputs": {},
                        "ops": DOMAIN + [["repeat", 15, _PLUME_STEP]], "save": ["density", "vel"], "exact": False,
                        # thresholds of test_2010_plume3d.py:72-73 are 1e-3 / 5e-3.  PcMIC runs the red-black IC stand-in here, so
                        # at cgAccuracy=1e-3 the two solvers stop at different iterates (MIC vs MG inside the
                        # reference itself differ by 8e-4 after ONE step); measured 1.4e-3 on density after 15 steps.
                        "abs_tol": {"density": 3e-3, "vel": 5e-3}}
_PLUME_STEP_MG = [op if op[1] != "solvePressure" else ["call", "solvePressure", dict(op[2], preconditioner="$PcMGStatic", zeroPressureFixing=True)] for op in _PLUME_STEP]
CASES["plume3d_mg"] = {"size": [24, 32, 20], "dim": 3, "dt": 1.0, "grids": dict(BASEGRIDS), "load": [], "inputs": {},
                       "ops": DOMAIN + [["repeat", 15, _PLUME_STEP_MG]], "save": ["density", "vel"], "exact": False,
                       "abs_tol": {"density": 1e-3, "vel": 5e-3}}


# ---- simple extrapolation plugins (fastmarch.cpp:337-553), SURVEY 8(f) rank 2: bit-exact ------------------------------
def _ex_case(size, dim, call, kw, load, inputs, save, grids):
    g = dict(BASEGRIDS); g.update(grids)
    return {"size": size, "dim": dim, "dt": 1.0, "grids": g, "load": load, "inputs": inputs, "ops": [["call", call, kw]], "save": save, "exact": True}


CASES["extrap_mac_simple"] = _ex_case(S3, 3, "extrapolateMACSimple", {"flags": "@flags", "vel": "@vel"}, ["vel", "flags"],
                                      {"vel": ["gen_vel", 120, 1.0, 3], "flags": ["gen_flags_liquid"]}, ["vel"], {})
CASES["extrap_mac_simple_d2"] = _ex_case(S3, 3, "extrapolateMACSimple", {"flags": "@flags", "vel": "@vel", "distance": 2}, ["vel", "flags"],
                                         {"vel": ["gen_vel", 121, 1.0, 3], "flags": ["gen_flags_liquid", 0.35]}, ["vel"], {})
CASES["extrap_mac_intoobs_phi"] = _ex_case(S3, 3, "extrapolateMACSimple", {"flags": "@flags", "vel": "@vel", "distance": 3, "phiObs": "@phiObs", "intoObs": True},
                                           ["vel", "flags", "phiObs"],
                                           {"vel": ["gen_vel", 122, 1.0, 3], "flags": ["gen_flags_liquid"], "phiObs": ["gen_phi_sphere", 7.0, 6.5, 6.0, 3.4]},
                                           ["vel"], {"phiObs": "LevelsetGrid"})
CASES["extrap_mac_simple_2d"] = _ex_case(S2, 2, "extrapolateMACSimple", {"flags": "@flags", "vel": "@vel", "distance": 5}, ["vel", "flags"],
                                         {"vel": ["gen_vel", 123, 1.0, 2], "flags": ["gen_flags_liquid", 0.5, 2]}, ["vel"], {})
CASES["extrap_mac_weight"] = _ex_case(S3, 3, "extrapolateMACFromWeight", {"vel": "@vel", "weight": "@weight", "distance": 3}, ["vel", "weight"],
                                      {"vel": ["gen_vel", 124, 1.0, 3], "weight": ["gen_weight", 125]}, ["vel", "weight"], {"weight": "VecGrid"})
CASES["extrap_mac_weight_2d"] = _ex_case(S2, 2, "extrapolateMACFromWeight", {"vel": "@vel", "weight": "@weight"}, ["vel", "weight"],
                                         {"vel": ["gen_vel", 126, 1.0, 2], "weight": ["gen_weight", 127]}, ["vel", "weight"], {"weight": "VecGrid"})
CASES["extrap_ls_simple"] = _ex_case(S3, 3, "extrapolateLsSimple", {"phi": "@phi", "distance": 5}, ["phi"],
                                     {"phi": ["gen_phi_sphere", 8.0, 10.5, 7.0, 3.7]}, ["phi"], {"phi": "LevelsetGrid"})
CASES["extrap_ls_inside_walls"] = _ex_case(S3, 3, "extrapolateLsSimple", {"phi": "@phi", "distance": 3, "inside": True, "include_walls": True}, ["phi"],
                                           {"phi": ["gen_phi_sphere", 8.0, 10.5, 7.0, 5.2]}, ["phi"], {"phi": "LevelsetGrid"})
CASES["extrap_ls_simple_2d"] = _ex_case(S2, 2, "extrapolateLsSimple", {"phi": "@phi", "distance": 4, "inside": True}, ["phi"],
                                        {"phi": ["gen_phi_sphere", 12.0, 9.5, 0.5, 5.5]}, ["phi"], {"phi": "LevelsetGrid"})
CASES["extrap_vec3_simple"] = _ex_case(S3, 3, "extrapolateVec3Simple", {"vel": "@vec", "phi": "@phi", "distance": 4}, ["vec", "phi"],
                                       {"vec": ["gen_vec", 128], "phi": ["gen_phi_sphere", 8.0, 10.5, 7.0, 3.7]}, ["vec"], {"vec": "VecGrid", "phi": "LevelsetGrid"})


# ---- scene-setup plugins inside the step loop: shapes, levelset flags, fractions -- bit-exact ----------
_BOX = {"p0": V(4.3, 5.1, 3.2), "p1": V(11.6, 12.4, 9.9)}
_BOX2 = {"p0": V(6.2, 4.4, 0), "p1": V(15.7, 12.3, 1)}
_SPH = {"center": V(8.3, 10.2, 6.7), "radius": 4.1}
_SPHS = {"center": V(8.3, 10.2, 6.7), "radius": 3.3, "scale": V(1.0, 1.7, 0.8)}
_CYL = {"center": V(8.3, 10.2, 6.7), "radius": 3.1, "z": V(0.6, 2.0, 0.9)}


def _shape_case(size, dim, ops, save, grids, load=(), inputs=None):
    g = dict(BASEGRIDS); g.update(grids)
    return {"size": size, "dim": dim, "dt": 1.0, "grids": g, "load": list(load), "inputs": inputs or {}, "ops": ops, "save": save, "exact": True}


CASES["shape_box_sdf"] = _shape_case(S3, 3, [["shape_ls", "Box", _BOX, "phi"]], ["phi"], {"phi": "LevelsetGrid"})
CASES["shape_box_sdf_2d"] = _shape_case(S2, 2, [["shape_ls", "Box", _BOX2, "phi"]], ["phi"], {"phi": "LevelsetGrid"})
CASES["shape_sphere_sdf"] = _shape_case(S3, 3, [["shape_ls", "Sphere", _SPHS, "phi"]], ["phi"], {"phi": "LevelsetGrid"})
CASES["shape_cylinder_sdf"] = _shape_case(S3, 3, [["shape_ls", "Cylinder", _CYL, "phi"]], ["phi"], {"phi": "LevelsetGrid"})
CASES["shape_apply_mix"] = _shape_case(S3, 3, DOMAIN + OBST3 + [
    ["shape", "Sphere", _SPH, "applyToGrid", {"grid": "@density", "value": 0.75, "respectFlags": "@flags"}],
    ["shape", "Sphere", _SPHS, "applyToGrid", {"grid": "@vel", "value": V(0.5, -1.5, 2.5)}],
    ["shape", "Cylinder", _CYL, "applyToGrid", {"grid": "@vec", "value": V(1.0, 2.0, 3.0), "respectFlags": "@flags"}],
    ["shape", "Cylinder", _CYL, "applyToGrid", {"grid": "@flags", "value": 4}],
    ["shape", "Box", _BOX, "applyToGrid", {"grid": "@pressure", "value": -2.0}],
    ["shape", "Box", {"center": V(9, 11, 7), "size": V(3.5, 2.5, 4.5)}, "applyToGrid", {"grid": "@vel", "value": V(7.0, 8.0, 9.0), "respectFlags": "@flags"}]],
    ["density", "vel", "vec", "flags", "pressure"], {"vec": "VecGrid"}, load=["vel"], inputs={"vel": ["gen_vel", 130, 1.0, 3]})
CASES["flags_from_levelset"] = _shape_case(S3, 3, [
    ["method", "flags", "initDomain", {"boundaryWidth": 0, "outflow": "Y"}], ["method", "flags", "fillGrid", {}]] + OBST3 + [
    ["shape_ls", "Sphere", _SPH, "phi"], ["shape_ls", "Box", {"p0": V(0, 0, 0), "p1": V(18, 5.5, 14)}, "phi2"],
    ["method", "phi", "join", {"o": "@phi2"}],
    ["method", "flags", "updateFromLevelset", {"levelset": "@phi"}]],
    ["flags", "phi"], {"phi": "LevelsetGrid", "phi2": "LevelsetGrid"})
_FRAC_OPS = [["method", "flags", "initDomain", {"boundaryWidth": 0, "open": "xY", "inflow": "Z", "outflow": "y"}], ["method", "flags", "fillGrid", {}],
             ["shape_ls", "Sphere", _SPH, "phiObs"]]
CASES["update_fractions"] = _shape_case(S3, 3, _FRAC_OPS + [
    ["call", "updateFractions", {"flags": "@flags", "phiObs": "@phiObs", "fractions": "@fractions"}]],
    ["fractions"], {"phiObs": "LevelsetGrid", "fractions": "MACGrid"})
CASES["update_fractions_bw1"] = _shape_case(S3, 3, [["method", "flags", "initDomain", {"boundaryWidth": 1, "open": "XyzZ"}], ["method", "flags", "fillGrid", {}],
    ["shape_ls", "Box", _BOX, "phiObs"],
    ["call", "updateFractions", {"flags": "@flags", "phiObs": "@phiObs", "fractions": "@fractions", "boundaryWidth": 1, "fracThreshold": 0.2}]],
    ["fractions"], {"phiObs": "LevelsetGrid", "fractions": "MACGrid"})
CASES["update_fractions_2d"] = _shape_case(S2, 2, [["method", "flags", "initDomain", {"boundaryWidth": 0, "open": "XY"}], ["method", "flags", "fillGrid", {}],
    ["shape_ls", "Sphere", {"center": V(12.4, 9.7, 0.5), "radius": 4.2}, "phiObs"],
    ["call", "updateFractions", {"flags": "@flags", "phiObs": "@phiObs", "fractions": "@fractions"}]],
    ["fractions"], {"phiObs": "LevelsetGrid", "fractions": "MACGrid"})
CASES["set_obstacle_flags"] = _shape_case(S3, 3, _FRAC_OPS + [
    ["call", "updateFractions", {"flags": "@flags", "phiObs": "@phiObs", "fractions": "@fractions"}],
    ["shape_ls", "Box", {"p0": V(0, 18.5, 0), "p1": V(18, 22, 14)}, "phiOut"], ["shape_ls", "Cylinder", _CYL, "phiIn"],
    ["call", "setObstacleFlags", {"flags": "@flags", "phiObs": "@phiObs", "fractions": "@fractions", "phiOut": "@phiOut", "phiIn": "@phiIn"}]],
    ["flags"], {"phiObs": "LevelsetGrid", "fractions": "MACGrid", "phiOut": "LevelsetGrid", "phiIn": "LevelsetGrid"})
CASES["set_obstacle_flags_phi"] = _shape_case(S3, 3, DOMAIN + [["shape_ls", "Sphere", _SPHS, "phiObs"],
    ["call", "setObstacleFlags", {"flags": "@flags", "phiObs": "@phiObs", "boundaryWidth": 2}]],
    ["flags"], {"phiObs": "LevelsetGrid"})


# ---- wavelet noise sources: densityInflow / addNoise with the reference's NoiseField ------------------
_NZ_PLUME = {"posScale": V(45, 45, 45), "clamp": True, "clampNeg": 0, "clampPos": 1, "valOffset": 0.75, "timeAnim": 0.2}   # simpleplume.py:22-28
_NZ_FREE = {"posScale": V(7.5, 11.0, 3.25), "posOffset": V(0.3, -1.7, 2.2), "valScale": 1.5, "valOffset": -0.1, "timeAnim": 1.3}
CASES["noise_inflow"] = _shape_case(S3, 3, DOMAIN + OBST3 + [
    ["object", "nz", "NoiseField", {}, _NZ_PLUME], ["object", "src", "Cylinder", _CYL],
    ["repeat", 3, [["call", "densityInflow", {"flags": "@flags", "density": "@density", "noise": "@nz", "shape": "@src", "scale": 1, "sigma": 0.5}]]]],
    ["density"], {})
CASES["noise_inflow_sigma0"] = _shape_case(S3, 3, DOMAIN + [
    ["object", "nz", "NoiseField", {"fixedSeed": 42}, _NZ_FREE], ["object", "src", "Sphere", _SPH],
    ["repeat", 2, [["call", "densityInflow", {"flags": "@flags", "density": "@density", "noise": "@nz", "shape": "@src", "scale": 0.8}]]]],
    ["density"], {}, load=["density"], inputs={"density": ["gen_real", 140, -0.5, 0.5]})
CASES["noise_inflow_2d"] = _shape_case(S2, 2, DOMAIN + [
    ["object", "nz", "NoiseField", {"fixedSeed": 7}, _NZ_PLUME], ["object", "src", "Box", {"p0": V(6.5, 3.0, 0), "p1": V(17.0, 9.5, 1)}],
    ["repeat", 2, [["call", "densityInflow", {"flags": "@flags", "density": "@density", "noise": "@nz", "shape": "@src", "scale": 1.25, "sigma": 1.5}]]]],
    ["density"], {})
CASES["noise_add"] = _shape_case(S3, 3, DOMAIN + OBST3 + [
    ["object", "nz", "NoiseField", {}, _NZ_FREE], ["shape_ls", "Sphere", _SPH, "phi"],
    ["repeat", 2, [["call", "addNoise", {"flags": "@flags", "density": "@density", "noise": "@nz", "sdf": "@phi", "scale": 0.6}],
                   ["call", "addNoise", {"flags": "@flags", "density": "@pressure", "noise": "@nz"}]]]],
    ["density", "pressure"], {"phi": "LevelsetGrid"}, load=["density"], inputs={"density": ["gen_real", 141]})
CASES["noise_add_2d"] = _shape_case(S2, 2, DOMAIN + [
    ["object", "nz", "NoiseField", {"fixedSeed": 99}, _NZ_FREE],
    ["repeat", 2, [["call", "addNoise", {"flags": "@flags", "density": "@density", "noise": "@nz", "scale": 2.0}]]]],
    ["density"], {})

# the reference's hello-world scene (scenes/simpleplume.py: noise-textured smoke source, MacCormack advection, buoyancy,
# pressure projection) on a 24 x 36 x 24 grid, 12 steps; tolerances as for plume3d_mic / plume3d_mg above
_SIMPLEPLUME_STEP = [
    ["call", "densityInflow", {"flags": "@flags", "density": "@density", "noise": "@nz", "shape": "@src", "scale": 1, "sigma": 0.5}],
    ["call", "advectSemiLagrange", {"flags": "@flags", "vel": "@vel", "grid": "@density", "order": 2}],
    ["call", "advectSemiLagrange", {"flags": "@flags", "vel": "@vel", "grid": "@vel", "order": 2, "strength": 1.0}],
    ["call", "setWallBcs", {"flags": "@flags", "vel": "@vel"}],
    ["call", "addBuoyancy", {"density": "@density", "vel": "@vel", "gravity": V(0, -6e-4, 0), "flags": "@flags"}],
    ["call", "solvePressure", {"flags": "@flags", "vel": "@vel", "pressure": "@pressure"}],
]
_SIMPLEPLUME_SETUP = [["object", "nz", "NoiseField", {"loadFromFile": True}, _NZ_PLUME],
                      ["object", "src", "Cylinder", {"center": V(12, 3.6, 12), "radius": 3.36, "z": V(0, 0.72, 0)}]] + DOMAIN
CASES["simpleplume_noise"] = {"size": [24, 36, 24], "dim": 3, "dt": 1.0, "grids": dict(BASEGRIDS), "load": [], "inputs": {},
                              "ops": _SIMPLEPLUME_SETUP + [["repeat", 12, _SIMPLEPLUME_STEP]], "save": ["density", "vel"], "exact": False,
                              "abs_tol": {"density": 3e-3, "vel": 5e-3}}
_SIMPLEPLUME_STEP_MG = [op if op[1] != "solvePressure" else ["call", "solvePressure", dict(op[2], preconditioner="$PcMGStatic", zeroPressureFixing=True)]
                        for op in _SIMPLEPLUME_STEP]
CASES["simpleplume_noise_mg"] = dict(CASES["simpleplume_noise"], ops=_SIMPLEPLUME_SETUP + [["repeat", 12, _SIMPLEPLUME_STEP_MG]],
                                     abs_tol={"density": 1e-3, "vel": 5e-3})

# the reference's von Karman vortex street scene (scenes/karman.py, 2-D, second-order obstacle boundaries: fractions from the
# obstacle levelset, extrapolation into the obstacle, noise-perturbed inflow) on a 48 x 24 grid, 8 steps
_KC, _KZ = V(12, 12, 0.5), V(0, 0, 1.0)
_KARMAN_SETUP = [
    ["method", "flags", "initDomain", {"inflow": "xX", "phiWalls": "@phiWalls", "boundaryWidth": 0}],
    ["shape_ls", "Cylinder", {"center": _KC, "radius": 4.8, "z": _KZ}, "phiObs"],
    ["object", "densInflow", "Cylinder", {"center": _KC, "radius": 5.04, "z": _KZ}],
    ["method", "phiObs", "join", {"o": "@phiWalls"}],
    ["call", "updateFractions", {"flags": "@flags", "phiObs": "@phiObs", "fractions": "@fractions"}],
    ["call", "setObstacleFlags", {"flags": "@flags", "phiObs": "@phiObs", "fractions": "@fractions"}],
    ["method", "flags", "fillGrid", {}],
    ["method", "vel", "setConst", {"s": V(0.9, 0, 0)}],
    ["object", "nz", "NoiseField", {"loadFromFile": True}, {"posScale": V(75, 75, 75), "clamp": True, "clampNeg": -1.0, "clampPos": 1.0}],
    ["method", "testall", "setConst", {"s": -1.0}],
    ["call", "addNoise", {"flags": "@flags", "density": "@density", "noise": "@nz", "sdf": "@testall", "scale": 0.1}],
    ["call", "setComponent", {"target": "@vel", "source": "@density", "component": 1}],
    ["method", "density", "setConst", {"s": 0.0}],
]
_KIN = {"vel": "@vel", "dir": "xX", "value": V(0.9, 0, 0)}
_KWALL = {"flags": "@flags", "vel": "@vel", "fractions": "@fractions", "phiObs": "@phiObs"}
_KARMAN_STEP = [
    ["method", "densInflow", "applyToGrid", {"grid": "@density", "value": 2.0}],
    ["call", "advectSemiLagrange", {"flags": "@flags", "vel": "@vel", "grid": "@density", "order": 2, "orderSpace": 1}],
    ["call", "advectSemiLagrange", {"flags": "@flags", "vel": "@vel", "grid": "@vel", "order": 2, "strength": 1.0}],
    ["call", "extrapolateMACSimple", {"flags": "@flags", "vel": "@vel", "distance": 2, "intoObs": True}],
    ["call", "setWallBcs", _KWALL],
    ["call", "setInflowBcs", _KIN],
    ["call", "solvePressure", {"flags": "@flags", "vel": "@vel", "pressure": "@pressure", "fractions": "@fractions", "cgAccuracy": 1e-4, "cgMaxIterFac": 5}],
    ["call", "extrapolateMACSimple", {"flags": "@flags", "vel": "@vel", "distance": 5, "intoObs": True}],
    ["call", "setWallBcs", _KWALL],
    ["call", "setInflowBcs", _KIN],
]
CASES["karman2d"] = {"size": [48, 24, 1], "dim": 2, "dt": 1.0,
                     "grids": dict(BASEGRIDS, fractions="MACGrid", phiWalls="LevelsetGrid", phiObs="LevelsetGrid", testall="RealGrid"),
                     "load": [], "inputs": {}, "ops": _KARMAN_SETUP + [["repeat", 8, _KARMAN_STEP]],
                     "save": ["density", "vel", "flags", "fractions"], "exact": False,
                     "abs_tol": {"density": 5e-3, "vel": 5e-3, "flags": 0, "fractions": 0}}


# ---- FLIP particles: sampling, RK advection, particle <-> MAC grid transfers, the flip01_simple scene --------
_PGRIDS = {"pp": "BasicParticleSystem", "pVel": "PdataVec3:pp", "velOld": "MACGrid", "tmpVec3": "VecGrid", "phi": "LevelsetGrid"}


def _flip_case(size, dim, ops, save, load=(), inputs=None, exact=True, abs_tol=None, grids=None):
    g = dict(BASEGRIDS); g.update(_PGRIDS); g.update(grids or {})
    c = {"size": size, "dim": dim, "dt": 0.5, "grids": g, "load": list(load), "inputs": inputs or {}, "ops": ops, "save": save, "exact": exact}
    if abs_tol:
        c["abs_tol"] = abs_tol
    return c


_DAM3 = [["method", "flags", "initDomain", {"boundaryWidth": 0}], ["shape_ls", "Box", {"p0": V(0, 0, 0), "p1": V(7.2, 13.2, 14)}, "phi"],
         ["method", "flags", "updateFromLevelset", {"levelset": "@phi"}]]
_DAM2 = [["method", "flags", "initDomain", {"boundaryWidth": 0}], ["shape_ls", "Box", {"p0": V(0, 0, 0), "p1": V(10.4, 12.0, 1)}, "phi"],
         ["method", "flags", "updateFromLevelset", {"levelset": "@phi"}]]
_SAMPLE3 = ["call", "sampleFlagsWithParticles", {"flags": "@flags", "parts": "@pp", "discretization": 2, "randomness": 0.4}]
_SAMPLE2 = ["call", "sampleFlagsWithParticles", {"flags": "@flags", "parts": "@pp", "discretization": 3, "randomness": 0.2}]
CASES["flip_sample"] = _flip_case(S3, 3, _DAM3 + OBST3 + [_SAMPLE3], ["pp", "flags"])
CASES["flip_sample_2d"] = _flip_case(S2, 2, _DAM2 + [_SAMPLE2], ["pp", "flags"])
_ADV = {"flags": "@flags", "vel": "@vel", "deleteInObstacle": False}
CASES["flip_advect_rk4"] = _flip_case(S3, 3, _DAM3 + OBST3 + [_SAMPLE3, ["repeat", 3, [["method", "pp", "advectInGrid", dict(_ADV, integrationMode="$IntRK4")]]]],
                                      ["pp"], load=["vel"], inputs={"vel": ["gen_vel", 150, 1.9, 3]})
CASES["flip_advect_rk2_stop"] = _flip_case(S3, 3, _DAM3 + OBST3 + [_SAMPLE3, ["repeat", 3, [
    ["method", "pp", "advectInGrid", {"flags": "@flags", "vel": "@vel", "integrationMode": "$IntRK2"}]]]],
    ["pp"], load=["vel"], inputs={"vel": ["gen_vel", 151, 2.6, 3]})
CASES["flip_advect_euler_2d"] = _flip_case(S2, 2, _DAM2 + [_SAMPLE2, ["repeat", 4, [
    ["method", "pp", "advectInGrid", dict(_ADV, integrationMode="$IntEuler", stopInObstacle=False)]]]],
    ["pp"], load=["vel"], inputs={"vel": ["gen_vel", 152, 2.2, 2]})
_TRANSFER = [["call", "mapMACToParts", {"flags": "@flags", "vel": "@vel", "parts": "@pp", "partVel": "@pVel"}],
             ["method", "pp", "advectInGrid", dict(_ADV, integrationMode="$IntRK4")],
             ["call", "mapPartsToMAC", {"vel": "@vel2", "flags": "@flags", "velOld": "@velOld", "parts": "@pp", "partVel": "@pVel", "weight": "@tmpVec3"}],
             ["call", "markFluidCells", {"parts": "@pp", "flags": "@flags"}],
             ["call", "flipVelocityUpdate", {"vel": "@vel", "velOld": "@velOld", "flags": "@flags", "parts": "@pp", "partVel": "@pVel", "flipRatio": 0.97}]]
CASES["flip_transfer"] = _flip_case(S3, 3, _DAM3 + OBST3 + [_SAMPLE3] + _TRANSFER, ["vel2", "velOld", "tmpVec3", "flags", "pVel", "pp"],
                                    load=["vel"], inputs={"vel": ["gen_vel", 153, 1.4, 3]}, grids={"vel2": "MACGrid"})
CASES["flip_transfer_2d"] = _flip_case(S2, 2, _DAM2 + [_SAMPLE2] + _TRANSFER, ["vel2", "velOld", "tmpVec3", "flags", "pVel", "pp"],
                                       load=["vel"], inputs={"vel": ["gen_vel", 154, 1.4, 2]}, grids={"vel2": "MACGrid"})
CASES["flip_map_noweight"] = _flip_case(S3, 3, _DAM3 + [_SAMPLE3, _TRANSFER[0],
    ["call", "mapPartsToMAC", {"vel": "@vel2", "flags": "@flags", "velOld": "@velOld", "parts": "@pp", "partVel": "@pVel"}]], ["vel2", "velOld"],
    load=["vel"], inputs={"vel": ["gen_vel", 155, 1.0, 3]}, grids={"vel2": "MACGrid"})

# the reference's scenes/flip01_simple.py (breaking dam, FLIP without levelset or resampling), 2-D as shipped and a 3-D variant
_FLIP01_STEP = [
    ["method", "pp", "advectInGrid", dict(_ADV, integrationMode="$IntRK4")],
    ["call", "mapPartsToMAC", {"vel": "@vel", "flags": "@flags", "velOld": "@velOld", "parts": "@pp", "partVel": "@pVel", "weight": "@tmpVec3"}],
    ["call", "extrapolateMACFromWeight", {"vel": "@vel", "distance": 2, "weight": "@tmpVec3"}],
    ["call", "markFluidCells", {"parts": "@pp", "flags": "@flags"}],
    ["call", "addGravity", {"flags": "@flags", "vel": "@vel", "gravity": V(0, -0.002, 0)}],
    ["call", "setWallBcs", {"flags": "@flags", "vel": "@vel"}],
    ["call", "solvePressure", {"flags": "@flags", "vel": "@vel", "pressure": "@pressure"}],
    ["call", "setWallBcs", {"flags": "@flags", "vel": "@vel"}],
    ["call", "extrapolateMACSimple", {"flags": "@flags", "vel": "@vel"}],
    ["call", "flipVelocityUpdate", {"vel": "@vel", "velOld": "@velOld", "flags": "@flags", "parts": "@pp", "partVel": "@pVel", "flipRatio": 0.97}],
]
CASES["flip01_2d"] = _flip_case([32, 32, 1], 2, [["method", "flags", "initDomain", {"boundaryWidth": 0}],
    ["shape_ls", "Box", {"p0": V(0, 0, 0), "p1": V(12.8, 19.2, 1)}, "phi"], ["method", "flags", "updateFromLevelset", {"levelset": "@phi"}],
    _SAMPLE2, ["repeat", 20, _FLIP01_STEP]], ["vel", "flags", "pp", "pVel"], exact=False,
    abs_tol={"vel": 2e-3, "flags": 0, "pp": 2e-3, "pVel": 2e-3})
CASES["flip01_3d"] = _flip_case([20, 20, 20], 3, [["method", "flags", "initDomain", {"boundaryWidth": 0}],
    ["shape_ls", "Box", {"p0": V(0, 0, 0), "p1": V(8, 12, 20)}, "phi"], ["method", "flags", "updateFromLevelset", {"levelset": "@phi"}],
    ["call", "sampleFlagsWithParticles", {"flags": "@flags", "parts": "@pp", "discretization": 2, "randomness": 0.2}],
    ["repeat", 10, _FLIP01_STEP]], ["vel", "flags", "pp", "pVel"], exact=False,
    abs_tol={"vel": 2e-3, "flags": 0, "pp": 2e-3, "pVel": 2e-3})


# ---- round 2: the kernels the bench times, pinned directly -----------------------------------------
# grids with nx % 4 == 0 take the 16-bit row-code SpMV (k_cg_spmv_dot_code) and the level-0 wavefront legs
# (k_mg_down0 / k_mg_up0w) that the 18-wide family cases above cannot reach
S4 = [32, 24, 20]
CASES["solve4_mgstatic_tight"] = _solve(dict(_TIGHT, preconditioner="$PcMGStatic"), seed=201, size=S4)
CASES["solve4_mgdynamic_tight"] = _solve(dict(_TIGHT, preconditioner="$PcMGDynamic"), seed=202, size=S4)
CASES["solve4_mic_tight"] = _solve(dict(_TIGHT), seed=203, size=S4)
CASES["solve4_liquid_gf"] = _solve({"cgAccuracy": 1e-6, "cgMaxIterFac": 99, "phi": "@phi", "gfClamp": 0.01}, seed=204, size=S4, setup=[], wb=True,
                                   load=("vel", "flags", "phi"), inputs={"flags": ["gen_flags_liquid"], "phi": ["gen_smooth", 205, 1.5]}, grids={"phi": "RealGrid"})
CASES["solve4_liquid_mg"] = _solve({"cgAccuracy": 1e-6, "cgMaxIterFac": 99, "preconditioner": "$PcMGDynamic"}, seed=206, size=S4, setup=[],
                                   load=("vel", "flags"), inputs={"flags": ["gen_flags_liquid"]})
CASES["solve4_liquid_mgstatic_gf"] = _solve({"cgAccuracy": 1e-6, "cgMaxIterFac": 99, "preconditioner": "$PcMGStatic", "phi": "@phi", "gfClamp": 0.01}, seed=207,
                                            size=S4, setup=[], load=("vel", "flags", "phi"),
                                            inputs={"flags": ["gen_flags_liquid"], "phi": ["gen_smooth", 208, 1.5]}, grids={"phi": "RealGrid"})
# PcMIC with the DEFAULT iteration budget (cgMaxIterFac 1.5 -> maxIter 33 on the 18x22x14 grid) and a tight accuracy
CASES["solve_mic_defaultiter"] = _solve({"cgAccuracy": 1e-6, "zeroPressureFixing": True}, seed=209)
# 48 x 40 x 32: three multigrid levels below level 0, obstacle, both MG modes
S5 = [48, 40, 32]
CASES["solve5_mgstatic_tight"] = _solve(dict(_TIGHT, preconditioner="$PcMGStatic"), seed=210, size=S5)

# ---- resetOutflow, open-boundary plume loop (extforces.cpp:134-161, tools/tests/test_2011_plume3d_open.py) -------------------
_OPENALL = [["call", "setOpenBound", {"flags": "@flags", "bWidth": 1, "openBound": "xXyYzZ", "type": 20}]]
CASES["reset_outflow"] = {"size": S3, "dim": 3, "dt": 1.0, "grids": dict(BASEGRIDS, phi="RealGrid"), "load": ["density", "phi"],
                          "inputs": {"density": ["gen_real", 220], "phi": ["gen_smooth", 221, 1.5]},
                          "ops": [["method", "flags", "initDomain", {"boundaryWidth": 1}], ["method", "flags", "fillGrid", {}]] + OBST3 + _OPENALL + [
                              ["call", "resetOutflow", {"flags": "@flags", "phi": "@phi", "real": "@density"}]],
                          "save": ["flags", "density", "phi"], "exact": True}
CASES["reset_outflow_2d"] = {"size": S2, "dim": 2, "dt": 1.0, "grids": dict(BASEGRIDS), "load": ["density"],
                             "inputs": {"density": ["gen_real", 222]},
                             "ops": DOMAIN + [["call", "setOpenBound", {"flags": "@flags", "bWidth": 0, "openBound": "xY", "type": 20}],
                                              ["call", "resetOutflow", {"flags": "@flags", "real": "@density"}]],
                             "save": ["flags", "density"], "exact": True}
_OPEN_PLUME_STEP = [
    ["call", "densityInflow", {"flags": "@flags", "density": "@density", "noise": "@nz", "shape": "@src", "scale": 1, "sigma": 0.5}],
    ["call", "advectSemiLagrange", {"flags": "@flags", "vel": "@vel", "grid": "@density", "order": 2, "clampMode": 1}],
    ["call", "resetOutflow", {"flags": "@flags", "real": "@density"}],
    ["call", "advectSemiLagrange", {"flags": "@flags", "vel": "@vel", "grid": "@vel", "order": 2, "clampMode": 1}],
    ["call", "setWallBcs", {"flags": "@flags", "vel": "@vel"}],
    ["call", "addBuoyancy", {"density": "@density", "vel": "@vel", "gravity": V(0, -5e-2, 0), "flags": "@flags"}],
    ["call", "solvePressure", {"flags": "@flags", "vel": "@vel", "pressure": "@pressure"}],
    ["call", "setWallBcs", {"flags": "@flags", "vel": "@vel"}],
]
_OPEN_PLUME_SETUP = [["object", "nz", "NoiseField", {"loadFromFile": True}, dict(_NZ_PLUME, valScale=1)]] + DOMAIN + _OPENALL + [
    ["object", "src", "Cylinder", {"center": V(12, 3.0, 12), "radius": 3.36, "z": V(0, 0.6, 0)}]]
CASES["plume3d_open_mic"] = {"size": [24, 30, 24], "dim": 3, "dt": 0.5, "grids": dict(BASEGRIDS), "load": [], "inputs": {},
                             "ops": _OPEN_PLUME_SETUP + [["repeat", 15, _OPEN_PLUME_STEP]], "save": ["density", "vel", "flags"], "exact": False,
                             "abs_tol": {"density": 1e-3, "vel": 5e-3, "flags": 0}}
_OPEN_PLUME_STEP_MG = [op if op[1] != "solvePressure" else ["call", "solvePressure", dict(op[2], preconditioner="$PcMGStatic")] for op in _OPEN_PLUME_STEP]
CASES["plume3d_open_mg"] = dict(CASES["plume3d_open_mic"], ops=_OPEN_PLUME_SETUP + [["repeat", 15, _OPEN_PLUME_STEP_MG]])

# ---- grid arithmetic the scenes use between plugins (grid.cpp:367-420) ------------------------------------------------------
CASES["grid_arith"] = {"size": S3, "dim": 3, "dt": 1.0, "grids": dict(BASEGRIDS, vec="VecGrid"), "load": ["density", "pressure", "vel", "vec"],
                       "inputs": {"density": ["gen_real", 230, -1.0, 1.0], "pressure": ["gen_smooth", 231, 2.0], "vel": ["gen_vel", 232, 1.0, 3], "vec": ["gen_vec", 233]},
                       "ops": [["method", "density", "addScaled", {"a": "@pressure", "factor": 0.37}],
                               ["method", "density", "multConst", {"s": -1.7}],
                               ["method", "pressure", "add", {"a": "@density"}],
                               ["method", "pressure", "sub", {"a": "@density"}],
                               ["method", "vel", "addScaled", {"a": "@vec", "factor": V(0.3, -1.1, 2.5)}],
                               ["method", "vec", "multConst", {"s": V(1.5, 0.25, -3.0)}]],
                       "save": ["density", "pressure", "vel", "vec"], "exact": True}
