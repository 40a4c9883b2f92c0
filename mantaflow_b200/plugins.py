"""PYTHON() plugin functions of the hot path, with the reference's keyword names and defaults.

advectSemiLagrange  source/plugin/advection.cpp:443
setWallBcs / addBuoyancy / addGravity / setInflowBcs / setOpenBound / resetOutflow  source/plugin/extforces.cpp
computePressureRhs / solvePressureSystem / correctVelocity / solvePressure / releaseMG  source/plugin/pressure.cpp
cgSolveDiffusion  source/conjugategrad.cpp:350
gridMaxDiff / gridMaxDiffInt / gridMaxDiffVec3  source/grid.cpp:437-467
extrapolateMACSimple / extrapolateMACFromWeight / extrapolateLsSimple / extrapolateVec3Simple  source/fastmarch.cpp:337-553
updateFractions / setObstacleFlags  source/plugin/initplugins.cpp:437,473
"""
import ctypes as C

from . import manta as M
from ._lib import lib, check, MantaError, PressureParams, SolveInfo

__all__ = ["advectSemiLagrange", "setWallBcs", "addBuoyancy", "addGravity", "addGravityNoScale", "setInflowBcs",
           "setOpenBound", "resetOutflow", "computePressureRhs", "solvePressureSystem", "correctVelocity",
           "solvePressure", "releaseMG", "cgSolveDiffusion", "gridMaxDiff", "gridMaxDiffInt", "gridMaxDiffVec3",
           "getSolveInfo", "extrapolateMACSimple", "extrapolateMACFromWeight", "extrapolateLsSimple", "extrapolateVec3Simple",
           "updateFractions", "setObstacleFlags", "densityInflow", "addNoise", "getComponent", "setComponent",
           "sampleFlagsWithParticles", "mapPartsToMAC", "markFluidCells", "mapMACToParts", "flipVelocityUpdate", "GridCg",
           "copyArrayToGridReal", "copyGridToArrayReal", "copyArrayToGridFlag", "copyGridToArrayFlag", "copyArrayToGridLevelset",
           "copyGridToArrayLevelset", "copyArrayToGridVec3", "copyGridToArrayVec3", "copyArrayToGridMAC", "copyGridToArrayMAC",
           "copyArrayToPdataInt", "copyPdataToArrayInt", "copyArrayToPdataReal", "copyPdataToArrayReal", "copyArrayToPdataVec3",
           "copyPdataToArrayVec3"]


def _h(g):
    return g._h if g is not None else None


def _req(g, cls, name, fn):
    if g is None:
        raise MantaError("Argument '%s' is not defined." % name)
    if not isinstance(g, cls):
        raise MantaError("%s: can't convert argument '%s' to type '%s'" % (fn, name, cls.__name__))
    return g


@M._plugin
def advectSemiLagrange(flags=None, vel=None, grid=None, order=1, strength=1.0, orderSpace=1, openBounds=False,
                       boundaryWidth=-1, clampMode=2, orderTrace=1, **kwargs):
    M._checkArgs("advectSemiLagrange", kwargs, ())
    _req(flags, M.FlagGrid, "flags", "advectSemiLagrange")
    _req(vel, M.MACGrid, "vel", "advectSemiLagrange")
    _req(grid, M.GridBase, "grid", "advectSemiLagrange")
    if boundaryWidth != -1 or openBounds:
        M.mantaMsg("Warning: boundaryWidth and openBounds parameters in AdvectSemiLagrange plugin are deprecated "
                   "(and have no more effect), please remove.", 0)
    if isinstance(grid, M.MACGrid) and orderSpace != 1:
        M.mantaMsg("Warning higher order for MAC grids not yet implemented...", 1)
    check(lib().mf_advect_semi_lagrange(flags._h, vel._h, grid._h, int(order), float(strength), int(orderSpace),
                                        int(clampMode), int(orderTrace)))


@M._plugin
def setWallBcs(flags=None, vel=None, obvel=None, fractions=None, phiObs=None, boundaryWidth=0, **kwargs):
    M._checkArgs("setWallBcs", kwargs, ())
    _req(flags, M.FlagGrid, "flags", "setWallBcs")
    _req(vel, M.MACGrid, "vel", "setWallBcs")
    check(lib().mf_set_wall_bcs(flags._h, vel._h, _h(obvel), _h(fractions), _h(phiObs), int(boundaryWidth)))


@M._plugin
def addBuoyancy(flags=None, density=None, vel=None, gravity=None, coefficient=1.0, scale=True, **kwargs):
    M._checkArgs("addBuoyancy", kwargs, ())
    _req(flags, M.FlagGrid, "flags", "addBuoyancy")
    _req(density, M.RealGrid, "density", "addBuoyancy")
    _req(vel, M.MACGrid, "vel", "addBuoyancy")
    if gravity is None:
        raise MantaError("Argument 'gravity' is not defined.")
    g = M._v3(gravity)
    check(lib().mf_add_buoyancy(flags._h, density._h, vel._h, float(g.x), float(g.y), float(g.z), float(coefficient), int(bool(scale))))


@M._plugin
def addGravity(flags=None, vel=None, gravity=None, exclude=None, scale=True, **kwargs):
    M._checkArgs("addGravity", kwargs, ())
    _req(flags, M.FlagGrid, "flags", "addGravity")
    _req(vel, M.MACGrid, "vel", "addGravity")
    if gravity is None:
        raise MantaError("Argument 'gravity' is not defined.")
    g = M._v3(gravity)
    check(lib().mf_add_gravity(flags._h, vel._h, float(g.x), float(g.y), float(g.z), _h(exclude), int(bool(scale))))


def addGravityNoScale(flags=None, vel=None, gravity=None, exclude=None, **kwargs):
    addGravity(flags=flags, vel=vel, gravity=gravity, exclude=exclude, scale=False, **kwargs)


@M._plugin
def setInflowBcs(vel=None, dir="", value=None, **kwargs):
    M._checkArgs("setInflowBcs", kwargs, ())
    _req(vel, M.MACGrid, "vel", "setInflowBcs")
    v = M._v3(value)
    check(lib().mf_set_inflow_bcs(vel._h, dir.encode(), float(v.x), float(v.y), float(v.z)))


@M._plugin
def setOpenBound(flags=None, bWidth=None, openBound="", type=M.FlagOutflow | M.FlagEmpty, **kwargs):
    M._checkArgs("setOpenBound", kwargs, ())
    _req(flags, M.FlagGrid, "flags", "setOpenBound")
    check(lib().mf_flags_set_open_bound(flags._h, int(bWidth), openBound.encode(), int(type)))


@M._plugin
def resetOutflow(flags=None, phi=None, parts=None, real=None, index=None, indexSys=None, **kwargs):
    M._checkArgs("resetOutflow", kwargs, ())
    _req(flags, M.FlagGrid, "flags", "resetOutflow")
    if parts is not None:
        raise MantaError("resetOutflow: particle systems are not part of the device grid path")
    check(lib().mf_reset_outflow(flags._h, _h(phi), _h(real)))


_PRESSURE_KW = ("cgAccuracy", "phi", "perCellCorr", "fractions", "obvel", "gfClamp", "cgMaxIterFac", "precondition",
                "preconditioner", "enforceCompatibility", "useL2Norm", "zeroPressureFixing", "curv", "surfTens", "retRhs")


def _params(fn, kw, allowObvel=True, allowRetRhs=False):
    p = PressureParams()
    lib().mf_pressure_params_default(C.byref(p))
    keep = []
    for k, v in kw.items():
        if k not in _PRESSURE_KW or (k == "obvel" and not allowObvel) or (k == "retRhs" and not allowRetRhs):
            raise MantaError("Argument '%s' unknown" % k)
        if k in ("phi", "perCellCorr", "curv", "retRhs"):
            if v is not None:
                _req(v, M.RealGrid, k, fn); keep.append(v)
            setattr(p, k, _h(v))
        elif k in ("fractions", "obvel"):
            if v is not None:
                _req(v, M.MACGrid, k, fn); keep.append(v)
            setattr(p, k, _h(v))
        elif k in ("precondition", "enforceCompatibility", "useL2Norm", "zeroPressureFixing"):
            setattr(p, k, int(bool(v)))
        elif k == "preconditioner":
            p.preconditioner = int(v)
        else:
            setattr(p, k, float(v))
    pc = p.preconditioner if p.precondition else M.PcNone
    if pc == M.PcNone and not M.allowPcNone and fn in ("solvePressure", "solvePressureSystem"):
        # setICPreconditioner(PC_None, ...) asserts in the reference (conjugategrad.cpp:312, reached from pressure.cpp:417)
        raise MantaError("GridCg<APPLYMAT>::setICPreconditioner: Invalid method specified.")
    return p, keep


def _publish(info, accuracy=None):
    converged = accuracy is None or info.resNorm < accuracy
    M.lastSolveInfo = dict(iterations=info.iterations, resNorm=info.resNorm, maxIter=info.maxIter,
                           fixedCell=info.fixedCell, mgLevels=info.mgLevels, mgRebuilt=info.mgRebuilt, converged=converged)
    M.mantaMsg("FluidSolver::solvePressure done. Iterations:%d, residual:%g" % (info.iterations, info.resNorm), 2)
    if not converged:     # the reference stops silently at maxIter (pressure.cpp:430-436); say so at its debug level
        M.mantaMsg("FluidSolver::solvePressure stopped at maxIter=%d with residual %g >= cgAccuracy %g" % (info.maxIter, info.resNorm, accuracy), 2)


def getSolveInfo():
    """Iterations / residual of the last pressure solve (the reference prints them at debug level 2, pressure.cpp:442)."""
    return M.lastSolveInfo


@M._plugin
def computePressureRhs(rhs=None, vel=None, pressure=None, flags=None, **kwargs):
    _req(rhs, M.RealGrid, "rhs", "computePressureRhs"); _req(vel, M.MACGrid, "vel", "computePressureRhs")
    _req(pressure, M.RealGrid, "pressure", "computePressureRhs"); _req(flags, M.FlagGrid, "flags", "computePressureRhs")
    p, keep = _params("computePressureRhs", kwargs)
    check(lib().mf_compute_pressure_rhs(rhs._h, vel._h, pressure._h, flags._h, C.byref(p)))


@M._plugin
def solvePressureSystem(rhs=None, vel=None, pressure=None, flags=None, **kwargs):
    _req(rhs, M.RealGrid, "rhs", "solvePressureSystem"); _req(vel, M.MACGrid, "vel", "solvePressureSystem")
    _req(pressure, M.RealGrid, "pressure", "solvePressureSystem"); _req(flags, M.FlagGrid, "flags", "solvePressureSystem")
    p, keep = _params("solvePressureSystem", kwargs, allowObvel=False)
    info = SolveInfo()
    check(lib().mf_solve_pressure_system(rhs._h, vel._h, pressure._h, flags._h, C.byref(p), C.byref(info)))
    _publish(info, p.cgAccuracy)


@M._plugin
def correctVelocity(vel=None, pressure=None, flags=None, **kwargs):
    _req(vel, M.MACGrid, "vel", "correctVelocity"); _req(pressure, M.RealGrid, "pressure", "correctVelocity")
    _req(flags, M.FlagGrid, "flags", "correctVelocity")
    p, keep = _params("correctVelocity", kwargs, allowObvel=False)
    check(lib().mf_correct_velocity(vel._h, pressure._h, flags._h, C.byref(p)))


@M._plugin
def solvePressure(vel=None, pressure=None, flags=None, **kwargs):
    _req(vel, M.MACGrid, "vel", "solvePressure"); _req(pressure, M.RealGrid, "pressure", "solvePressure")
    _req(flags, M.FlagGrid, "flags", "solvePressure")
    p, keep = _params("solvePressure", kwargs, allowRetRhs=True)
    info = SolveInfo()
    check(lib().mf_solve_pressure(vel._h, pressure._h, flags._h, C.byref(p), C.byref(info)))
    _publish(info, p.cgAccuracy)


def releaseMG(solver=None, **kwargs):
    M._checkArgs("releaseMG", kwargs, ())
    if solver is not None:
        check(lib().mf_release_mg(solver._h))


@M._plugin
def cgSolveDiffusion(flags=None, grid=None, alpha=0.25, cgMaxIterFac=1.0, cgAccuracy=1e-4, **kwargs):
    M._checkArgs("cgSolveDiffusion", kwargs, ())
    _req(flags, M.FlagGrid, "flags", "cgSolveDiffusion"); _req(grid, M.GridBase, "grid", "cgSolveDiffusion")
    check(lib().mf_cg_solve_diffusion(flags._h, grid._h, float(alpha), float(cgMaxIterFac), float(cgAccuracy)))


class GridCg(object):
    """GridCg<ApplyMatrix> (conjugategrad.h:65-116) for C++-style clients of the solver: the reference does not expose this class to
    Python, its plugins drive it directly (pressure.cpp:372-452, waves.cpp:129, vortexplugins.cpp:279, fluidguiding.cpp:276).  Same
    method names and semantics; the grids stay owned by the caller."""
    PC_None, PC_ICP, PC_mICP, PC_MGP = 0, 1, 2, 3

    def __init__(self, dst, rhs, residual, search, flags, tmp, A0, Ai, Aj, Ak):
        self._keep = (dst, rhs, residual, search, flags, tmp, A0, Ai, Aj, Ak)
        self._h = lib().mf_cg_create(dst._h, rhs._h, residual._h, search._h, flags._h, tmp._h, A0._h, Ai._h, Aj._h, Ak._h)
        if not self._h:
            check(1)

    def __del__(self):
        try:
            if getattr(self, "_h", None):
                lib().mf_cg_destroy(self._h)
                self._h = None
        except Exception:
            pass

    def setICPreconditioner(self, method, A0=None, Ai=None, Aj=None, Ak=None): check(lib().mf_cg_set_ic_preconditioner(self._h, int(method)))
    def setMGPreconditioner(self, method, MG=None, rebuild=False): check(lib().mf_cg_set_mg_preconditioner(self._h, int(method), 1 if rebuild else 0))
    def setAccuracy(self, a): check(lib().mf_cg_set_accuracy(self._h, float(a)))
    def setUseL2Norm(self, b): check(lib().mf_cg_set_use_l2_norm(self._h, 1 if b else 0))
    def forceReinit(self): check(lib().mf_cg_force_reinit(self._h))

    def iterate(self):
        r = C.c_int()
        check(lib().mf_cg_iterate(self._h, C.byref(r)))
        return bool(r.value)

    def solve(self, maxIter): check(lib().mf_cg_solve(self._h, int(maxIter)))

    def _get(self, i):
        v = [C.c_float() for _ in range(4)]
        check(lib().mf_cg_get(self._h, *[C.byref(x) for x in v]))
        return v[i].value

    def getSigma(self): return self._get(0)
    def getIterations(self): return self._get(1)
    def getResNorm(self): return self._get(2)
    def getAccuracy(self): return self._get(3)


def _maxDiff(g1, g2):
    out = C.c_double()
    check(lib().mf_grid_max_diff(g1._h, g2._h, C.byref(out)))
    return out.value


def gridMaxDiff(g1, g2): return _maxDiff(g1, g2)
def gridMaxDiffInt(g1, g2): return _maxDiff(g1, g2)
def gridMaxDiffVec3(g1, g2): return _maxDiff(g1, g2)


# ---- simple extrapolation (source/fastmarch.cpp) ---------------------------------------------------------------------
@M._plugin
def extrapolateMACSimple(flags=None, vel=None, distance=4, phiObs=None, intoObs=False, **kwargs):
    M._checkArgs("extrapolateMACSimple", kwargs, ())
    _req(flags, M.FlagGrid, "flags", "extrapolateMACSimple")
    _req(vel, M.MACGrid, "vel", "extrapolateMACSimple")
    check(lib().mf_extrapolate_mac_simple(flags._h, vel._h, int(distance), _h(phiObs), 1 if intoObs else 0))


@M._plugin
def extrapolateMACFromWeight(vel=None, weight=None, distance=2, **kwargs):
    M._checkArgs("extrapolateMACFromWeight", kwargs, ())
    _req(vel, M.MACGrid, "vel", "extrapolateMACFromWeight")
    _req(weight, M.VecGrid, "weight", "extrapolateMACFromWeight")
    check(lib().mf_extrapolate_mac_from_weight(vel._h, weight._h, int(distance)))


@M._plugin
def extrapolateLsSimple(phi=None, distance=4, inside=False, include_walls=False, **kwargs):
    M._checkArgs("extrapolateLsSimple", kwargs, ())
    _req(phi, M.RealGrid, "phi", "extrapolateLsSimple")
    check(lib().mf_extrapolate_ls_simple(phi._h, int(distance), 1 if inside else 0, 1 if include_walls else 0))


@M._plugin
def extrapolateVec3Simple(vel=None, phi=None, distance=4, inside=False, **kwargs):
    M._checkArgs("extrapolateVec3Simple", kwargs, ())
    _req(vel, M.VecGrid, "vel", "extrapolateVec3Simple")
    _req(phi, M.RealGrid, "phi", "extrapolateVec3Simple")
    check(lib().mf_extrapolate_vec3_simple(vel._h, phi._h, int(distance), 1 if inside else 0))


# ---- obstacle fractions / flags from levelsets (source/plugin/initplugins.cpp) -----------------------------------------
@M._plugin
def updateFractions(flags=None, phiObs=None, fractions=None, boundaryWidth=0, fracThreshold=0.01, **kwargs):
    M._checkArgs("updateFractions", kwargs, ())
    _req(flags, M.FlagGrid, "flags", "updateFractions")
    _req(phiObs, M.RealGrid, "phiObs", "updateFractions")
    _req(fractions, M.MACGrid, "fractions", "updateFractions")
    check(lib().mf_update_fractions(flags._h, phiObs._h, fractions._h, int(boundaryWidth), float(fracThreshold)))


@M._plugin
def setObstacleFlags(flags=None, phiObs=None, fractions=None, phiOut=None, phiIn=None, boundaryWidth=1, **kwargs):
    M._checkArgs("setObstacleFlags", kwargs, ())
    _req(flags, M.FlagGrid, "flags", "setObstacleFlags")
    _req(phiObs, M.RealGrid, "phiObs", "setObstacleFlags")
    check(lib().mf_set_obstacle_flags(flags._h, phiObs._h, _h(fractions), _h(phiOut), _h(phiIn), int(boundaryWidth)))


# ---- noise-driven sources (source/plugin/initplugins.cpp:26-50) -------------------------------------------------------
@M._plugin
def densityInflow(flags=None, density=None, noise=None, shape=None, scale=1.0, sigma=0, **kwargs):
    M._checkArgs("densityInflow", kwargs, ())
    _req(flags, M.FlagGrid, "flags", "densityInflow")
    _req(density, M.RealGrid, "density", "densityInflow")
    if not isinstance(noise, M.NoiseField):
        raise MantaError("densityInflow: argument 'noise' is missing or not a NoiseField")
    if not isinstance(shape, M.Shape):
        raise MantaError("densityInflow: argument 'shape' is missing or not a Shape")
    p = noise._params()
    d = shape._struct()
    check(lib().mf_density_inflow(flags._h, density._h, noise._h, C.byref(p), C.byref(d), float(scale), float(sigma),
                                  float(flags.parent.timeTotal)))


@M._plugin
def addNoise(flags=None, density=None, noise=None, sdf=None, scale=1.0, **kwargs):
    M._checkArgs("addNoise", kwargs, ())
    _req(flags, M.FlagGrid, "flags", "addNoise")
    _req(density, M.RealGrid, "density", "addNoise")
    if not isinstance(noise, M.NoiseField):
        raise MantaError("addNoise: argument 'noise' is missing or not a NoiseField")
    p = noise._params()
    check(lib().mf_add_noise(flags._h, density._h, noise._h, C.byref(p), _h(sdf), float(scale), float(flags.parent.timeTotal)))


# ---- Real <-> Vec3 component transfer (source/grid.cpp:741-751) ---------------------------------------------------------
@M._plugin
def getComponent(source=None, target=None, component=None, **kwargs):
    M._checkArgs("getComponent", kwargs, ())
    _req(source, M.VecGrid, "source", "getComponent")
    _req(target, M.RealGrid, "target", "getComponent")
    check(lib().mf_grid_get_component(source._h, target._h, int(component)))


@M._plugin
def setComponent(source=None, target=None, component=None, **kwargs):
    M._checkArgs("setComponent", kwargs, ())
    _req(source, M.RealGrid, "source", "setComponent")
    _req(target, M.VecGrid, "target", "setComponent")
    check(lib().mf_grid_set_component(source._h, target._h, int(component)))


# ---- FLIP (source/plugin/flip.cpp) --------------------------------------------------------------------------------------
def _parts(p, name, fn):
    if not isinstance(p, M.BasicParticleSystem):
        raise MantaError("%s: argument '%s' is missing or not a BasicParticleSystem" % (fn, name))
    return p


def _pvec(pd, parts, name, fn):
    if not isinstance(pd, M.PdataVec3) or pd.parts is not parts:
        raise MantaError("%s: argument '%s' is missing or not a PdataVec3 of 'parts'" % (fn, name))
    return pd._index


def _noPtype(fn, ptype):
    if ptype is not None:
        raise MantaError("%s: ptype / exclude are not supported" % fn)


@M._plugin
def sampleFlagsWithParticles(flags=None, parts=None, discretization=None, randomness=None, **kwargs):
    M._checkArgs("sampleFlagsWithParticles", kwargs, ())
    _req(flags, M.FlagGrid, "flags", "sampleFlagsWithParticles")
    _parts(parts, "parts", "sampleFlagsWithParticles")
    check(lib().mf_sample_flags_with_particles(flags._h, parts._h, int(discretization), float(randomness)))


@M._plugin
def mapPartsToMAC(flags=None, vel=None, velOld=None, parts=None, partVel=None, weight=None, ptype=None, exclude=0, **kwargs):
    M._checkArgs("mapPartsToMAC", kwargs, ())
    _noPtype("mapPartsToMAC", ptype)
    _req(flags, M.FlagGrid, "flags", "mapPartsToMAC")
    _req(vel, M.MACGrid, "vel", "mapPartsToMAC")
    _req(velOld, M.MACGrid, "velOld", "mapPartsToMAC")
    _parts(parts, "parts", "mapPartsToMAC")
    check(lib().mf_map_parts_to_mac(flags._h, vel._h, velOld._h, parts._h, _pvec(partVel, parts, "partVel", "mapPartsToMAC"), _h(weight)))


@M._plugin
def markFluidCells(parts=None, flags=None, phiObs=None, ptype=None, exclude=0, **kwargs):
    M._checkArgs("markFluidCells", kwargs, ())
    _noPtype("markFluidCells", ptype)
    _req(flags, M.FlagGrid, "flags", "markFluidCells")
    _parts(parts, "parts", "markFluidCells")
    check(lib().mf_mark_fluid_cells(parts._h, flags._h, _h(phiObs)))


@M._plugin
def mapMACToParts(flags=None, vel=None, parts=None, partVel=None, ptype=None, exclude=0, **kwargs):
    M._checkArgs("mapMACToParts", kwargs, ())
    _noPtype("mapMACToParts", ptype)
    _req(vel, M.MACGrid, "vel", "mapMACToParts")
    _parts(parts, "parts", "mapMACToParts")
    check(lib().mf_map_mac_to_parts(_h(flags), vel._h, parts._h, _pvec(partVel, parts, "partVel", "mapMACToParts")))


@M._plugin
def flipVelocityUpdate(flags=None, vel=None, velOld=None, parts=None, partVel=None, flipRatio=None, ptype=None, exclude=0, **kwargs):
    M._checkArgs("flipVelocityUpdate", kwargs, ())
    _noPtype("flipVelocityUpdate", ptype)
    _req(vel, M.MACGrid, "vel", "flipVelocityUpdate")
    _req(velOld, M.MACGrid, "velOld", "flipVelocityUpdate")
    _parts(parts, "parts", "flipVelocityUpdate")
    check(lib().mf_flip_velocity_update(_h(flags), vel._h, velOld._h, parts._h, _pvec(partVel, parts, "partVel", "flipVelocityUpdate"), float(flipRatio)))


# ---- numpy bridge (source/plugin/numpyconvert.cpp:145-222): C-order arrays, z slowest; the grid side is device memory -------------
def _np_to_grid(fn, cls, ncomp, int_grid=False):
    import numpy as np

    def plugin(source=None, target=None, **kwargs):
        M._checkArgs(fn, kwargs, ())
        if source is None:
            raise MantaError("Argument 'source' is not defined.")
        _req(target, cls, "target", fn)
        a = np.asarray(source)
        nx, ny, nz = target.parent._size
        if a.size != nx * ny * nz * ncomp:
            raise MantaError("The size of the numpy array doesn't match the size of the grid!")
        if a.ndim >= 3 and tuple(a.shape[:3]) != (nz, ny, nx):
            raise MantaError("The dimensions of source grid (%d, %d, %d) and target numpy array (%d, %d, %d) do not match!"
                             % (nx, ny, nz, a.shape[2], a.shape[1], a.shape[0]))
        ok = ("i", "u") if int_grid else ("f",)
        if a.dtype.kind not in ok or a.dtype.itemsize not in (4, 8):
            raise MantaError("unknown/unsupported type of Numpy array")
        target.from_numpy(a)           # converts to the grid's element type like the N_DOUBLE / N_INT branches
    plugin.__name__ = fn
    return M._plugin(plugin)


def _grid_to_np(fn, cls, ncomp):
    import numpy as np

    def plugin(source=None, target=None, **kwargs):
        M._checkArgs(fn, kwargs, ())
        _req(source, cls, "source", fn)
        if not isinstance(target, np.ndarray):
            raise MantaError("Argument 'target' is not defined.")
        nx, ny, nz = source.parent._size
        if target.size != nx * ny * nz * ncomp:
            raise MantaError("The size of the numpy array doesn't match the size of the grid!")
        if target.ndim >= 3 and tuple(target.shape[:3]) != (nz, ny, nx):
            raise MantaError("The dimensions of source grid (%d, %d, %d) and target numpy array (%d, %d, %d) do not match!"
                             % (nx, ny, nz, target.shape[2], target.shape[1], target.shape[0]))
        if not target.flags["C_CONTIGUOUS"] or not target.flags["WRITEABLE"]:
            raise MantaError("numpy array must be C-contiguous and writeable")
        target.reshape(source._shape())[...] = source.to_numpy()
    plugin.__name__ = fn
    return M._plugin(plugin)


copyArrayToGridReal = _np_to_grid("copyArrayToGridReal", M.RealGrid, 1)
copyGridToArrayReal = _grid_to_np("copyGridToArrayReal", M.RealGrid, 1)
copyArrayToGridLevelset = _np_to_grid("copyArrayToGridLevelset", M.LevelsetGrid, 1)
copyGridToArrayLevelset = _grid_to_np("copyGridToArrayLevelset", M.LevelsetGrid, 1)
copyArrayToGridFlag = _np_to_grid("copyArrayToGridFlag", M.FlagGrid, 1, int_grid=True)
copyGridToArrayFlag = _grid_to_np("copyGridToArrayFlag", M.FlagGrid, 1)
copyArrayToGridVec3 = _np_to_grid("copyArrayToGridVec3", M.VecGrid, 3)
copyGridToArrayVec3 = _grid_to_np("copyGridToArrayVec3", M.VecGrid, 3)
copyArrayToGridMAC = _np_to_grid("copyArrayToGridMAC", M.MACGrid, 3)
copyGridToArrayMAC = _grid_to_np("copyGridToArrayMAC", M.MACGrid, 3)


def _np_to_pdata(fn, cls):
    import numpy as np

    def plugin(source=None, target=None, **kwargs):
        M._checkArgs(fn, kwargs, ())
        _req(target, cls, "target", fn)
        a = np.asarray(source)
        if a.size != target.parts.size() * (3 if cls._ncomp == 3 else 1):
            raise MantaError("The size of the numpy array doesn't match the size of the pdata field!")
        target.from_numpy(a)
    plugin.__name__ = fn
    return M._plugin(plugin)


def _pdata_to_np(fn, cls):
    def plugin(source=None, target=None, **kwargs):
        M._checkArgs(fn, kwargs, ())
        _req(source, cls, "source", fn)
        if target is None or target.size != source.parts.size() * (3 if cls._ncomp == 3 else 1):
            raise MantaError("The size of the numpy array doesn't match the size of the pdata field!")
        target.reshape(-1)[...] = source.to_numpy().reshape(-1)
    plugin.__name__ = fn
    return M._plugin(plugin)


copyArrayToPdataInt = _np_to_pdata("copyArrayToPdataInt", M.PdataInt)
copyPdataToArrayInt = _pdata_to_np("copyPdataToArrayInt", M.PdataInt)
copyArrayToPdataReal = _np_to_pdata("copyArrayToPdataReal", M.PdataReal)
copyPdataToArrayReal = _pdata_to_np("copyPdataToArrayReal", M.PdataReal)
copyArrayToPdataVec3 = _np_to_pdata("copyArrayToPdataVec3", M.PdataVec3)
copyPdataToArrayVec3 = _pdata_to_np("copyPdataToArrayVec3", M.PdataVec3)
