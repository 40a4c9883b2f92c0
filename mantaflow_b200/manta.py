"""Host-side mirror of the reference's Python scene API for the Eulerian grid hot path.

`from mantaflow_b200.manta import *` gives the names a Mantaflow scene uses for the path
(`Solver`/`FluidSolver`, `FlagGrid`, `MACGrid`, `RealGrid`, `vec3`, `advectSemiLagrange`, `setWallBcs`,
`addBuoyancy`, `solvePressure`, constants of source/python/defines.py ...) with the reference's keyword
names, defaults and error behaviour (unknown keyword -> RuntimeError "Argument 'x' unknown",
source/pwrapper/pconvert.cpp:460-474).  Every call goes through the C-ABI of libmantacuda.so; grids live
in device memory and only cross to the host in `to_numpy`/`save`/reductions.
"""
import ctypes as C
import gzip
import struct
import time as _time

import numpy as np

from ._lib import lib, check, MantaError, PressureParams, SolveInfo

# ---- module constants (source/python/defines.py:24-50, pwrapper/registry.cpp:398-420) ---------------
FlagFluid, FlagObstacle, FlagEmpty, FlagInflow, FlagOutflow, FlagOpen, FlagStick, FlagReserved = 1, 2, 4, 8, 16, 32, 64, 256
TypeFluid, TypeObstacle, TypeEmpty, TypeInflow, TypeOutflow, TypeStick, TypeReserved = 1, 2, 4, 8, 16, 64, 256
IntEuler, IntRK2, IntRK4 = 0, 1, 2
PcNone, PcMIC, PcMGDynamic, PcMGStatic = 0, 1, 2, 3
GUI = False
DEBUG = False
MT = True
DOUBLEPRECISION = False
CUDA = True

_TYPE_REAL, _TYPE_INT, _TYPE_VEC3, _TYPE_MAC, _TYPE_LEVELSET, _TYPE_FLAGS = 1, 2, 4, 8, 16, 32

_UNI_HEADER_BYTES = 288


def _unifyGridType(t):
    """unifyGridType (fileio/iogrids.cpp:213-221): real <> levelset, vec3 <> MAC"""
    if t & _TYPE_REAL:
        t |= _TYPE_LEVELSET
    if t & _TYPE_LEVELSET:
        t |= _TYPE_REAL
    if t & _TYPE_VEC3:
        t |= _TYPE_MAC
    if t & _TYPE_MAC:
        t |= _TYPE_VEC3
    return t


_debugLevel = 1
# the reference's solvePressure(preconditioner=PcNone) throws in setICPreconditioner (conjugategrad.cpp:312);
# set allowPcNone = True to run plain CG instead
allowPcNone = False
lastSolveInfo = None


def setDebugLevel(level=1):
    global _debugLevel
    _debugLevel = int(level)


def mantaMsg(out, level=1):
    if level <= _debugLevel:
        print(out)


def printBuildInfo():
    s = lib().mf_build_info().decode()
    mantaMsg("Build info: %s " % s, 1)
    return s


def _f32(v):
    return np.float32(v)


class vec3(object):
    """Python `vec3` (source/pwrapper/pvec3.cpp): three Reals (float32), component-wise arithmetic."""
    __slots__ = ("x", "y", "z")

    def __init__(self, x=0.0, y=None, z=None):
        if isinstance(x, vec3):
            x, y, z = x.x, x.y, x.z
        elif isinstance(x, (tuple, list, np.ndarray)):
            x, y, z = x[0], x[1], x[2]
        elif y is None:
            y = z = x
        elif z is None:
            z = 0.0
        self.x, self.y, self.z = _f32(x), _f32(y), _f32(z)

    def _bin(self, o, op):
        o = o if isinstance(o, vec3) else vec3(o)
        return vec3(op(self.x, o.x), op(self.y, o.y), op(self.z, o.z))

    def __add__(self, o): return self._bin(o, lambda a, b: a + b)
    def __radd__(self, o): return self._bin(o, lambda a, b: b + a)
    def __sub__(self, o): return self._bin(o, lambda a, b: a - b)
    def __rsub__(self, o): return self._bin(o, lambda a, b: b - a)
    def __mul__(self, o): return self._bin(o, lambda a, b: a * b)
    def __rmul__(self, o): return self._bin(o, lambda a, b: b * a)
    def __truediv__(self, o): return self._bin(o, lambda a, b: a / b)
    def __neg__(self): return vec3(-self.x, -self.y, -self.z)
    def __getitem__(self, i): return (self.x, self.y, self.z)[i]
    def __iter__(self): return iter((self.x, self.y, self.z))
    def __eq__(self, o): return isinstance(o, vec3) and (self.x, self.y, self.z) == (o.x, o.y, o.z)
    def __repr__(self): return "[%g,%g,%g]" % (self.x, self.y, self.z)
    def max(self): return max(self.x, self.y, self.z)
    def min(self): return min(self.x, self.y, self.z)


Vec3 = vec3


def _v3(v):
    return v if isinstance(v, vec3) else vec3(v)


def _checkArgs(fn, kwargs, allowed):
    """PbArgs::check (pconvert.cpp:460-474): unknown keyword arguments are an error; `notiming` is implicit."""
    for k in kwargs:
        if k not in allowed and k != "notiming":
            raise MantaError("Argument '%s' unknown" % k)


class _Timing(object):
    """TimingData (source/timing.cpp:23-112): wall-clock per plugin, the stream is synchronised before stop()."""
    def __init__(self):
        self.data = {}
        self.enabled = False

    def add(self, name, ms):
        self.data.setdefault(name, []).append(ms)


_timing = _Timing()


class Timings(object):
    def __init__(self):
        _timing.enabled = True

    def display(self):
        total = 0.0
        for name, v in _timing.data.items():
            ms = sum(v)
            total += ms
            print("[%8.2f ms] %s (%d calls)" % (ms, name, len(v)))
        print("Total : %.2f ms" % total)
        _timing.data.clear()

    def step(self):
        pass


def _plugin(fn):
    def wrapper(*args, **kwargs):
        if _timing.enabled and not kwargs.get("notiming", False):
            solver = None
            for a in list(args) + list(kwargs.values()):
                if isinstance(a, GridBase):
                    solver = a.parent
                    break
            t0 = _time.perf_counter()
            kwargs.pop("notiming", None)
            r = fn(*args, **kwargs)
            if solver is not None:
                solver.sync()
            _timing.add(fn.__name__, (_time.perf_counter() - t0) * 1e3)
            return r
        kwargs.pop("notiming", None)
        return fn(*args, **kwargs)
    wrapper.__name__ = fn.__name__
    wrapper.__doc__ = fn.__doc__
    return wrapper


# ---- multi-GPU: the team of ranks a solver is partitioned over (include/mantacuda.h "multi-GPU") ---------------
class Comm(object):
    """One process per GPU (`Comm.from_env()` under torchrun), or all ranks emulated in this process on one GPU
    (`Comm.emulated(n)`, for testing the partitioned path).  Scenes stay unchanged: `setDefaultComm(comm)` makes every
    Solver created afterwards a z-slab partitioned one."""

    def __init__(self, handle, rank, nranks, emulated):
        self._h, self.rank, self.nranks, self.emulated = handle, rank, nranks, emulated

    @staticmethod
    def emulated(nranks, device=0):
        h = lib().mf_comm_create_emulated(int(nranks), int(device))
        if not h:
            raise MantaError(lib().mf_last_error().decode())
        return Comm(h, 0, int(nranks), True)

    @staticmethod
    def from_env(session=None):
        import os
        rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
        local = int(os.environ.get("LOCAL_RANK", str(rank)))
        session = session or ("job%s" % os.environ.get("MASTER_PORT", "0"))
        h = lib().mf_comm_create(rank, world, local, session.encode())
        if not h:
            raise MantaError(lib().mf_last_error().decode())
        return Comm(h, rank, world, False)

    def destroy(self):
        h, self._h = self._h, None
        if h:
            lib().mf_comm_destroy(h)


_defaultComm = None


def setDefaultComm(comm):
    global _defaultComm
    _defaultComm = comm


# ---- FluidSolver -----------------------------------------------------------------------------------------
class FluidSolver(object):
    """FluidSolver / Solver (source/fluidsolver.h:26-119, fluidsolver.cpp:107-166)."""

    def __init__(self, name="solver", gridSize=None, dim=3, device=0, comm=None, **kwargs):
        _checkArgs("Solver", kwargs, ())
        comm = comm or _defaultComm
        if gridSize is None:
            raise MantaError("Argument 'gridSize' is not defined.")
        gs = _v3(gridSize)
        self.name = name
        nx, ny, nz = int(gs.x), int(gs.y), int(gs.z)
        if dim not in (2, 3):
            raise MantaError("Only 2D and 3D solvers allowed.")
        if dim == 2 and nz != 1:
            raise MantaError("Trying to create 2D solver with size.z != 1")
        self._size = (nx, ny, nz)
        self._dim = dim
        self.comm = comm
        if comm is not None and comm.nranks > 1:
            self._h = lib().mf_solver_create_dist(comm._h, nx, ny, nz, dim)
        else:
            self._h = lib().mf_solver_create(nx, ny, nz, dim, device)
        if not self._h:
            raise MantaError(lib().mf_last_error().decode())
        self._dt = np.float32(1.0)
        self.timeTotal = np.float32(0.0)
        self.frame = 0
        self._count = 0
        self.cfl = np.float32(1000)
        self.timestepMin = np.float32(1.0)
        self.timestepMax = np.float32(1.0)
        self.frameLength = np.float32(1.0)
        self.timePerFrame = np.float32(0.0)
        self._lockDt = False

    @property
    def timestep(self):
        return self._dt

    @timestep.setter
    def timestep(self, v):
        self._dt = np.float32(v)
        check(lib().mf_solver_set_dt(self._h, float(self._dt)))

    def getGridSize(self): return vec3(*self._size)
    def is2D(self): return self._dim == 2
    def is3D(self): return self._dim == 3
    def getDx(self): return np.float32(1.0 / max(self._size))

    def create(self, type=None, name="", **kwargs):
        if type is None:
            raise MantaError("Need to specify object type. Use e.g. Solver.create(FlagGrid, ...) or Solver.create(type=FlagGrid, ...)")
        return type(self, name=name, **kwargs)

    def step(self, frame=-1):
        """fluidsolver.cpp:143-166"""
        eps = np.float32(1e-6)
        self.timePerFrame = np.float32(self.timePerFrame + self._dt)
        self.timeTotal = np.float32(self.timeTotal + self._dt)
        self._count += 1
        if (self.timePerFrame + eps) > self.frameLength:
            self.frame += 1
            self.timeTotal = np.float32(float(self.frame) * float(self.frameLength))
            self.timePerFrame = np.float32(0.0)
            self._lockDt = False
        if frame >= 0:
            self.frame = frame

    def adaptTimestep(self, maxVel):
        """fluidsolver.cpp:184-205"""
        f = np.float32
        mvt = f(maxVel) * self._dt
        if not self._lockDt:
            dt = max(min(self._dt * f(float(self.cfl) / (float(mvt) + 1e-05)), self.timestepMax), self.timestepMin)
            if (float(self.timePerFrame) + float(dt) * 1.05) > float(self.frameLength):
                dt = f(float(self.frameLength - self.timePerFrame) + 1e-04)
            elif (self.timePerFrame + dt + self.timestepMin) > self.frameLength or (float(self.timePerFrame) + float(dt) * 1.25) > float(self.frameLength):
                dt = f((float(self.frameLength - self.timePerFrame) + 1e-04) * 0.5)
                self._lockDt = True
            self.timestep = dt
        if not (self._dt > self.timestepMin / 2.0):
            raise MantaError("Invalid dt encountered! Shouldnt happen...")

    def transfersSync(self):
        """wait for the solver stream and all streaming transfers"""
        check(lib().mf_solver_transfers_sync(self._h))

    def sync(self):
        check(lib().mf_solver_sync(self._h))

    def launchCount(self):
        return int(lib().mf_solver_launch_count(self._h))

    def slab(self, rank=-1):
        """z planes [k0, k1) of this process's (or `rank`'s) slab"""
        k0, k1 = C.c_int(), C.c_int()
        check(lib().mf_solver_slab(self._h, rank, C.byref(k0), C.byref(k1)))
        return k0.value, k1.value

    def printMemInfo(self):
        print("Device temp-grid pool: %d bytes" % lib().mf_solver_pool_bytes(self._h))

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            try:
                lib().mf_solver_destroy(h)
            except Exception:
                pass


Solver = FluidSolver


# ---- grids -----------------------------------------------------------------------------------------------
class GridBase(object):
    _type = 0
    _dtype = np.float32
    _ncomp = 1

    def __init__(self, parent, show=True, name="", **kwargs):
        _checkArgs("Grid", kwargs, ())
        if not isinstance(parent, FluidSolver):
            raise MantaError("Grid: need a Solver as parent")
        self.parent = parent
        self._name = name
        self._h = lib().mf_grid_create(parent._h, self._type)
        if not self._h:
            raise MantaError(lib().mf_last_error().decode())

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h and getattr(self.parent, "_h", None):
            try:
                lib().mf_grid_destroy(h)
            except Exception:
                pass

    # --- reference Python methods (source/grid.h:27-247) ---
    def get_name(self): return self._name
    def set_name(self, s): self._name = s
    def getSizeX(self): return self.parent._size[0]
    def getSizeY(self): return self.parent._size[1]
    def getSizeZ(self): return self.parent._size[2]
    def getSize(self): return vec3(*self.parent._size)
    def is3D(self): return self.parent.is3D()
    def is4D(self): return False
    def getGridType(self): return self._type
    def getParent(self): return self.parent

    def clear(self): check(lib().mf_grid_clear(self._h))

    def copyFrom(self, a, copyType=True):
        check(lib().mf_grid_copy(self._h, a._h))
        return self

    def swap(self, other): check(lib().mf_grid_swap(self._h, other._h))

    def getMaxAbs(self):
        out = C.c_float()
        check(lib().mf_grid_max_abs(self._h, C.byref(out)))
        return out.value

    def getMax(self):
        mn, mx = C.c_float(), C.c_float()
        check(lib().mf_grid_min_max(self._h, C.byref(mn), C.byref(mx)))
        return mx.value

    def getMin(self):
        mn, mx = C.c_float(), C.c_float()
        check(lib().mf_grid_min_max(self._h, C.byref(mn), C.byref(mx)))
        return mn.value

    # --- host <-> device (reference layout: x fastest; Vec3 = 3 packed floats) ---
    def _shape(self):
        nx, ny, nz = self.parent._size
        return (nz, ny, nx) if self._ncomp == 1 else (nz, ny, nx, 3)

    def to_numpy(self):
        a = np.empty(self._shape(), dtype=self._dtype)
        check(lib().mf_grid_download(self._h, a.ctypes.data_as(C.c_void_p)))
        return a

    def from_numpy(self, a):
        a = np.ascontiguousarray(a, dtype=self._dtype)
        if a.shape != self._shape():
            a = a.reshape(self._shape())
        check(lib().mf_grid_upload(self._h, a.ctypes.data_as(C.c_void_p)))
        return self

    # --- streaming transfers (pinned host slabs, double-buffered staging, copies overlap the solver stream) ---
    def stageUpload(self, host_ptr):
        """start the host->device copy of this process' slab (pointer from mf_host_alloc / a pinned array)"""
        check(lib().mf_grid_stage_upload(self._h, C.c_void_p(host_ptr)))

    def commitUpload(self):
        """the solver stream takes over the oldest staged upload"""
        check(lib().mf_grid_commit_upload(self._h))

    def downloadAsync(self, host_ptr):
        """snapshot the grid now (in stream order) and copy it to the host in the background"""
        check(lib().mf_grid_download_async(self._h, C.c_void_p(host_ptr)))

    def get(self, i, j, k):
        v = self.to_numpy()[k, j, i]
        return vec3(v) if self._ncomp == 3 else v.item()

    # --- file I/O on the host mirror (source/grid.cpp:136-179, fileio/iogrids.cpp:255-300,386-513) ---
    def save(self, name):
        a = self.to_numpy()
        if name.endswith(".raw"):
            with gzip.open(name, "wb", compresslevel=1) as f:
                f.write(a.tobytes())
        elif name.endswith(".uni"):
            # "MNT3" + UniHeader v4 (fileio/iogrids.cpp:38-44): 6 ints, info[252], int dimT, 8-byte timestamp = 288 bytes
            nx, ny, nz = self.parent._size
            et = 0 if self._dtype == np.int32 else (2 if self._ncomp == 3 else 1)
            info = lib().mf_build_info()[:251]
            head = struct.pack("<iiiiii252siQ", nx, ny, nz, self._type, et, a.itemsize * self._ncomp,
                               info, 0, int(_time.time()))
            assert len(head) == _UNI_HEADER_BYTES
            with gzip.open(name, "wb", compresslevel=1) as f:
                f.write(b"MNT3")
                f.write(head)
                f.write(a.tobytes())
        elif name.endswith(".npy"):
            np.save(name, a)
        else:
            raise MantaError("file '%s' filetype not supported" % name)
        return 1

    def load(self, name):
        if name.endswith(".raw"):
            with gzip.open(name, "rb") as f:
                a = np.frombuffer(f.read(), dtype=self._dtype)
        elif name.endswith(".uni"):
            with gzip.open(name, "rb") as f:
                ident = f.read(4)
                if ident not in (b"MNT3", b"MNT2"):     # MNT2 (UniLegacyHeader3: info[256], no dimT) has the same 288 bytes
                    raise MantaError("readGridUni: Unknown header '%s' " % ident.decode("ascii", "replace"))
                head = f.read(_UNI_HEADER_BYTES)
                if len(head) != _UNI_HEADER_BYTES:
                    raise MantaError("can't read file, no header present")
                nx, ny, nz, gt, et, bpe = struct.unpack("<iiiiii", head[:24])
                if (nx, ny, nz) != self.parent._size:
                    raise MantaError("grid dim doesn't match, [%d,%d,%d] vs %s" % (nx, ny, nz, self.getSize()))
                if _unifyGridType(gt) != _unifyGridType(self._type):
                    raise MantaError("grid type doesn't match %d vs %d" % (gt, self._type))
                if bpe != np.dtype(self._dtype).itemsize * self._ncomp:
                    raise MantaError("grid element size doesn't match %d vs %d" % (bpe, np.dtype(self._dtype).itemsize * self._ncomp))
                a = np.frombuffer(f.read(), dtype=self._dtype)
        elif name.endswith(".npy"):
            a = np.load(name)
        else:
            raise MantaError("file '%s' filetype not supported" % name)
        self.from_numpy(a.reshape(self._shape()))
        return 1


class RealGrid(GridBase):
    _type = _TYPE_REAL

    def setConst(self, s): check(lib().mf_grid_set_const_real(self._h, float(s)))
    def addScaled(self, a, factor): check(lib().mf_grid_add_scaled(self._h, a._h, float(factor), 0.0, 0.0))
    def add(self, a): check(lib().mf_grid_add_scaled(self._h, a._h, 1.0, 0.0, 0.0))
    def sub(self, a): check(lib().mf_grid_add_scaled(self._h, a._h, -1.0, 0.0, 0.0))
    def multConst(self, s): check(lib().mf_grid_mult_const(self._h, float(s), 0.0, 0.0))


class LevelsetGrid(RealGrid):
    _type = _TYPE_REAL | _TYPE_LEVELSET

    def join(self, o):
        """levelset union (levelset.cpp:111)"""
        check(lib().mf_levelset_join(self._h, o._h))


class IntGrid(GridBase):
    _type = _TYPE_INT
    _dtype = np.int32

    def setConst(self, s): check(lib().mf_grid_set_const_int(self._h, int(s)))


class VecGrid(GridBase):
    _type = _TYPE_VEC3
    _ncomp = 3

    def setConst(self, s):
        s = _v3(s)
        check(lib().mf_grid_set_const_vec3(self._h, float(s.x), float(s.y), float(s.z)))

    def addScaled(self, a, factor):
        f = _v3(factor)
        check(lib().mf_grid_add_scaled(self._h, a._h, float(f.x), float(f.y), float(f.z)))

    def multConst(self, s):
        f = _v3(s)
        check(lib().mf_grid_mult_const(self._h, float(f.x), float(f.y), float(f.z)))


Vec3Grid = VecGrid


class MACGrid(VecGrid):
    _type = _TYPE_MAC | _TYPE_VEC3


class FlagGrid(IntGrid):
    _type = _TYPE_FLAGS | _TYPE_INT

    def __init__(self, parent, dim=3, show=True, name="", **kwargs):
        IntGrid.__init__(self, parent, show=show, name=name, **kwargs)

    def initDomain(self, boundaryWidth=0, wall="xXyYzZ", open="      ", inflow="      ", outflow="      ", phiWalls=None):
        """FlagGrid::initDomain (source/grid.cpp:798-888)"""
        check(lib().mf_flags_init_domain(self._h, int(boundaryWidth), wall.encode(), open.encode(), inflow.encode(),
                                         outflow.encode(), phiWalls._h if phiWalls is not None else None))

    def fillGrid(self, type=TypeFluid):
        check(lib().mf_flags_fill_grid(self._h, int(type)))

    def updateFromLevelset(self, levelset):
        """fluid / empty from the sign of a levelset (grid.cpp:910)"""
        check(lib().mf_flags_update_from_levelset(self._h, levelset._h))

    def countCells(self, flag, bnd=0, mask=None):
        out = C.c_int()
        check(lib().mf_flags_count_cells(self._h, int(flag), int(bnd), C.byref(out)))
        return out.value


# ---- shapes used by the config scenes (source/shapes.cpp) -------------------------------------------------
class Shape(object):
    """Shape base (source/shapes.h:27): applyToGrid / computeLevelset run on the device"""
    _kind = -1

    def __init__(self, parent):
        self.parent = parent

    def _desc(self):
        raise MantaError("Shape: abstract")

    def _struct(self):
        from ._lib import ShapeDesc
        a, b, r = self._desc()
        d = ShapeDesc()
        d.type = self._kind
        d.a[0], d.a[1], d.a[2] = float(a.x), float(a.y), float(a.z)
        d.b[0], d.b[1], d.b[2] = float(b.x), float(b.y), float(b.z)
        d.r = float(r)
        return d

    def applyToGrid(self, grid=None, value=0, respectFlags=None, **kwargs):
        _checkArgs("Shape::applyToGrid", kwargs, ())
        if grid is None:
            raise MantaError("Argument 'grid' is not defined.")
        d = self._struct()
        rf = respectFlags._h if respectFlags is not None else None
        if isinstance(grid, IntGrid):
            check(lib().mf_shape_apply(C.byref(d), grid._h, 0.0, 0.0, 0.0, int(value), rf))
        elif isinstance(grid, RealGrid):
            check(lib().mf_shape_apply(C.byref(d), grid._h, float(value), 0.0, 0.0, 0, rf))
        elif isinstance(grid, VecGrid):
            v = _v3(value)
            check(lib().mf_shape_apply(C.byref(d), grid._h, float(v.x), float(v.y), float(v.z), 0, rf))
        else:
            raise MantaError("Shape::applyToGrid(): unknown grid type")

    def computeLevelset(self):
        phi = LevelsetGrid(self.parent)
        d = self._struct()
        check(lib().mf_shape_levelset(C.byref(d), phi._h))
        return phi


class Box(Shape):
    _kind = 0

    def __init__(self, parent, center=None, p0=None, p1=None, size=None, **kwargs):
        _checkArgs("Box", kwargs, ("name",))
        Shape.__init__(self, parent)
        if center is not None and size is not None:
            self.p0, self.p1 = _v3(center) - _v3(size), _v3(center) + _v3(size)
        elif p0 is not None and p1 is not None:
            self.p0, self.p1 = _v3(p0), _v3(p1)
        else:
            raise MantaError("Box: specify either p0,p1 or size,center")

    def _desc(self):
        return self.p0, self.p1, 0.0


class Sphere(Shape):
    _kind = 1

    def __init__(self, parent, center=None, radius=None, scale=None, **kwargs):
        _checkArgs("Sphere", kwargs, ("name",))
        Shape.__init__(self, parent)
        self.center, self.radius = _v3(center), np.float32(radius)
        self.scale = _v3(scale) if scale is not None else vec3(1, 1, 1)

    def _desc(self):
        return self.center, self.scale, self.radius


class Cylinder(Shape):
    _kind = 2

    def __init__(self, parent, center=None, radius=None, z=None, **kwargs):
        _checkArgs("Cylinder", kwargs, ("name",))
        Shape.__init__(self, parent)
        self.center, self.radius, self.z = _v3(center), np.float32(radius), _v3(z)

    def _desc(self):
        return self.center, self.z, self.radius


# ---- wavelet noise (source/noisefield.h) ----------------------------------------------------------------------------
class NoiseField(object):
    """WaveletNoiseField (noisefield.h:26-97): python-visible members as attributes, evaluation on the device"""

    def __init__(self, parent, fixedSeed=-1, loadFromFile=False, name="", **kwargs):
        _checkArgs("NoiseField", kwargs, ())
        self.parent = parent
        self._name = name
        self.posOffset = vec3(0.0)
        self.posScale = vec3(1.0)
        self.valOffset = 0.0
        self.valScale = 1.0
        self.clamp = False
        self.clampNeg = 0.0
        self.clampPos = 1.0
        self.timeAnim = 0.0
        self._h = lib().mf_noise_create(parent._h, int(fixedSeed), 1 if loadFromFile else 0)
        if not self._h:
            check(1)

    def __del__(self):
        try:
            if getattr(self, "_h", None):
                lib().mf_noise_destroy(self._h)
                self._h = None
        except Exception:
            pass

    def _params(self):
        from ._lib import NoiseParams
        p = NoiseParams()
        po, ps = _v3(self.posOffset), _v3(self.posScale)
        p.posOffset[0], p.posOffset[1], p.posOffset[2] = float(po.x), float(po.y), float(po.z)
        p.posScale[0], p.posScale[1], p.posScale[2] = float(ps.x), float(ps.y), float(ps.z)
        p.valOffset, p.valScale = float(self.valOffset), float(self.valScale)
        p.clamp = 1 if self.clamp else 0
        p.clampNeg, p.clampPos = float(self.clampNeg), float(self.clampPos)
        p.timeAnim = float(self.timeAnim)
        return p

    def evaluate(self, positions):
        """WaveletNoiseField::evaluate at an (n, 3) array of positions (a test hook: the reference evaluates inside its kernels)"""
        pos = np.ascontiguousarray(positions, np.float32).reshape(-1, 3)
        out = np.empty(pos.shape[0], np.float32)
        p = self._params()
        fp = C.POINTER(C.c_float)
        check(lib().mf_noise_evaluate(self._h, C.byref(p), float(self.parent.timeTotal), pos.ctypes.data_as(fp), pos.shape[0], out.ctypes.data_as(fp)))
        return out


# ---- FLIP particles (source/particle.h) -------------------------------------------------------------------------------
IntEuler, IntRK2, IntRK4 = 0, 1, 2           # IntegrationMode (util/integrator.h:23)


class ParticleDataBase(object):
    """ParticleDataImpl<T> (particle.h:310): one value per particle, stored on the device next to the positions"""
    _ncomp = 0

    def __init__(self, parts, index, name=""):
        self.parts, self._index, self._name = parts, index, name

    def to_numpy(self):
        n = self.parts.size()
        nc = 3 if self._ncomp == 3 else 1
        out = np.empty((n, nc), np.int32 if self._ncomp == -1 else np.float32)
        if n:
            check(lib().mf_pdata_download(self.parts._h, self._index, out.ctypes.data_as(C.c_void_p)))
        return out if nc == 3 else out[:, 0]

    def from_numpy(self, a):
        n = self.parts.size()
        nc = 3 if self._ncomp == 3 else 1
        a = np.ascontiguousarray(a, np.int32 if self._ncomp == -1 else np.float32).reshape(n, nc)
        if n:
            check(lib().mf_pdata_upload(self.parts._h, self._index, a.ctypes.data_as(C.c_void_p)))


class PdataReal(ParticleDataBase):
    _ncomp = 1


class PdataVec3(ParticleDataBase):
    _ncomp = 3


class PdataInt(ParticleDataBase):
    _ncomp = -1


class BasicParticleSystem(object):
    """BasicParticleSystem (particle.h:198): positions + status flags on the device"""

    def __init__(self, parent, name="", **kwargs):
        _checkArgs("BasicParticleSystem", kwargs, ())
        self.parent, self._name = parent, name
        self._h = lib().mf_parts_create(parent._h)
        if not self._h:
            check(1)

    def __del__(self):
        try:
            if getattr(self, "_h", None):
                lib().mf_parts_destroy(self._h)
                self._h = None
        except Exception:
            pass

    def create(self, type=None, name="", **kwargs):
        if type is None or not issubclass(type, ParticleDataBase) or type is ParticleDataBase:
            raise MantaError("ParticleSystem::create: need a particle data type (PdataReal, PdataVec3, PdataInt)")
        idx = C.c_int(-1)
        check(lib().mf_parts_create_pdata(self._h, type._ncomp, C.byref(idx)))
        return type(self, idx.value, name)

    def size(self): return int(lib().mf_parts_size(self._h))
    def pySize(self): return self.size()

    def advectInGrid(self, flags=None, vel=None, integrationMode=None, deleteInObstacle=True, stopInObstacle=True, skipNew=False,
                     ptype=None, exclude=0, **kwargs):
        _checkArgs("advectInGrid", kwargs, ())
        if ptype is not None:
            raise MantaError("advectInGrid: ptype / exclude are not supported")
        if integrationMode is None:
            raise MantaError("Argument 'integrationMode' is not defined.")
        check(lib().mf_parts_advect_in_grid(self._h, flags._h, vel._h, int(integrationMode), 1 if deleteInObstacle else 0,
                                            1 if stopInObstacle else 0, 1 if skipNew else 0))

    def getPosPdata(self, target=None):
        pos, _ = self.to_numpy()
        target.from_numpy(pos)

    def to_numpy(self):
        """(positions (n, 3) float32, status flags (n,) int32)"""
        n = self.size()
        pos = np.empty((n, 3), np.float32); flag = np.empty(n, np.int32)
        if n:
            check(lib().mf_parts_download(self._h, pos.ctypes.data_as(C.POINTER(C.c_float)), flag.ctypes.data_as(C.POINTER(C.c_int))))
        return pos, flag

    def from_numpy(self, pos, flag=None):
        pos = np.ascontiguousarray(pos, np.float32).reshape(-1, 3)
        fp = None
        if flag is not None:
            flag = np.ascontiguousarray(flag, np.int32)
            fp = flag.ctypes.data_as(C.POINTER(C.c_int))
        check(lib().mf_parts_upload(self._h, pos.ctypes.data_as(C.POINTER(C.c_float)), fp, pos.shape[0]))


from .plugins import *  # noqa: E402,F401,F403  (plugin functions of the hot path)
