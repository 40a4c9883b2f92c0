"""development: level-1 operator format and solve time of the bench configuration's pressure solve (plume, zeroPressureFixing)"""
import ctypes as Ct, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mantaflow_b200 import manta
from mantaflow_b200._lib import lib, check
n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
s = manta.Solver(name="s", gridSize=manta.vec3(n, n, n), dim=3)
flags = s.create(manta.FlagGrid); flags.initDomain(); flags.fillGrid()
vel = s.create(manta.MACGrid); p = s.create(manta.RealGrid)
rs = np.random.RandomState(3)
vel.from_numpy((rs.rand(n, n, n, 3).astype(np.float32) - np.float32(0.5)) * np.float32(0.1))
for rep in range(3):
    s.sync(); t0 = time.perf_counter()
    manta.solvePressure(flags=flags, vel=vel, pressure=p, cgAccuracy=1e-3, preconditioner=manta.PcMGStatic, zeroPressureFixing=True)
    s.sync(); t1 = time.perf_counter()
    fmt = Ct.c_int(0)
    check(lib().mf_mg_level_format(s._h, 1, Ct.byref(fmt)))
    print("solve %.2f ms, iterations %d, level-1 bytes/entry %d" % ((t1 - t0) * 1e3, manta.getSolveInfo()["iterations"], fmt.value))
