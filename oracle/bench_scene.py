# TEST / BASELINE INFRASTRUCTURE ONLY -- scene file executed BY THE REFERENCE binary
# (oracle/_ref/manta oracle/bench_scene.py); never imported by the product.
#
# Runs the synthetic smoke-plume step of bench.py (Cylinder source via applyToGrid,
# MacCormack advection of density and velocity, setWallBcs, addBuoyancy, solvePressure) through the
# reference's own plugins on the host cores and prints the wall time of every step:
#   BENCH_JSON {"n": [nx,ny,nz], "step_s": [...], "iterations": [...]}
# Settings come from the environment: MF_BENCH_SIZE "nx,ny,nz", MF_BENCH_STEPS, MF_BENCH_WARMUP,
# MF_BENCH_PC (1 = PcMIC, 2 = PcMGDynamic, 3 = PcMGStatic), MF_BENCH_DUMP (directory for final grids, optional).
import json
import os
import sys
import time
from manta import *

nx, ny, nz = [int(v) for v in os.environ.get("MF_BENCH_SIZE", "64,64,64").split(",")]
steps = int(os.environ.get("MF_BENCH_STEPS", "2"))
warmup = int(os.environ.get("MF_BENCH_WARMUP", "1"))
pc = int(os.environ.get("MF_BENCH_PC", "3"))
acc = float(os.environ.get("MF_BENCH_ACC", "1e-3"))
dump = os.environ.get("MF_BENCH_DUMP", "")

gs = vec3(nx, ny, nz)
s = Solver(name="main", gridSize=gs, dim=3)
s.timestep = 1.0
flags = s.create(FlagGrid)
vel = s.create(MACGrid)
density = s.create(RealGrid)
pressure = s.create(RealGrid)
flags.initDomain(boundaryWidth=0)
flags.fillGrid()
source = Cylinder(parent=s, center=gs * vec3(0.5, 0.1, 0.5), radius=nx * 0.14, z=gs * vec3(0, 0.02, 0))
setDebugLevel(0)

times = []
for t in range(warmup + steps):
    t0 = time.time()
    source.applyToGrid(grid=density, value=1)
    advectSemiLagrange(flags=flags, vel=vel, grid=density, order=2)
    advectSemiLagrange(flags=flags, vel=vel, grid=vel, order=2)
    setWallBcs(flags=flags, vel=vel)
    addBuoyancy(density=density, vel=vel, gravity=vec3(0, -6e-4, 0), flags=flags)
    # closed box, all-Neumann system: pin one cell as tools/tests/test_0110_mgsolve.py does (without it GridMg's
    # coarsest-level CG never converges on the singular operator and the solve runs to maxIter)
    solvePressure(flags=flags, vel=vel, pressure=pressure, preconditioner=pc, cgAccuracy=acc, zeroPressureFixing=True)
    s.step()
    times.append(time.time() - t0)

if dump:
    density.save(os.path.join(dump, "ref_density.raw"))
    vel.save(os.path.join(dump, "ref_vel.raw"))
print("BENCH_JSON " + json.dumps({"n": [nx, ny, nz], "step_s": times, "warmup": warmup, "steps": steps, "pc": pc}))
sys.stdout.flush()
