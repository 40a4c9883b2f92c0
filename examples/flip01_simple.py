"""Breaking dam with FLIP particles (no levelset, no resampling) -- the calls of the reference's scenes/flip01_simple.py against
mantaflow_b200.manta.  Particles, transfers and the pressure projection all stay on the GPU.

    python examples/flip01_simple.py [--res 64] [--dim 2] [--steps 250]
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mantaflow_b200.manta import *  # noqa: E402,F401,F403

ap = argparse.ArgumentParser()
ap.add_argument("--res", type=int, default=64)
ap.add_argument("--dim", type=int, default=2)
ap.add_argument("--steps", type=int, default=250)
a = ap.parse_args()

dim, res = a.dim, a.res
particleNumber = 3 if dim == 2 else 2
gs = vec3(res, res, res if dim == 3 else 1)
s = Solver(name="main", gridSize=gs, dim=dim)
s.timestep = 0.5

flags = s.create(FlagGrid)
vel = s.create(MACGrid)
velOld = s.create(MACGrid)
pressure = s.create(RealGrid)
tmpVec3 = s.create(VecGrid)
pp = s.create(BasicParticleSystem)
pVel = pp.create(PdataVec3)

flags.initDomain(boundaryWidth=0)
fluidbox = Box(parent=s, p0=gs * vec3(0, 0, 0), p1=gs * vec3(0.4, 0.6, 1))
phiInit = fluidbox.computeLevelset()
flags.updateFromLevelset(phiInit)
sampleFlagsWithParticles(flags=flags, parts=pp, discretization=particleNumber, randomness=0.2)

for t in range(a.steps):
    pp.advectInGrid(flags=flags, vel=vel, integrationMode=IntRK4, deleteInObstacle=False)
    mapPartsToMAC(vel=vel, flags=flags, velOld=velOld, parts=pp, partVel=pVel, weight=tmpVec3)
    extrapolateMACFromWeight(vel=vel, distance=2, weight=tmpVec3)
    markFluidCells(parts=pp, flags=flags)
    addGravity(flags=flags, vel=vel, gravity=vec3(0, -0.002, 0))
    setWallBcs(flags=flags, vel=vel)
    solvePressure(flags=flags, vel=vel, pressure=pressure)
    setWallBcs(flags=flags, vel=vel)
    extrapolateMACSimple(flags=flags, vel=vel)
    flipVelocityUpdate(vel=vel, velOld=velOld, flags=flags, parts=pp, partVel=pVel, flipRatio=0.97)
    s.step()

s.sync()
pos, _ = pp.to_numpy()
print("steps %d  particles %d  max |vel| %.6f  mean height %.4f" % (a.steps, pp.size(), vel.getMaxAbs(), float(pos[:, 1].mean())))
