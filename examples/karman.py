"""Von Karman vortex street behind a cylinder, 2-D, second-order obstacle boundaries -- the calls of the reference's
scenes/karman.py against mantaflow_b200.manta (fractions from the obstacle levelset, velocity extrapolation into the obstacle,
noise-perturbed inflow).  Every plugin runs on the GPU.

    python examples/karman.py [--res 64] [--steps 400]
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mantaflow_b200.manta import *  # noqa: E402,F401,F403

ap = argparse.ArgumentParser()
ap.add_argument("--res", type=int, default=64)
ap.add_argument("--steps", type=int, default=400)
a = ap.parse_args()

res = a.res
gs = vec3(2 * res, res, 1)
s = FluidSolver(name="main", gridSize=gs, dim=2)
s.timestep = 1.0

flags = s.create(FlagGrid)
vel = s.create(MACGrid)
density = s.create(RealGrid)
pressure = s.create(RealGrid)
fractions = s.create(MACGrid)
phiWalls = s.create(LevelsetGrid)

flags.initDomain(inflow="xX", phiWalls=phiWalls, boundaryWidth=0)
obstacle = Cylinder(parent=s, center=gs * vec3(0.25, 0.5, 0.5), radius=res * 0.2, z=gs * vec3(0, 0, 1.0))
phiObs = obstacle.computeLevelset()
densInflow = Cylinder(parent=s, center=gs * vec3(0.25, 0.5, 0.5), radius=res * 0.21, z=gs * vec3(0, 0, 1.0))

phiObs.join(phiWalls)
updateFractions(flags=flags, phiObs=phiObs, fractions=fractions)
setObstacleFlags(flags=flags, phiObs=phiObs, fractions=fractions)
flags.fillGrid()

velInflow = vec3(0.9, 0, 0)
vel.setConst(velInflow)

noise = s.create(NoiseField, loadFromFile=True)
noise.posScale = vec3(75)
noise.clamp = True
noise.clampNeg = -1.0
noise.clampPos = 1.0
everywhere = s.create(RealGrid)
everywhere.setConst(-1.0)
addNoise(flags=flags, density=density, noise=noise, sdf=everywhere, scale=0.1)
setComponent(target=vel, source=density, component=1)
density.setConst(0.0)

for t in range(a.steps):
    densInflow.applyToGrid(grid=density, value=2.0)
    advectSemiLagrange(flags=flags, vel=vel, grid=density, order=2, orderSpace=1)
    advectSemiLagrange(flags=flags, vel=vel, grid=vel, order=2, strength=1.0)
    extrapolateMACSimple(flags=flags, vel=vel, distance=2, intoObs=True)
    setWallBcs(flags=flags, vel=vel, fractions=fractions, phiObs=phiObs)
    setInflowBcs(vel=vel, dir="xX", value=velInflow)
    solvePressure(flags=flags, vel=vel, pressure=pressure, fractions=fractions, cgAccuracy=1e-4, cgMaxIterFac=5)
    extrapolateMACSimple(flags=flags, vel=vel, distance=5, intoObs=True)
    setWallBcs(flags=flags, vel=vel, fractions=fractions, phiObs=phiObs)
    setInflowBcs(vel=vel, dir="xX", value=velInflow)
    s.step()

s.sync()
print("steps %d  max density %.6f  max |vel| %.6f" % (a.steps, density.getMax(), vel.getMaxAbs()))
