"""Buoyant smoke plume with a noise-textured source -- the calls of the reference's hello-world scene (scenes/simpleplume.py)
against mantaflow_b200.manta.  Every plugin runs on the GPU; nothing is copied to the host inside the loop.

    python examples/simpleplume.py [--res 64] [--steps 250]
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mantaflow_b200.manta import *  # noqa: E402,F401,F403

ap = argparse.ArgumentParser()
ap.add_argument("--res", type=int, default=64)
ap.add_argument("--steps", type=int, default=250)
a = ap.parse_args()

res = a.res
gs = vec3(res, int(1.5 * res), res)
s = FluidSolver(name="main", gridSize=gs)

flags = s.create(FlagGrid)
vel = s.create(MACGrid)
density = s.create(RealGrid)
pressure = s.create(RealGrid)

noise = s.create(NoiseField, loadFromFile=True)
noise.posScale = vec3(45)
noise.clamp = True
noise.clampNeg = 0
noise.clampPos = 1
noise.valOffset = 0.75
noise.timeAnim = 0.2

source = s.create(Cylinder, center=gs * vec3(0.5, 0.1, 0.5), radius=res * 0.14, z=gs * vec3(0, 0.02, 0))

flags.initDomain()
flags.fillGrid()

for t in range(a.steps):
    if t < 100:
        densityInflow(flags=flags, density=density, noise=noise, shape=source, scale=1, sigma=0.5)
    advectSemiLagrange(flags=flags, vel=vel, grid=density, order=2)
    advectSemiLagrange(flags=flags, vel=vel, grid=vel, order=2, strength=1.0)
    setWallBcs(flags=flags, vel=vel)
    addBuoyancy(density=density, vel=vel, gravity=vec3(0, -6e-4, 0), flags=flags)
    solvePressure(flags=flags, vel=vel, pressure=pressure)
    s.step()

s.sync()
print("steps %d  max density %.6f  max |vel| %.6f" % (a.steps, density.getMax(), vel.getMaxAbs()))
