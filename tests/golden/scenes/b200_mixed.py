# Test scene for the manta_b200 binding (NOT a reference file): the plume loop of scenes/simpleplume.py with one plugin that
# is not ported to the device (vorticityConfinement, source/plugin/extforces.cpp:419) between ported ones, and host-side grid
# methods, so that grids travel device -> host -> device through the mirror's dirty bits.  Runs unchanged on the reference.
from manta import *

res = 32
gs = vec3(res, int(1.5 * res), res)
s = FluidSolver(name='main', gridSize=gs)
flags = s.create(FlagGrid)
vel = s.create(MACGrid)
density = s.create(RealGrid)
pressure = s.create(RealGrid)
flags.initDomain()
flags.fillGrid()
source = s.create(Cylinder, center=gs * vec3(0.5, 0.1, 0.5), radius=res * 0.14, z=gs * vec3(0, 0.02, 0))
stats = []
for t in range(12):
	source.applyToGrid(grid=density, value=1)                      # host (Shape::applyToGrid)
	advectSemiLagrange(flags=flags, vel=vel, grid=density, order=2)  # device
	advectSemiLagrange(flags=flags, vel=vel, grid=vel, order=2)      # device
	if t % 2 == 1:
		vorticityConfinement(vel=vel, flags=flags, strength=0.4)     # host, every other step
	setWallBcs(flags=flags, vel=vel)                               # device
	addBuoyancy(density=density, vel=vel, gravity=vec3(0, -6e-4, 0), flags=flags)
	solvePressure(flags=flags, vel=vel, pressure=pressure, preconditioner=PcMGStatic, zeroPressureFixing=True)
	maxVel = vel.getMax()                                           # host-side reduction on a device-written grid
	s.step()
	if 'b200Stats' in globals():
		stats.append(b200Stats())
if stats:
	print('B200STATS ' + ' | '.join(stats))
