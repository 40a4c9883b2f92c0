# Executed by the reference binary (oracle/_ref/manta tests/golden/uni_driver.py): .uni interoperability check.
#   in_<name>.raw  -> grid.load -> grid.save("ref_<name>.uni")        (a file the REFERENCE wrote)
#   our_<name>.uni -> grid.load -> grid.save("back_<name>.raw")       (a file mantaflow_b200 wrote, read by the reference)
import os
from manta import *

d = os.environ["MF_CASE_DIR"]
nx, ny, nz = [int(x) for x in os.environ["MF_UNI_SIZE"].split(",")]
s = Solver(name="main", gridSize=vec3(nx, ny, nz), dim=3 if nz > 1 else 2)
types = {"RealGrid": RealGrid, "MACGrid": MACGrid, "FlagGrid": FlagGrid, "LevelsetGrid": LevelsetGrid, "VecGrid": VecGrid, "IntGrid": IntGrid}
for spec in os.environ["MF_UNI_GRIDS"].split(","):
    name, gt = spec.split(":")
    g = s.create(types[gt])
    if os.path.exists(os.path.join(d, "in_%s.raw" % name)):
        g.load(os.path.join(d, "in_%s.raw" % name))
        g.save(os.path.join(d, "ref_%s.uni" % name))
    if os.path.exists(os.path.join(d, "our_%s.uni" % name)):
        g.load(os.path.join(d, "our_%s.uni" % name))
        g.save(os.path.join(d, "back_%s.raw" % name))
