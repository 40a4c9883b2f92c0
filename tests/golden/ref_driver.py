# Scene file executed BY THE REFERENCE binary (oracle/_ref/manta tests/golden/ref_driver.py) in generate mode.
# Reads $MF_CASE_DIR/case.json + in_<grid>.raw, runs the case through the reference's own plugins,
# writes out_<grid>.raw + log.json.  See tests/golden/make_golden.py.
import json
import os
import sys
from manta import *

sys.path.insert(0, os.environ["MF_TESTS_DIR"])
import scene_runner as sr

d = os.environ["MF_CASE_DIR"]
case = json.load(open(os.path.join(d, "case.json")))
setDebugLevel(2)

def loader(name, grid, gtype):
    grid.load(os.path.join(d, "in_%s.raw" % name))

def saver(name, grid, gtype):
    if gtype == "BasicParticleSystem":       # positions through a particle-data copy (the particle .uni writer is unusable in this fork)
        tmp = grid.create(PdataVec3)
        grid.getPosPdata(target=tmp)
        tmp.save(os.path.join(d, "out_%s.uni" % name))
    elif sr.is_particle_type(gtype):
        grid.save(os.path.join(d, "out_%s.uni" % name))
    else:
        grid.save(os.path.join(d, "out_%s.raw" % name))
    return None

out, log = sr.run_case(case, globals(), loader, saver)
json.dump(log, open(os.path.join(d, "log.json"), "w"))
