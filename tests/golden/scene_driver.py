# Executed by BOTH sides with the same environment:
#     oracle/_ref/manta tests/golden/scene_driver.py            (the unmodified reference)
#     python -m mantaflow_b200 tests/golden/scene_driver.py     (device path)
# Runs the UNMODIFIED bytes of $MF_SCENE (a reference scene file), with the frame loop cut to $MF_SCENE_STEPS steps by shadowing
# the builtin `range` in the scene's globals (the file itself is not touched), then writes the objects named in $MF_SCENE_SAVE
# ("density:RealGrid,vel:MACGrid,pp:BasicParticleSystem,...") to $MF_SCENE_OUT.
import builtins
import os

_scene = os.environ["MF_SCENE"]
_limit = int(os.environ.get("MF_SCENE_STEPS", "5"))
_out = os.environ["MF_SCENE_OUT"]
_save = [x.split(":") for x in os.environ["MF_SCENE_SAVE"].split(",") if x]


def _short_range(*a):
    r = builtins.range(*a)
    return r[:_limit] if len(r) > _limit else r


# the reference injects source/python/defines.py (IntRK4, FlagFluid, ...) into the globals of the file it runs: pass them on
_g = dict(globals())
_g.update({"__name__": "__main__", "__file__": _scene, "range": _short_range})
with open(_scene, "rb") as _f:
    _src = _f.read()
exec(compile(_src, _scene, "exec"), _g)

for _name, _gtype in _save:
    _o = _g[_name]
    if hasattr(_o, "to_numpy"):                      # mantaflow_b200: straight from device memory
        import numpy as _np
        _a = _o.to_numpy()
        _np.save(os.path.join(_out, "out_%s.npy" % _name), _a[0] if _gtype == "BasicParticleSystem" else _a)
    elif _gtype == "BasicParticleSystem":            # reference: positions through a particle-data copy
        _tmp = _o.create(_g["PdataVec3"])
        _o.getPosPdata(target=_tmp)
        _tmp.save(os.path.join(_out, "out_%s.uni" % _name))
    elif _gtype.startswith("Pdata"):
        _o.save(os.path.join(_out, "out_%s.uni" % _name))
    else:
        _o.save(os.path.join(_out, "out_%s.raw" % _name))
