"""The reference-side binding, built and run (SURVEY 8 rows a18 / b): integration/_build/manta_b200 is the
reference's own executable -- its prep-generated Python API, pwrapper, and every un-ported plugin, compiled from the unmodified
sources -- with the hot-path plugin bodies replaced by C-ABI calls into libmantacuda.so (integration/b200_plugins.cpp) and the
host-mirror / dirty-bit bookkeeping interposed at link time (integration/b200_mirror.cpp).  The unmodified scene files run
through it as `manta_b200 scene.py` and must agree with `oracle/_ref/manta scene.py`; a mixed scene runs an un-ported plugin
(vorticityConfinement) and host-side grid methods between ported plugins and checks the transfers the mirror made."""
import os
import tempfile

import numpy as np
import pytest

import test_scenes_gpu as TS

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
MANTA_B200 = os.path.join(ROOT, "integration", "_build", "manta_b200")


def _need():
    if not os.path.exists(TS.MANTA_REF) or not os.path.exists(MANTA_B200):
        pytest.skip("oracle/_ref/manta or integration/_build/manta_b200 not built (needs the reference checkout at build time)")


@pytest.mark.gpu
@pytest.mark.parametrize("name", list(TS.RUNS))
def test_manta_b200_runs_the_unmodified_scenes(name, parity_report):
    _need()
    scene, steps, save, thr, size = TS.RUNS[name]
    with tempfile.TemporaryDirectory() as cwd:
        ref = TS.run_reference(scene, steps, save, size, cwd=cwd)
        ours = TS.run_reference(scene, steps, save, size, cwd=cwd, binary=MANTA_B200)
    rep = parity_report.setdefault("binding_" + name, {"steps": steps, "size": list(size)})
    errors = []
    for g in save:
        d = TS.grid_max_diff(ours[g], ref[g])
        rep[g] = {"gridMaxDiff": d, "threshold": thr[g], "bit_identical": bool(np.array_equal(ours[g], ref[g]))}
        if d > thr[g]:
            errors.append("%s: gridMaxDiff %.3g > %.3g" % (g, d, thr[g]))
    assert not errors, "%s: %s" % (name, "; ".join(errors))


@pytest.mark.gpu
def test_manta_b200_host_plugins_between_device_plugins(parity_report):
    _need()
    save = {"density": "RealGrid", "vel": "MACGrid", "pressure": "RealGrid", "flags": "FlagGrid"}
    size = (32, 48, 32)
    with tempfile.TemporaryDirectory() as cwd:
        ref = TS.run_reference("b200_mixed.py", 12, save, size, cwd=cwd, extra_env={"OMP_NUM_THREADS": "1"})
        ours = TS.run_reference("b200_mixed.py", 12, save, size, cwd=cwd, binary=MANTA_B200, extra_env={"OMP_NUM_THREADS": "1"})
    rep = parity_report.setdefault("binding_mixed", {})
    for g in save:
        rep[g] = {"gridMaxDiff": TS.grid_max_diff(ours[g], ref[g]), "bit_identical": bool(np.array_equal(ours[g], ref[g]))}
        assert np.array_equal(ours[g], ref[g]), "%s differs from the reference (gridMaxDiff %.3g)" % (g, rep[g]["gridMaxDiff"])
    line = [l for l in ours["_stdout"].splitlines() if l.startswith("B200STATS ")]
    assert line, ours["_stdout"][-2000:]
    stats = [[int(x) for x in part.split()] for part in line[0][len("B200STATS "):].split(" | ")]
    rep["stats_per_step"] = stats
    up = [b[0] - a[0] for a, b in zip(stats, stats[1:])]        # uploads per step (from step 1 on)
    down = [b[1] - a[1] for a, b in zip(stats, stats[1:])]
    ported = [b[2] - a[2] for a, b in zip(stats, stats[1:])]
    assert all(p >= 5 for p in ported), ported                   # the five ported plugins of the loop ran on the device
    # even steps (t = 2, 4, ...): density goes to the host for applyToGrid and back (1 down + 1 up), vel to the host for
    # getMax of the previous step (already counted there) -- odd steps add vel down + up around vorticityConfinement
    even = [i for i in range(len(up)) if (i + 1) % 2 == 0]
    odd = [i for i in range(len(up)) if (i + 1) % 2 == 1]
    assert all(up[i] < up[j] for i in even for j in odd), (up, down)
    assert max(up) <= 4 and max(down) <= 4, (up, down)           # never a blanket re-upload of every grid: flags / pressure stay put
