"""No GPU here: the reference-side binding (integration/_build/manta_b200) must fail LOUDLY in its ported plugins -- there is no
CPU fallback behind the C-ABI -- while everything un-ported keeps running on the host; and the binding's sources must not have
drifted from the plugin list they replace."""
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
MANTA_B200 = os.path.join(ROOT, "integration", "_build", "manta_b200")
PORTED = ["advectSemiLagrange", "addGravity", "addGravityNoScale", "addBuoyancy", "setOpenBound", "resetOutflow", "setInflowBcs", "setWallBcs",
          "releaseMG", "computePressureRhs", "solvePressureSystem", "correctVelocity", "solvePressure"]


def test_sources_agree_on_the_ported_plugins():
    plug = open(os.path.join(ROOT, "integration", "b200_plugins.cpp")).read()
    mirror = open(os.path.join(ROOT, "integration", "b200_mirror.cpp")).read()
    build = open(os.path.join(ROOT, "integration", "build_manta_b200.sh")).read()
    for n in PORTED:
        assert re.search(r"PYTHON\(\)\s+void\s+%s\s*\(" % n, plug), n
        assert '"%s"' % n in mirror, n
        assert re.search(r"\b%s\b" % n, build), n


def test_ported_plugins_fail_loudly_without_a_device(tmp_path):
    if not os.path.exists(MANTA_B200):
        pytest.skip("integration/_build/manta_b200 not built (needs the reference checkout)")
    try:
        import ctypes
        if ctypes.CDLL("libcuda.so.1").cuInit(0) == 0:
            pytest.skip("a CUDA device is present")
    except OSError:
        pass
    scene = tmp_path / "s.py"
    scene.write_text(
        "from manta import *\n"
        "s = Solver(name='main', gridSize=vec3(16,16,16))\n"
        "flags = s.create(FlagGrid); vel = s.create(MACGrid); density = s.create(RealGrid)\n"
        "flags.initDomain(); flags.fillGrid()\n"
        "try:\n"
        "    advectSemiLagrange(flags=flags, vel=vel, grid=density, order=2)\n"
        "    print('NO-ERROR')\n"
        "except RuntimeError as e:\n"
        "    print('LOUD:', str(e).splitlines()[0])\n"
        "vorticityConfinement(vel=vel, flags=flags, strength=0.1)\n"
        "print('HOST-OK', b200Stats())\n")
    p = subprocess.run([MANTA_B200, str(scene)], env=dict(os.environ, MANTA_DISABLE_UI="1"), capture_output=True, text=True, timeout=120, cwd=str(tmp_path))
    assert "LOUD: mantacuda: no CUDA device available" in p.stdout, p.stdout + p.stderr
    assert "NO-ERROR" not in p.stdout and "HOST-OK" in p.stdout
