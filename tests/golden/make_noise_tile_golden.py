#!/usr/bin/env python
"""Golden data for the wavelet-noise tile: runs the UNMODIFIED reference (oracle/_ref/manta) on a scene that creates a
NoiseField with loadFromFile=True, which makes it write its 3 x 128^3 tile to ./waveletNoiseTile.bin (noisefield.cpp:172-179),
and stores the tile's sha256, its float64 sum and every 1021st value in tests/golden/noise_tile_ref.npz.
Run in the build container only:   python tests/golden/make_noise_tile_golden.py
"""
import hashlib
import os
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
MANTA = os.path.join(ROOT, "oracle", "_ref", "manta")
SCENE = """from manta import *
s = Solver(name='main', gridSize=vec3(16, 16, 16), dim=3)
n = s.create(NoiseField, loadFromFile=True)
"""


def main():
    if not os.path.exists(MANTA):
        sys.exit("oracle/_ref/manta missing: run oracle/build_ref.sh first")
    with tempfile.TemporaryDirectory() as d:
        open(os.path.join(d, "scene.py"), "w").write(SCENE)
        subprocess.run([MANTA, "scene.py"], cwd=d, check=True, capture_output=True, env=dict(os.environ, MANTA_DISABLE_UI="1"))
        raw = open(os.path.join(d, "waveletNoiseTile.bin"), "rb").read()
    t = np.frombuffer(raw, np.float32)
    assert t.size == 3 * 128 ** 3
    np.savez_compressed(os.path.join(HERE, "noise_tile_ref.npz"), sha256=np.array(hashlib.sha256(raw).hexdigest()),
                        total=np.array(t.astype(np.float64).sum()), sample=t[::1021].copy())
    print("tile sha256", hashlib.sha256(raw).hexdigest())


if __name__ == "__main__":
    main()
