"""Host<->device copy rates through the streaming-transfer API: upload only, download only, both directions at once."""
import ctypes as C
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mantaflow_b200 import manta  # noqa: E402
from mantaflow_b200._lib import lib, check  # noqa: E402

L = lib()
n = 256
s = manta.Solver(name="s", gridSize=manta.vec3(n, n, n), dim=3)
a = s.create(manta.MACGrid); b = s.create(manta.MACGrid)
nb = 12 * n ** 3
pa = L.mf_host_alloc(nb); pb = L.mf_host_alloc(nb)
np.ctypeslib.as_array((C.c_byte * nb).from_address(pa))[:] = 1
np.ctypeslib.as_array((C.c_byte * nb).from_address(pb))[:] = 2
R = 8


def run(up, down):
    s.transfersSync()
    t0 = time.perf_counter()
    for _ in range(R):
        if up:
            a.stageUpload(pa); a.commitUpload()
        if down:
            b.downloadAsync(pb)
    s.transfersSync()
    return time.perf_counter() - t0


for _ in range(2):
    run(True, True)
tu, td, tb = run(True, False), run(False, True), run(True, True)
gb = nb * R / 1e9
print("upload %.1f GB/s  download %.1f GB/s  both: %.1f + %.1f GB/s in %.1f ms (serial sum would be %.1f ms)" % (
    gb / tu, gb / td, gb / tb, gb / tb, 1e3 * tb, 1e3 * (tu + td)))
