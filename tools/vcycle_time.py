"""Time one multigrid V-cycle (mf_mg_vcycle) on an n^3 closed box with one Dirichlet cell (no trivial rows, so the level-1
operator is stored in binary16 unless MF_MG_NOHALF is set).  A tuning aid, not a bench number.

  python tools/vcycle_time.py --size 256 --reps 20
"""
import argparse
import ctypes as Ct
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--size", type=int, default=256)
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--levels", action="store_true", help="per-level cost with programmatic dependent launches on: V-cycles that leave out every level >= l (MF_MG_DEVSKIP)")
    a = ap.parse_args()
    import torch
    from mantaflow_b200 import manta
    from mantaflow_b200._lib import lib, check
    n = a.size
    s = manta.Solver(name="s", gridSize=manta.vec3(n, n, n), dim=3)
    flags = s.create(manta.FlagGrid); flags.initDomain(); flags.fillGrid()
    f = flags.to_numpy(); f[n // 2, n // 2, n // 2] = 4; flags.from_numpy(f)
    A0, Ai, Aj, Ak, R, D = [s.create(manta.RealGrid) for _ in range(6)]
    R.from_numpy(np.random.RandomState(1).rand(n, n, n).astype(np.float32) - np.float32(0.5))
    check(lib().mf_make_laplace_matrix(flags._h, A0._h, Ai._h, Aj._h, Ak._h, None))
    check(lib().mf_mg_vcycle(s._h, A0._h, Ai._h, Aj._h, Ak._h, R._h, D._h, 1e-7))
    def cyc():
        check(lib().mf_mg_vcycle(s._h, None, None, None, None, R._h, D._h, 1e-7))
    for _ in range(3):
        cyc()
    s.sync()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    # the library launches on its own stream: bracket with device-wide synchronisation and host timers as a cross-check
    import time
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(a.reps):
        cyc()
    s.sync(); torch.cuda.synchronize(); t1 = time.perf_counter()
    if a.levels:
        nl = Ct.c_int(0); fb = Ct.c_int(0)
        check(lib().mf_mg_info(s._h, Ct.byref(nl), Ct.byref(fb)))
        prev = None
        for skip in list(range(1, nl.value + 1)) + [1000]:
            os.environ["MF_MG_DEVSKIP"] = str(skip)
            for _ in range(3):
                cyc()
            s.sync(); torch.cuda.synchronize(); t2 = time.perf_counter()
            for _ in range(a.reps):
                cyc()
            s.sync(); torch.cuda.synchronize(); t3 = time.perf_counter()
            us = (t3 - t2) / a.reps * 1e6
            print("levels < %d: %.1f us%s" % (skip, us, "" if prev is None else "  (+%.1f)" % (us - prev)))
            prev = us
        os.environ.pop("MF_MG_DEVSKIP")
    b = Ct.c_int(0)
    check(lib().mf_mg_level_format(s._h, 1, Ct.byref(b)))
    print("size %d: %.1f us per V-cycle (incl. setRhs), level-1 operator %d bytes/entry" % (n, (t1 - t0) / a.reps * 1e6, b.value))


if __name__ == "__main__":
    main()
