"""Double-precision ground truth (VERDICT r1 missing 7): the reference built with -DFLOATINGPOINT_PRECISION=2
(oracle/_ref/mantad, `DOUBLE=1 oracle/build_ref.sh`) runs PcMIC cases in double; against that truth the error of
  * the reference's own float build (oracle/_ref/manta),
  * this library on one GPU (the reference's sequential MIC(0): bit-identical to the float reference),
  * this library on 2 emulated z-slab ranks (the red-black IC stand-in that partitioned solvers use)
is reported per grid, so that "GPU-float error <= CPU-float error" can be read off for the red-black path.

    python tools/double_oracle.py > profiles/<tag>_double_oracle.json          (on the GPU box)
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import cases as C            # noqa: E402
import make_golden as MG     # noqa: E402
import test_parity_gpu as T  # noqa: E402


def main():
    from mantaflow_b200 import manta
    mantad = os.path.join(ROOT, "oracle", "_ref", "mantad")
    out = {}
    for name in ("solve_mic_tight", "solve4_mic_tight", "plume3d_mic", "simpleplume_noise", "plume3d_open_mic"):
        case = C.CASES[name]
        truth = MG.run_reference(name, case, binary=mantad, real_bytes=8)
        ref32 = MG.run_reference(name, case)
        _, one, info1 = T.run_ours(case)
        comm = manta.Comm.emulated(2)
        manta.setDefaultComm(comm)
        try:
            _, two, info2 = T.run_ours(case)
        finally:
            manta.setDefaultComm(None)
            comm.destroy()
        rep = {"iterations": {"double": truth["ref_iterations"].tolist(), "reference_float": ref32["ref_iterations"].tolist(),
                              "gpu_1": [i["iterations"] for i in info1], "gpu_2_ranks_redblack": [i["iterations"] for i in info2]}}
        for g in case["save"]:
            if case["grids"][g] in ("FlagGrid", "IntGrid"):
                continue
            t = truth["out_" + g].astype(np.float64)
            scale = float(np.abs(t).max()) or 1.0
            err = lambda a: float(np.abs(a.astype(np.float64) - t).max())      # noqa: E731
            rep[g] = {"max_abs_truth": scale, "err_reference_float": err(ref32["out_" + g]), "err_gpu_1": err(one[g]),
                      "err_gpu_2_ranks_redblack": err(two[g]), "gpu_1_equals_reference_float": bool(np.array_equal(one[g], ref32["out_" + g]))}
        out[name] = rep
        sys.stderr.write("%s done\n" % name)
    json.dump(out, sys.stdout, indent=1)


if __name__ == "__main__":
    main()
