"""Runs parity cases of tests/cases.py on a z-slab partitioned solver, one process per GPU (launch with torch.distributed.run).
Every rank executes the same scene calls (the library is SPMD); every rank downloads the full grids and compares them with the
reference outputs in tests/golden/; rank 0 writes the report.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 \
      tools/dist_cases.py --out gpurun_out/dist_cases.json adv_real_o2_c2 solve_mgstatic_tight plume3d_mg
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="")
    ap.add_argument("names", nargs="+")
    a = ap.parse_args()
    import cases as C
    import test_parity_gpu as T
    from mantaflow_b200 import manta
    comm = manta.Comm.from_env()
    manta.setDefaultComm(comm)
    passed, failed, detail = [], [], {}
    for name in a.names:
        case = C.CASES[name]
        gold = np.load(os.path.join(ROOT, "tests", "golden", name + ".npz"))
        rep = {}
        try:
            ins, out, infos = T.run_ours(case)
            for g in case["save"]:
                T.compare(name, case, g, out[g], gold["out_" + g], rep)
            rep["iterations"] = [i["iterations"] for i in infos]
            passed.append(name)
        except Exception as e:  # noqa: BLE001  (report and keep the ranks in lock-step)
            rep["error"] = "%s: %s" % (type(e).__name__, e)
            failed.append(name)
        detail[name] = rep
        if comm.rank == 0:
            print("rank0 %-28s %s" % (name, "ok" if name in passed else rep.get("error")), flush=True)
    manta.setDefaultComm(None)
    if a.out:
        path = a.out if comm.rank == 0 else a.out + ".rank%d" % comm.rank
        with open(path, "w") as f:
            json.dump({"rank": comm.rank, "world": comm.nranks, "passed": passed, "failed": failed, "detail": detail}, f, indent=1)
    import gc
    gc.collect()
    comm.destroy()
    sys.exit(1 if failed else 0)


if __name__ == "__main__":
    main()
