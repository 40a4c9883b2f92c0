#!/usr/bin/env python
"""Generate tests/golden/*.npz by running every case of tests/cases.py through the UNMODIFIED reference binary
(oracle/_ref/manta, built from /root/reference by oracle/build_ref.sh).  Run in the build container only:

    python tests/golden/make_golden.py [case-name-prefix ...]

Each npz holds the reference's output grids (out_<grid>), the crc32 of the regenerated inputs (in_crc), and the
'Iterations:' / 'residual:' lines the reference printed at debug level 2 (pressure.cpp:442).
"""
import json
import os
import re
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
TESTS = os.path.dirname(HERE)
ROOT = os.path.dirname(TESTS)
sys.path.insert(0, TESTS)
import cases as C          # noqa: E402
import scene_runner as sr  # noqa: E402

MANTA = os.path.join(ROOT, "oracle", "_ref", "manta")


def run_reference(name, case, threads="1", binary=None, real_bytes=4):
    """binary: oracle/_ref/manta (default) or oracle/_ref/mantad, the double-precision build (-DFLOATINGPOINT_PRECISION=2, real_bytes=8:
    its .raw files hold doubles; inputs are written as doubles too)"""
    ins = C.inputs_for(case)
    with tempfile.TemporaryDirectory() as d:
        json.dump(case, open(os.path.join(d, "case.json"), "w"))
        for k, a in ins.items():
            sr.write_raw(os.path.join(d, "in_%s.raw" % k), a.astype(np.float64) if (real_bytes == 8 and a.dtype == np.float32) else a)
        env = dict(os.environ, MF_CASE_DIR=d, MF_TESTS_DIR=TESTS, OMP_NUM_THREADS=threads, MANTA_DISABLE_UI="1")
        p = subprocess.run([binary or MANTA, os.path.join(HERE, "ref_driver.py")], env=env, cwd=d, capture_output=True, text=True, timeout=1800)
        if p.returncode != 0 or "Error" in p.stdout or not os.path.exists(os.path.join(d, "log.json")):
            print(p.stdout[-3000:], p.stderr[-3000:])
            raise RuntimeError("reference failed on case %s" % name)
        out = {}
        for g in case["save"]:
            gt = case["grids"][g]
            if sr.is_particle_type(gt):
                out["out_" + g] = sr.read_pdata_uni(os.path.join(d, "out_%s.uni" % g), 1 if gt.startswith("PdataReal") else 3)
                continue
            dt = np.int32 if gt in ("FlagGrid", "IntGrid") else (np.float64 if real_bytes == 8 else np.float32)
            out["out_" + g] = sr.read_raw(os.path.join(d, "out_%s.raw" % g), dt, sr.grid_shape(case, gt))
        its = [int(m) for m in re.findall(r"solvePressure done\. Iterations:(\d+)", p.stdout)]
        res = [float(m) for m in re.findall(r"solvePressure done\. Iterations:\d+, residual:([-+0-9.eE]+|nan|inf)", p.stdout)]
        out["ref_iterations"] = np.array(its, np.int32)
        out["ref_residuals"] = np.array(res, np.float64)
        out["in_crc"] = np.array([C.checksum(ins)], np.uint32)
        return out


def main():
    if not os.path.exists(MANTA):
        sys.exit("oracle/_ref/manta missing: run oracle/build_ref.sh first")
    sel = sys.argv[1:]
    for name, case in C.CASES.items():
        if sel and not any(name.startswith(s) for s in sel):
            continue
        out = run_reference(name, case)
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
        print("%-32s its=%s" % (name, out["ref_iterations"].tolist()[:20]))


if __name__ == "__main__":
    main()
