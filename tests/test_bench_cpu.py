"""bench.py's reference arm on CPU (no GPU needed): it times the unmodified reference at the size it prints (the arm
used to shrink the grid silently), alone on rank 0 under a 2-process launch, with the contract's keys."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
MANTA_REF = os.path.join(ROOT, "oracle", "_ref", "manta")


def _lines(out):
    return [json.loads(l) for l in out.splitlines() if l.startswith("{")]


@pytest.mark.skipif(not os.path.exists(MANTA_REF), reason="oracle/_ref/manta not built")
def test_reference_arm_runs_the_size_it_prints():
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--size", "32", "--steps", "2", "--warmup", "1"],
                       capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert p.returncode == 0, p.stderr[-2000:]
    (d,) = _lines(p.stdout)
    assert d["impl"] == "reference" and d["steps"] == 2 and d["warmup"] == 1 and d["higher_is_better"] is True
    assert d["config"]["grid"] == [32, 32, 32] and "32x32x32" in d["config"]["workload"] and "sample_grid" not in d["config"]
    assert "32x32x32 plume step x2" in d["cpu_baseline"]["sample"] and d["cpu_baseline"]["kind"] == "reference"
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert abs(d["value"] - 32 ** 3 / (d["ms_per_step"] * 1e-3)) < 1e-6 * d["value"]


@pytest.mark.skipif(not os.path.exists(MANTA_REF), reason="oracle/_ref/manta not built")
def test_reference_arm_under_a_two_process_launch():
    """N > 1 (weak scaling): rank 0 alone runs and prints; the line names the bounded sample it really ran (one GPU's slab), not the job"""
    p = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                        "--master-port", "29577", os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--size", "32",
                        "--steps", "1", "--warmup", "1"], capture_output=True, text=True, timeout=900, cwd=ROOT)
    assert p.returncode == 0, p.stderr[-2000:]
    (d,) = _lines(p.stdout)
    assert d["n_gpus"] == 2 and d["config"]["grid"] == [32, 32, 32] and d["config"]["job_grid"] == [32, 32, 64]
    assert "bounded sample" in d["config"]["workload"] and "32x32x32" in d["config"]["workload"]


def test_algo_bytes_follow_the_operator_format():
    """roofline bookkeeping: a 27-point colour sweep moves 27 coefficients of the level's entry size + x, b, x-out per vertex of
    its colour; level sizes follow the reference's coarsening rule (multigrid.cpp:256-276); a level whose operator and vectors fit
    the L2 is labelled latency-bound, not given an HBM fraction"""
    sys.path.insert(0, ROOT)
    import bench
    sizes = bench.mg_level_sizes((256, 256, 256))
    assert [tuple(s) for s in sizes[:3]] == [(256, 256, 256), (129, 129, 129), (65, 65, 65)]
    assert sizes[-1][0] * sizes[-1][1] * sizes[-1][2] <= 1000
    nv1 = 129 ** 3
    b4, bound4 = bench.algo_bytes("k_mg_smooth@L1", (256, 256, 256), 256 ** 3, {1: 4})
    b2, bound2 = bench.algo_bytes("k_mg_smooth@L1", (256, 256, 256), 256 ** 3, {1: 2})
    assert b4 == (27 * 4 + 12) / 8.0 * nv1 and bound4 == "hbm"
    assert b2 == (27 * 2 + 12) / 8.0 * nv1 and bound2 in ("hbm", "latency")
    b, bound = bench.algo_bytes("k_mg_smooth@L3", (256, 256, 256), 256 ** 3, {})
    assert bound == "latency" and b == (27 * 4 + 12) / 8.0 * 33 ** 3
    assert bench.algo_bytes("k_mc_mac_f", (256, 256, 256), 256 ** 3)[0] == 40.0 * 256 ** 3
    assert bench.algo_bytes("k_not_a_kernel", (256, 256, 256), 256 ** 3) == (None, None)
