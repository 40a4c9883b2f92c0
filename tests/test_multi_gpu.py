"""z-slab partitioned solver: the same cases as test_parity_gpu.py, executed with the grids split over
several ranks, against the same reference outputs (tests/golden/*.npz).

* emulated ranks (mf_comm_create_emulated): all ranks run inside one process on one GPU, launch group by launch group --
  exercises the z ranges of every kernel, the solo coarse levels of the multigrid and the cross-rank reductions on the 1-GPU box;
* real ranks: one process per GPU under torch.distributed.run (tools/dist_cases.py), CUDA-VMM global address space over NVLink and
  the device flag barrier; skipped with fewer than 2 GPUs.
Streaming kernels stay bit-exact (nothing about a cell's arithmetic depends on the partition); CG-based results keep the case's
tolerance (the dot products are summed per rank first, so the last bits of alpha/beta may differ from the single-GPU run).
"""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

import cases as C
import test_parity_gpu as T

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")


def _cases_for(nranks):
    # particle systems are replicated on every rank, rank 0 writes the particle -> grid transfers (csrc/particles.cu)
    return [n for n, c in C.CASES.items() if c.get("dim", 3) == 3 and c["size"][2] >= 2 * nranks]


def _run_partitioned(name, comm):
    from mantaflow_b200 import manta
    manta.setDefaultComm(comm)
    try:
        return T.run_ours(C.CASES[name])
    finally:
        manta.setDefaultComm(None)


@pytest.mark.gpu
@pytest.mark.parametrize("nranks", [2, 3])
@pytest.mark.parametrize("name", _cases_for(3))
def test_emulated_ranks_match_reference(name, nranks):
    from mantaflow_b200 import manta
    case = C.CASES[name]
    gold = np.load(os.path.join(GOLDEN, name + ".npz"))
    comm = manta.Comm.emulated(nranks)
    try:
        ins, out, infos = _run_partitioned(name, comm)
    finally:
        comm.destroy()
    errors = []
    rep = {}
    for g in case["save"]:
        try:
            T.compare(name, case, g, out[g], gold["out_" + g], rep)
        except AssertionError as e:
            errors.append(str(e))
    assert not errors, "; ".join(errors)
    if "mg" in name and name.startswith("solve") and len(infos) == len(gold["ref_iterations"]):
        assert [i["iterations"] for i in infos] == gold["ref_iterations"].tolist()


@pytest.mark.gpu
def test_emulated_slab_transfers():
    """mf_grid_upload_slab / mf_grid_download_slab move exactly the rank's planes"""
    import ctypes as Ct
    from mantaflow_b200 import manta
    from mantaflow_b200._lib import lib, check
    comm = manta.Comm.emulated(1)     # one rank: the slab is the whole grid
    try:
        s = manta.Solver(name="s", gridSize=manta.vec3(10, 6, 8), dim=3, comm=comm)
        k0, k1 = s.slab()
        assert (k0, k1) == (0, 8)
        v = s.create(manta.MACGrid)
        a = np.random.RandomState(3).rand(8, 6, 10, 3).astype(np.float32)
        check(lib().mf_grid_upload_slab(v._h, a.ctypes.data_as(Ct.c_void_p)))
        b = np.empty_like(a)
        check(lib().mf_grid_download_slab(v._h, b.ctypes.data_as(Ct.c_void_p)))
        assert np.array_equal(a, b)
        assert np.array_equal(v.to_numpy(), a)
    finally:
        comm.destroy()


def _ngpus():
    from mantaflow_b200._lib import lib
    return lib().mf_device_count()


@pytest.mark.gpu
def test_two_processes_two_gpus(tmp_path):
    """real multi-process path: torch.distributed.run, one rank per GPU"""
    if _ngpus() < 2:
        pytest.skip("needs 2 GPUs")
    out = tmp_path / "dist.json"
    names = ["adv_real_o2_c2", "adv_mac_o2_c2", "adv_mac_o2_outflow", "wallbcs", "buoyancy", "rhs_plain", "correct_vel",
             "solve_mic_tight", "solve_mgstatic_tight", "solve_mg_default", "solve_l2norm", "plume3d_mg", "flags_domain",
             "flip_transfer", "flip_advect_rk4", "flip01_3d"]
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29533", os.path.join(ROOT, "tools", "dist_cases.py"), "--out", str(out)] + names
    p = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, p.stdout[-3000:] + p.stderr[-3000:]
    rep = json.loads(out.read_text())
    assert rep["failed"] == [], rep
    assert sorted(rep["passed"]) == sorted(names)


@pytest.mark.gpu
@pytest.mark.parametrize("nranks", [1, 2])
def test_streaming_transfers(nranks):
    """mf_grid_stage_upload / _commit_upload / _download_async: a pipelined loop returns, per step, exactly what a synchronous
    upload -> kernel -> download returns"""
    import ctypes as Ct
    from mantaflow_b200 import manta
    from mantaflow_b200._lib import lib, check
    L = lib()
    size = (12, 10, 8)
    comm = manta.Comm.emulated(nranks)
    bufs = []

    def pinned(shape):
        n = int(np.prod(shape)) * 4
        p = L.mf_host_alloc(n)
        assert p
        bufs.append(p)
        return p, np.ctypeslib.as_array((Ct.c_byte * n).from_address(p)).view(np.float32).reshape(shape)
    try:
        s = manta.Solver(name="s", gridSize=manta.vec3(*size), dim=3, comm=comm)
        v = s.create(manta.MACGrid); d = s.create(manta.RealGrid)
        rs = np.random.RandomState(1)
        steps = 5
        vin = [rs.rand(size[2], size[1], size[0], 3).astype(np.float32) for _ in range(steps)]
        din = [rs.rand(size[2], size[1], size[0]).astype(np.float32) for _ in range(steps)]
        hv = [pinned(vin[0].shape) for _ in range(2)]
        hd = [pinned(din[0].shape) for _ in range(2)]
        ov = [pinned(vin[0].shape) for _ in range(steps)]
        od = [pinned(din[0].shape) for _ in range(steps)]
        hv[0][1][...] = vin[0]; hd[0][1][...] = din[0]
        v.stageUpload(hv[0][0]); d.stageUpload(hd[0][0])
        for i in range(steps):
            v.commitUpload(); d.commitUpload()
            if i + 1 < steps:
                if i >= 1:
                    s.sync()          # the host buffer about to be refilled was consumed by the commit two steps ago
                hv[(i + 1) & 1][1][...] = vin[i + 1]; hd[(i + 1) & 1][1][...] = din[i + 1]
                v.stageUpload(hv[(i + 1) & 1][0]); d.stageUpload(hd[(i + 1) & 1][0])
            v.multConst(manta.vec3(2.0, 3.0, 4.0)); d.multConst(0.5)
            v.downloadAsync(ov[i][0]); d.downloadAsync(od[i][0])
        s.transfersSync()
        for i in range(steps):
            assert np.array_equal(ov[i][1], vin[i] * np.array([2.0, 3.0, 4.0], np.float32)), i
            assert np.array_equal(od[i][1], din[i] * np.float32(0.5)), i
        # staging more than two uploads without a commit is refused
        v.stageUpload(hv[0][0]); v.stageUpload(hv[1][0])
        assert L.mf_grid_stage_upload(v._h, Ct.c_void_p(hv[0][0])) != 0
        v.commitUpload(); v.commitUpload()
        assert L.mf_grid_commit_upload(v._h) != 0
        s.transfersSync()
        del v, d, s
    finally:
        comm.destroy()
        for p in bufs:
            L.mf_host_free(p)
