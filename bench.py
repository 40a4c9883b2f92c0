#!/usr/bin/env python
"""bench.py -- throughput of the Eulerian grid hot path (advect + project) on synthetic smoke-plume grids.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--size 256] [--precond mg|mic] [--impl ours|reference]

One "step" = one pass of the hot path over the whole grid, exactly the per-frame loop of scenes/simpleplume.py:43-55
with the noise source replaced by Cylinder.applyToGrid:

    source.applyToGrid(density, 1); advectSemiLagrange(density, order=2); advectSemiLagrange(vel, order=2);
    setWallBcs; addBuoyancy; solvePressure(preconditioner=PcMGStatic | PcMIC, zeroPressureFixing=True)

Metric (BASELINE.json): cell-updates/s per full step = cells * steps / time.  Workload at N=1: BASELINE.json configs[1],
the 256^3 smoke plume with MacCormack advection and GridMg-preconditioned PCG.

Prints ONE JSON line (rank 0).  Keys beyond the base contract:
  roofline      dominant kernel: algorithmic bytes per launch / CUDA-event duration, against MEASURED_PEAKS.json
  cpu_baseline  the UNMODIFIED reference (oracle/_ref/manta, OpenMP) timed on this box's host cores on a bounded sample
  e2e           same metric through the C-ABI with HOST (pinned) buffers: per step, upload vel+density, run the step,
                download vel+density+pressure
  kernels       per-kernel share of a step from CUDA events (mf_solver_profile_*), for DESIGN.md / profiles/
  strong_1024   BASELINE configs[4]: the 1024^3 plume on the N GPUs of this run (strong scaling), >= 5 timed steps, and the same
                grid on ONE GPU of this run, so that the speed-up lands in the driver's own SCALE file
  parity        (N > 1) the golden case plume3d_mg and one partitioned step of the bench workload, executed by the N real
                processes before the timed region and compared with tests/golden/ resp. the emulated-rank result
`--impl reference` times the reference's own CPU implementation of the same step on the host cores, at the size it prints.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "cell-updates/sec per full step (advect+project)"
UNIT = "cell-updates/s"

# algorithmic HBM bytes per cell and launch (perfect neighbour reuse), DESIGN.md "kernels"; SURVEY.md 8(d)
ALGO_BYTES_PER_CELL = {
    "k_sl": 20.0,              # vel 12 + src 4 + dst 4
    "k_mc_centered": 28.0,     # vel 12 + fwd 4 + orig 4 + flags 4 + dst 4
    "k_sl_mac": 24.0,          # self-advection: vel(=src) 12 + dst 12
    "k_mc_mac": 40.0,          # vel(=orig) 12 + fwd 12 + flags 4 + dst 12
    "k_sl_f": 20.0, "k_mc_centered_f": 28.0, "k_sl_mac_f": 24.0, "k_mc_mac_f": 40.0,   # the 32-bit-index kernels (advect_fast.cuh)
    "k_set_wall_bcs": 4.0,     # streams the flags; velocities are only touched in the thin shell where a wall condition applies
    "k_add_buoyancy": 21.4,    # reference sweep: flags 4 + density 4 + vel 12 r + 12 w = 32; the kernel reads / stores a velocity component only
                               # where the force has one: 359.6 MB per launch measured (ncu) = 21.4 B/cell -- the figure the fraction is taken on
    "k_wall_buoyancy": 32.0,
    "k_make_rhs": 20.0,        # flags 4 + vel 12 + rhs 4
    "k_make_rhs_plain": 20.0,
    "k_flag_stats": 4.0,
    "k_make_laplace_code": 6.0,  # flags 4 + row code 2
    "k_make_laplace_code_stats": 6.0,  # the same pass also counts the empty cells / finds the first fluid cell (zeroPressureFixing)
    "k_make_laplace": 22.0,    # flags 4 + 4 coefficient planes 16 + row code 2 (only when a preconditioner is (re)built)
    "k_cg_spmv_dot": 28.0,     # stored-A: flags 4 + src 4 + dst 4 + A 16
    "k_cg_spmv_dot_code": 10.0,  # 16-bit row code 2 + src 4 + dst 4
    "k_cg_update": 24.0,       # x r/w 8 + search 4 + residual r/w 8 + tmp 4
    "k_cg_dot_beta": 8.0, "k_cg_dot_sigma": 8.0,
    "k_cg_search": 12.0,
    "k_correct_velocity": 32.0,  # flags 4 + p 4 + vel 12 r + 12 w
    # level-0 multigrid sweeps with the 16-bit operator code (sector granularity: a colour sweep touches every sector of the
    # arrays it reads/writes): code 2 + x in 4 + b 4 + x out 4 = 14; the first sweep of a V-cycle (x == 0) has no x in: 10;
    # 3 of the 4 sweeps per V-cycle are of the first kind -> 13 on average
    "k_mg_smooth0": 13.0,
    "k_mg_residual0": 14.0,    # code 2 + x 4 + b 4 + r 4
    "k_mg_presmooth0": 10.0,   # first sweep pair of a V-cycle (x == 0): code 2 + b 4 + x out 4
    "k_mg_up0": 14.5,          # interpolate + add + both colour sweeps in one pass: code 2 + x 4 + b 4 + coarse x 0.5 + x out 4
    "k_mg_up0w": 14.5,         # the same as a z-marching wavefront (mg_wave0.cuh)
    "k_mg_down0": 10.5,        # both colours + residual + restriction in one pass: code 2 + b 4 + x out 4 + coarse b 0.5
    # k_mg_restrict / k_mg_interp_add / k_mg_smooth / k_mg_residual run on every level: no single bytes-per-cell figure
    "k_apply_cylinder": 4.0,
}


def mg_level_sizes(size):
    """sizes of the GridMg levels (multigrid.cpp:256-261: (size + 2) / 2 per axis until every axis is <= 5 or the level has <= 1000
    vertices); only used to turn per-level kernel times into bytes"""
    out = [tuple(size)]
    while max(out[-1]) > 5 and out[-1][0] * out[-1][1] * out[-1][2] > 1000 and len(out) < 101:     # multigrid.cpp:256-261
        out.append(tuple((n + 2) // 2 for n in out[-1]))
    return out


L2_BYTES = 126e6


def algo_bytes(kernel, size_gpu, cells_gpu, entry_bytes=None):
    """algorithmic bytes ONE launch of `kernel` moves and the roofline that bounds it.
    Multigrid kernels on the 27-point levels are tagged with their level ("k_mg_smooth@L2"): a colour sweep reads, per vertex of
    its colour, the 27 coefficients of the row, x of the 27 stencil points (perfect reuse: one x per vertex), b, and writes x;
    a level whose operator + vectors fit the L2 never reaches HBM in steady state: bound "latency".
    entry_bytes: {level: bytes per stored operator entry} as the library reports them (mf_mg_level_format: 4 = float planes,
    2 = binary16 planes, the format level 1 takes when its entries convert exactly or the exceptions fit the float-row lists)."""
    base, _, lv = kernel.partition("@L")
    if base in ALGO_BYTES_PER_CELL and (lv in ("", "0")):
        return ALGO_BYTES_PER_CELL[base] * cells_gpu, "hbm"
    if lv == "":
        return None, None
    lv = int(lv)
    sizes = mg_level_sizes(size_gpu)
    if lv >= len(sizes):
        return None, None
    nv = sizes[lv][0] * sizes[lv][1] * sizes[lv][2]
    eb = (entry_bytes or {}).get(lv, 4)
    level_bytes = nv * (14 * eb + 3 * 4)
    bound = "hbm" if level_bytes > L2_BYTES else "latency"
    per_vertex = {"k_mg_smooth": (27 * eb + 12) / 8.0,        # one of 8 colours per launch
                  "k_mg_smooth_ml": (27 * eb + 12) / 8.0,
                  "k_mg_residual": 27 * eb + 12, "k_mg_restrict": 4 + 0.5, "k_mg_interp_add": 8 + 0.5, "k_mg_interp_add_b": 8 + 0.5, "k_mg_interp_add_c": 8 + 0.5,
                  "k_mg_level_sweep": 14 * eb + 12}.get(base)
    if per_vertex is None:
        return None, None
    return per_vertex * nv, bound


def _parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--size", type=int, default=256, help="grid edge (BASELINE configs[1]: 256); with N GPUs the grid is size x size x (size*N) "
                                                       "(weak scaling, a size^3 z-slab per GPU) unless --strong")
    ap.add_argument("--strong", action="store_true", help="N GPUs share ONE size^3 grid (BASELINE configs[4]: --size 1024 --strong)")
    ap.add_argument("--precond", default="mg", choices=["mg", "mic"])
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--code-cache", action="store_true", help="let solvePressure reuse the CG's Laplace row codes while the flag grid is unchanged "
                    "(library default; the bench rebuilds the matrix every step like the reference does, so that no work of the step is skipped)")
    ap.add_argument("--profile-steps", type=int, default=3)
    ap.add_argument("--no-strong1024", action="store_true", help="skip the BASELINE configs[4] leg (1024^3 strong scaling)")
    ap.add_argument("--no-parity", action="store_true", help="skip the real-rank parity leg at N > 1")
    ap.add_argument("--kernels-out", default="", help="write the complete per-kernel table (JSON) to this file")
    return ap.parse_args()


# ---------------------------------------------------------------------------------------------------------------
# clocks: sampled DURING the timed region (B200_PROFILING.md "clocks line"), via NVML
class ClockSampler(object):
    def __init__(self, index):
        self.samples, self.reasons, self.stop_flag, self.thread, self.max_mhz = [], set(), False, None, None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _run(self):
        nv = self.nv
        names = {
            "hw_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8),
            "hw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40),
            "sw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20),
            "sw_power_cap": getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4),
            "hw_power_brake_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwPowerBrakeSlowdown", 0x80),
        }
        while not self.stop_flag:
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            time.sleep(0.02)

    def start(self):
        if self.nv:
            self.stop_flag = False
            self.thread = threading.Thread(target=self._run, daemon=True)
            self.thread.start()

    def stop(self):
        if self.thread:
            self.stop_flag = True
            self.thread.join()
        s = sorted(self.samples)
        return {"sm_mhz": (s[len(s) // 2] if s else None), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


# ---------------------------------------------------------------------------------------------------------------
def run_reference_binary(size, steps, warmup, pc, threads=None, timeout=3000):
    """Times the unmodified reference (oracle/_ref/manta) on the host cores.  Returns (cells/s, info dict)."""
    manta = os.path.join(ROOT, "oracle", "_ref", "manta")
    if not os.path.exists(manta):
        raise RuntimeError("oracle/_ref/manta missing (built by __graft_entry__.build() where /root/reference exists)")
    threads = threads or os.cpu_count() or 1
    env = dict(os.environ, MF_BENCH_SIZE="%d,%d,%d" % tuple(size), MF_BENCH_STEPS=str(steps), MF_BENCH_WARMUP=str(warmup),
               MF_BENCH_PC=str(pc), OMP_NUM_THREADS=str(threads), OMP_PROC_BIND="close", MANTA_DISABLE_UI="1")
    p = subprocess.run([manta, os.path.join(ROOT, "oracle", "bench_scene.py")], env=env, capture_output=True, text=True,
                       timeout=timeout, cwd=os.path.join(ROOT, "oracle", "_ref"))
    line = [l for l in p.stdout.splitlines() if l.startswith("BENCH_JSON ")]
    if p.returncode != 0 or not line:
        raise RuntimeError("reference bench scene failed: " + (p.stdout[-1500:] + p.stderr[-1500:]))
    d = json.loads(line[-1][len("BENCH_JSON "):])
    timed = d["step_s"][warmup:]
    cells = size[0] * size[1] * size[2]
    return cells * len(timed) / sum(timed), {"step_s": timed, "threads": threads}


def cpu_model():
    try:
        for l in open("/proc/cpuinfo"):
            if l.startswith("model name"):
                return l.split(":", 1)[1].strip()
    except OSError:
        pass
    return "unknown"


# ---------------------------------------------------------------------------------------------------------------
class Plume(object):
    """The synthetic smoke plume of SURVEY.md 8(d) on one GPU, driven through mantaflow_b200.manta (C-ABI -> CUDA)."""

    def __init__(self, size, device, precond, comm=None):
        from mantaflow_b200 import manta as M
        from mantaflow_b200._lib import lib, check
        self.M, self.L, self.check = M, lib(), check
        nx, ny, nz = size
        gs = M.vec3(nx, ny, nz)
        self.s = M.Solver(name="main", gridSize=gs, dim=3, device=device, comm=comm)
        self.s.timestep = 1.0
        self.flags = self.s.create(M.FlagGrid)
        self.vel = self.s.create(M.MACGrid)
        self.density = self.s.create(M.RealGrid)
        self.pressure = self.s.create(M.RealGrid)
        self.flags.initDomain(boundaryWidth=0)
        self.flags.fillGrid()
        self.source = M.Cylinder(parent=self.s, center=gs * M.vec3(0.5, 0.1, 0.5), radius=nx * 0.14, z=gs * M.vec3(0, 0.02, 0))
        self.pc = M.PcMGStatic if precond == "mg" else M.PcMIC
        self.iterations = []
        self.solve_ms = None

    def step(self):
        M = self.M
        self.source.applyToGrid(grid=self.density, value=1)
        M.advectSemiLagrange(flags=self.flags, vel=self.vel, grid=self.density, order=2)
        M.advectSemiLagrange(flags=self.flags, vel=self.vel, grid=self.vel, order=2)
        M.setWallBcs(flags=self.flags, vel=self.vel)
        M.addBuoyancy(density=self.density, vel=self.vel, gravity=M.vec3(0, -6e-4, 0), flags=self.flags)
        if self.solve_ms is not None:      # PCG iteration time (SURVEY 8d metric ii): events around the projection only
            self.check(self.L.mf_solver_event_record(self.s._h, 6))
        M.solvePressure(flags=self.flags, vel=self.vel, pressure=self.pressure, preconditioner=self.pc, zeroPressureFixing=True)
        if self.solve_ms is not None:
            self.check(self.L.mf_solver_event_record(self.s._h, 7))
            ms = C.c_float()
            self.s.sync()
            self.check(self.L.mf_solver_event_elapsed_ms(self.s._h, 6, 7, C.byref(ms)))
            self.solve_ms.append(ms.value)
        self.s.step()
        self.iterations.append(M.getSolveInfo()["iterations"])


def workload_text(size, pcname):
    return ("smoke plume %dx%dx%d, MacCormack (order 2) advection of density+velocity, setWallBcs, addBuoyancy, "
            "solvePressure %s, cgAccuracy 1e-3, zeroPressureFixing" % (tuple(size) + (pcname,)))


def time_steps(P, L, check, steps, barrier, max_over_ranks, ev=(0, 1)):
    """K steps between two events on the solver stream, barrier + sync on both sides; returns (device seconds max over ranks, wall seconds)"""
    s = P.s
    barrier(); s.sync()
    check(L.mf_solver_event_record(s._h, ev[0]))
    t0 = time.perf_counter()
    for _ in range(steps):
        P.step()
    check(L.mf_solver_event_record(s._h, ev[1]))
    s.sync()
    wall = time.perf_counter() - t0
    barrier()
    ms = C.c_float()
    check(L.mf_solver_event_elapsed_ms(s._h, ev[0], ev[1], C.byref(ms)))
    return max_over_ranks(ms.value * 1e-3), wall


def parity_leg(M, comm, world, rank, local, size, precond, gather):
    """Real-rank correctness inside the driver-run artefact (N > 1): (1) the golden case plume3d_mg (15 steps, reference output
    committed under tests/golden/) on the N real processes; (2) two steps of the bench workload on the N real processes against
    the same two steps on N EMULATED ranks inside rank 0's GPU (same kernels and reductions, barriers = stream order): per-slab
    CRC32 of density / velocity / pressure must agree."""
    import zlib
    import numpy as np
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import cases as CS
    import test_parity_gpu as T
    out = {}
    name = "plume3d_mg"
    case = CS.CASES[name]
    gold = np.load(os.path.join(ROOT, "tests", "golden", name + ".npz"))
    M.setDefaultComm(comm)
    try:
        ins, res, infos = T.run_ours(case)
    finally:
        M.setDefaultComm(None)
    d = {g: float(np.abs(res[g].astype(np.float64) - gold["out_" + g].astype(np.float64)).max()) for g in case["save"]}
    out.update({"case": name, "max_abs_diff": max(d.values()), "per_grid": d, "bit_identical": all(np.array_equal(res[g], gold["out_" + g]) for g in case["save"]),
                "iterations": [i["iterations"] for i in infos], "ref_iterations": gold["ref_iterations"].tolist(),
                "thresholds": case["abs_tol"]})
    out["ok"] = bool(all(d[g] <= case["abs_tol"][g] for g in case["save"]) and out["iterations"] == out["ref_iterations"])
    # (2) one partitioned bench-size step pair, real ranks vs emulated ranks
    def crcs(P, k0, k1):
        return [zlib.crc32(np.ascontiguousarray(g.to_numpy()[k0:k1]).tobytes()) for g in (P.density, P.vel, P.pressure)]
    P = Plume(size, local, precond, comm)
    for _ in range(2):
        P.step()
    k0, k1 = P.s.slab()
    mine = crcs(P, k0, k1)
    its = list(P.iterations)
    slabs = [P.s.slab(r) for r in range(world)]
    del P
    gc_collect()
    allc = gather(mine)
    emu = None
    if rank == 0:
        ec = M.Comm.emulated(world, local)
        try:
            Pe = Plume(size, local, precond, ec)
            for _ in range(2):
                Pe.step()
            emu = [crcs(Pe, a, b) for (a, b) in slabs]
            emu_its = list(Pe.iterations)
            del Pe
            gc_collect()
        finally:
            ec.destroy()
        out["step_vs_emulated"] = {"grid": list(size), "steps": 2, "slabs_identical": bool(emu == allc), "iterations": its, "emulated_iterations": emu_its,
                                   "what": "CRC32 per z-slab of density, velocity, pressure: N real processes vs N emulated ranks on one GPU"}
        out["ok"] = bool(out["ok"] and emu == allc and its == emu_its)
    return out


def gc_collect():
    import gc
    gc.collect()


def strong_leg(M, L, check, comm, world, rank, local, precond, barrier, max_over_ranks, steps=5, warm=2):
    """BASELINE configs[4]: the 1024^3 plume, z-slab partitioned over the N GPUs of this run (strong scaling), and on ONE GPU."""
    size = (1024, 1024, 1024)
    cells = size[0] * size[1] * size[2]
    out = {"grid": list(size), "steps": steps, "warmup": warm}
    def state_sums(P):      # size-independent check of the state after the warm-up steps: sum of squares of density and pressure
        out = {}
        for name, g in (("density", P.density), ("pressure", P.pressure)):
            v = C.c_double()
            check(L.mf_grid_sum_sqr(g._h, C.byref(v)))
            out[name + "_sumsqr"] = v.value
        out["pcg_iterations"] = list(P.iterations)
        return out

    one = None
    sums1 = None
    if rank == 0:     # the single-GPU time of the same grid, measured in this very run (needs ~110 GB of the 180 GB)
        try:
            P1 = Plume(size, local, precond, None)
            for _ in range(warm):
                P1.step()
            sums1 = state_sums(P1)
            n1 = 3 if world > 1 else steps
            t1, _ = time_steps(P1, L, check, n1, lambda: None, lambda x: x, ev=(8, 9))
            one = {"ms_per_step": 1e3 * t1 / n1, "value": cells * n1 / t1, "steps": n1, "pcg_iterations": P1.iterations[-n1:]}
            del P1
            gc_collect()
        except Exception as e:      # noqa: BLE001
            one = {"error": str(e)[:300]}
    if world == 1:
        out.update(one or {})
        out["speedup_vs_1gpu"] = 1.0 if one and "value" in one else None
        out["n_gpus"] = 1
        return out
    barrier()
    P = Plume(size, local, precond, comm)
    for _ in range(warm):
        P.step()
    sumsN = state_sums(P)       # collective: every rank calls it
    if sums1 is not None:
        rel = {k: abs(sumsN[k] - sums1[k]) / max(abs(sums1[k]), 1e-300) for k in ("density_sumsqr", "pressure_sumsqr")}
        out["check"] = {"after_steps": warm, "one_gpu": sums1, "n_gpus": sumsN, "rel_diff": rel,
                        "ok": bool(max(rel.values()) < 1e-4 and sums1["pcg_iterations"] == sumsN["pcg_iterations"]),
                        "what": "sum of squares of density / pressure after the warm-up steps: 1024^3 on one GPU vs z-slab partitioned over the N GPUs "
                                "(the partitioned PCG sums its dot products per rank first: equal to rounding, not bit for bit)"}
    t, _ = time_steps(P, L, check, steps, barrier, max_over_ranks, ev=(8, 9))
    out.update({"n_gpus": world, "ms_per_step": 1e3 * t / steps, "value": cells * steps / t, "pcg_iterations": P.iterations[-steps:], "one_gpu": one,
                "speedup_vs_1gpu": (one["ms_per_step"] / (1e3 * t / steps)) if one and "ms_per_step" in one else None})
    del P
    gc_collect()
    return out


def main():
    a = _parse()
    if not a.code_cache:
        os.environ["MF_CG_NOCODECACHE"] = "1"     # every step builds its matrix, as the reference's solvePressureSystem does
    os.environ["MF_PROFILE_LEVELS"] = "1"         # the per-kernel profile names multigrid kernels per level (k_mg_smooth@L1 ...)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    strong = a.strong or world == 1
    size = (a.size, a.size, a.size if strong else a.size * world)
    cells = size[0] * size[1] * size[2]            # whole job
    pcname = "GridMg-preconditioned PCG (PcMGStatic)" if a.precond == "mg" else "MIC(0)-preconditioned PCG (PcMIC)"
    config = {"workload": workload_text(size, pcname),
              "grid": list(size), "cells_per_gpu": cells // world, "dt": 1.0,
              "matrix": "row codes reused while the flags are unchanged" if a.code_cache else "rebuilt every step (MF_CG_NOCODECACHE=1)",
              "l2": "working set per step > 1 GiB (>> 126 MB L2), no flush needed" if a.size >= 256 else "L2-resident working set",
              "parallelism": ("z-slab x%d (one global address space over NVLink, device flag barrier per launch group), %s scaling"
                              % (world, "strong" if a.strong else "weak")) if world > 1 else "single GPU"}

    if a.impl == "reference":
        if rank != 0:
            return 0
        pc = 3 if a.precond == "mg" else 1
        # the reference's CPU step is timed at the size it prints.  N = 1: the bench configuration itself.  N > 1 (weak scaling):
        # the N x taller grid would take N x 2.7 s per step on the host, so ONE GPU's share (size^3) is timed as the bounded sample
        # and the line SAYS so: workload and grid name the sample, not the N-GPU job.
        ref_size = (a.size, a.size, a.size)
        if tuple(ref_size) != tuple(size):
            config = dict(config, workload=workload_text(ref_size, pcname) + " [bounded sample: one GPU's z-slab of the %dx%dx%d job]" % size,
                          grid=list(ref_size), job_grid=list(size), cells_per_gpu=ref_size[0] * ref_size[1] * ref_size[2],
                          parallelism="host cores only (OpenMP)")
        else:
            config = dict(config, parallelism="host cores only (OpenMP)")
        t0 = time.time()
        v, info = run_reference_binary(ref_size, a.steps, a.warmup, pc)
        wall = time.time() - t0
        sample = "%dx%dx%d plume step x%d (+%d warm-up), reference binary oracle/_ref/manta (OpenMP, -O3, float), %d threads on %s" % (
            ref_size + (a.steps, a.warmup, info["threads"], cpu_model()))
        out = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup,
               "ms_per_step": 1e3 * sum(info["step_s"]) / len(info["step_s"]), "higher_is_better": True, "scaling": "weak",
               "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config,
               "cpu_baseline": {"value": v, "unit": UNIT, "cores": info["threads"], "kind": "reference", "sample": sample},
               "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
               "gpu_launches": 0, "wall_s": wall}
        print(json.dumps(out))
        return 0

    # ---- our arm ------------------------------------------------------------------------------------------------
    from mantaflow_b200._lib import lib, check
    from mantaflow_b200 import manta as M
    L = lib()
    dist = None
    comm = None
    if world > 1:
        # torch.distributed is plumbing only (host barrier + max over ranks); the data path is the library's own
        # peer-mapped address space and device barrier
        import torch.distributed as dist
        dist.init_process_group("gloo")
        comm = M.Comm.from_env()

    def barrier():
        if dist is not None:
            dist.barrier()

    def max_over_ranks(x):
        if dist is None:
            return x
        import torch
        t = torch.tensor([x], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def gather(obj):
        if dist is None:
            return [obj]
        lst = [None] * world
        dist.all_gather_object(lst, obj)
        return lst

    # ---- real-rank parity before anything is timed (N > 1) ------------------------------------------------------------
    parity = None
    if world > 1 and not a.no_parity:
        try:
            parity = parity_leg(M, comm, world, rank, local, size, a.precond, gather)
        except Exception as e:      # noqa: BLE001  reported in the line; the ranks stay in lock-step (every rank raises alike)
            parity = {"ok": False, "error": "%s: %s" % (type(e).__name__, str(e)[:300])}
        barrier()

    P = Plume(size, local, a.precond, comm)
    s = P.s
    cells_gpu = cells // world
    k0, k1 = s.slab()
    size_gpu = (size[0], size[1], k1 - k0)
    for _ in range(max(a.warmup, 3)):
        P.step()
    s.sync()

    # ---- timed region: K steps, CUDA events on the solver stream, barrier + sync on both sides ---------------------
    clocks = ClockSampler(local)
    barrier(); s.sync()
    clocks.start()
    l0 = s.launchCount()
    b0 = L.mf_solver_barrier_count(s._h) if world > 1 else 0
    P.iterations = []
    dev_s, wall = time_steps(P, L, check, a.steps, barrier, max_over_ranks)
    clk = clocks.stop()
    launches = s.launchCount() - l0
    barriers_per_step = ((L.mf_solver_barrier_count(s._h) - b0) / float(a.steps)) if world > 1 else 0
    value = cells * a.steps / dev_s
    iters = list(P.iterations)
    ms = C.c_float()

    # ---- per-kernel CUDA-event profile (separate pass so the event pairs do not perturb `value`) -----------------
    kernels, roofline = [], None
    if a.profile_steps > 0:
        check(L.mf_solver_profile_begin(s._h))
        check(L.mf_solver_event_record(s._h, 2))
        for _ in range(a.profile_steps):
            P.step()
        check(L.mf_solver_event_record(s._h, 3))
        buf = C.create_string_buffer(1 << 16)
        check(L.mf_solver_profile_end(s._h, buf, len(buf)))
        check(L.mf_solver_event_elapsed_ms(s._h, 2, 3, C.byref(ms)))
        prof_ms = ms.value
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback (B200_PROFILING.md 6.65 TB/s)"
        entry_bytes = {}
        if a.precond == "mg":
            nl, fb, fmt = C.c_int(0), C.c_int(0), C.c_int(0)
            if L.mf_mg_info(s._h, C.byref(nl), C.byref(fb)) == 0:
                for lv in range(1, nl.value):
                    if L.mf_mg_level_format(s._h, lv, C.byref(fmt)) == 0:
                        entry_bytes[lv] = fmt.value
        for line in buf.value.decode().splitlines():
            name, n, tot = line.split()
            n, tot = int(n), float(tot)
            e = {"kernel": name, "launches_per_step": n / a.profile_steps, "ms_per_step": tot / a.profile_steps,
                 "share_of_step": tot / prof_ms, "avg_us": 1e3 * tot / n}
            nbytes, bound = algo_bytes(name, size_gpu, cells_gpu, entry_bytes)
            if nbytes is not None:
                e["algo_bytes_per_launch"] = nbytes
                e["achieved_gbs"] = nbytes / (tot / n * 1e-3) / 1e9
                e["frac_of_peak"] = e["achieved_gbs"] / peak
                e["bound"] = bound
                if name.partition("@L")[0] in ALGO_BYTES_PER_CELL:
                    e["algo_bytes_per_cell"] = ALGO_BYTES_PER_CELL[name.partition("@L")[0]]
            kernels.append(e)
        if a.kernels_out and rank == 0:
            with open(a.kernels_out, "w") as f:
                json.dump({"ms_per_step_profiled": prof_ms / a.profile_steps, "kernels": kernels}, f, indent=1)
        top = kernels[0] if kernels else None      # the list is sorted by total time: the kernel with the largest share of the step
        if top:
            traffic = None
            try:   # dram bytes per launch of the dominant kernel from the committed ncu capture, if any
                tr = json.load(open(os.path.join(ROOT, "profiles", "dram_traffic.json")))
                t256 = tr.get(str(a.size), {})
                traffic = t256.get(top["kernel"], t256.get(top["kernel"].split("@")[0]))
            except Exception:
                pass
            roofline = {"bound": top.get("bound", "latency"), "kernel": top["kernel"], "achieved": top.get("achieved_gbs"), "peak": peak, "unit": "GB/s",
                        "frac": top.get("frac_of_peak"), "traffic": traffic, "peak_source": peak_src,
                        "algo_bytes_per_launch": top.get("algo_bytes_per_launch"), "avg_launch_us": top["avg_us"],
                        "operator_entry_bytes": entry_bytes.get(int(top["kernel"].partition("@L")[2])) if "@L" in top["kernel"] and top["kernel"].partition("@L")[2] != "0" else None,
                        "launches_per_step": top["launches_per_step"],
                        "share_of_step": top["share_of_step"], "timed": "CUDA events around every launch, %d extra steps after the timed region" % a.profile_steps,
                        "note": "the kernel with the largest share of the step; bound 'latency' = the level's operator and vectors are L2-resident, "
                                "the figure is then L2-level bytes over time, not an HBM fraction to be maximised"}
            # the largest HBM-bound kernel is named beside it when the top one is latency-bound
            hb = next((k for k in kernels if k.get("bound") == "hbm"), None)
            if hb is not None and hb is not top:
                roofline["top_hbm_kernel"] = {"kernel": hb["kernel"], "achieved": hb["achieved_gbs"], "frac": hb["frac_of_peak"],
                                              "share_of_step": hb["share_of_step"], "avg_launch_us": hb["avg_us"],
                                              "algo_bytes_per_launch": hb["algo_bytes_per_launch"]}

    # ---- PCG iteration time: events around the iteration loop alone (MF_TIME_PCG -> event slots 14 / 15), 3 extra steps ------
    pcg_ms = None
    if a.profile_steps > 0:
        os.environ["MF_TIME_PCG"] = "1"
        loop_ms, solve_ms = [], []
        P.solve_ms = []
        n0 = len(P.iterations)
        for _ in range(3):
            P.step()
            check(L.mf_solver_event_elapsed_ms(s._h, 14, 15, C.byref(ms)))
            loop_ms.append(ms.value)
        del os.environ["MF_TIME_PCG"]
        its = sum(P.iterations[n0:])
        pcg_ms = {"ms_per_iteration": sum(loop_ms) / max(its, 1), "loop_ms_per_step": sum(loop_ms) / 3.0, "solve_ms_per_step": sum(P.solve_ms) / 3.0,
                  "iterations_per_step": its / 3.0,
                  "what": "CUDA events around the PCG iteration loop alone (SpMV + update + norm, V-cycle, dot, search update per iteration; "
                          "the converged iteration stops after its norm), divided by the iterations; solve_ms = all of mf_solve_pressure"}
        P.solve_ms = None

    # ---- e2e: host buffers, copies inside the timed region -------------------------------------------------------
    e2e = None
    if not a.no_e2e:
        import numpy as np
        slab_cells = size[0] * size[1] * (k1 - k0)
        nb_vel, nb_real = 12 * slab_cells, 4 * slab_cells

        def pinned(nbytes, dtype, shape):
            p = L.mf_host_alloc(nbytes)
            if not p:
                raise RuntimeError("mf_host_alloc failed")
            return p, np.ctypeslib.as_array((C.c_byte * nbytes).from_address(p)).view(dtype).reshape(shape)
        shp = (k1 - k0, size[1], size[0])
        # two sets of pinned input and output buffers per rank (the slab of the rank; the whole grid at N=1)
        inb = [(pinned(nb_vel, np.float32, shp + (3,)), pinned(nb_real, np.float32, shp)) for _ in range(2)]
        outb = [(pinned(nb_vel, np.float32, shp + (3,)), pinned(nb_real, np.float32, shp), pinned(nb_real, np.float32, shp)) for _ in range(2)]
        for (pv, _), (pd, _) in inb:
            check(L.mf_grid_download_slab(P.vel._h, pv)); check(L.mf_grid_download_slab(P.density._h, pd))
        ke = max(2, min(a.steps, 10))
        # (1) serial: upload, step, download, one after the other on the solver stream
        ks = max(2, min(a.steps, 3))
        barrier(); s.sync()
        t0 = time.perf_counter()
        for i in range(ks):
            (pv, _), (pd, _) = inb[i & 1]
            check(L.mf_grid_upload_slab(P.vel._h, pv)); check(L.mf_grid_upload_slab(P.density._h, pd))
            P.step()
            (ov, _), (od, _), (op, _) = outb[i & 1]
            check(L.mf_grid_download_slab(P.vel._h, ov)); check(L.mf_grid_download_slab(P.density._h, od)); check(L.mf_grid_download_slab(P.pressure._h, op))
        s.sync()
        ts = max_over_ranks(time.perf_counter() - t0)
        # (2) streaming API: the upload of step i+1 and the download of step i-1 overlap the kernels of step i
        def stream_steps(n):
            P.vel.stageUpload(inb[0][0][0]); P.density.stageUpload(inb[0][1][0])
            for i in range(n):
                P.vel.commitUpload(); P.density.commitUpload()
                if i + 1 < n:
                    P.vel.stageUpload(inb[(i + 1) & 1][0][0]); P.density.stageUpload(inb[(i + 1) & 1][1][0])
                P.step()
                (ov, _), (od, _), (op, _) = outb[i & 1]
                P.vel.downloadAsync(ov); P.density.downloadAsync(od); P.pressure.downloadAsync(op)
            s.transfersSync()
        stream_steps(2)                      # untimed: allocates the staging slots
        hp = outb[(ke - 1) & 1][2][1]
        barrier(); s.sync()
        t0 = time.perf_counter()
        stream_steps(ke)
        te = max_over_ranks(time.perf_counter() - t0)
        barrier()
        e2e = {"value": cells * ke / te, "unit": UNIT, "h2d_bytes_per_step": (nb_vel + nb_real) * world, "d2h_bytes_per_step": (nb_vel + 2 * nb_real) * world,
               "steps": ke, "ms_per_step": 1e3 * te / ke, "checksum": float(hp[(k1 - k0) // 2, size[1] // 2, size[0] // 2]),
               "path": "per rank and step: mf_grid_stage_upload + mf_grid_commit_upload (vel, density) from pinned host -> plume step through "
                       "mantaflow_b200.manta -> mf_grid_download_async (vel, density, pressure) to pinned host; copies run on copy streams beside "
                       "the solver stream (double-buffered staging); all %d uploads and downloads lie inside the timed region" % ke,
               "serial": {"value": cells * ks / ts, "ms_per_step": 1e3 * ts / ks, "steps": ks,
                          "path": "mf_grid_upload_slab -> step -> mf_grid_download_slab, no overlap"}}
        for bufs in inb + outb:
            for p, _ in bufs:
                L.mf_host_free(p)

    # ---- BASELINE configs[4]: 1024^3, strong scaling over the N GPUs of this run -----------------------------------------
    del P
    s = None
    gc_collect()
    strong_1024 = None
    if not a.no_strong1024 and a.size == 256 and not a.strong:
        try:
            strong_1024 = strong_leg(M, L, check, comm, world, rank, local, a.precond, barrier, max_over_ranks)
        except Exception as e:      # noqa: BLE001
            strong_1024 = {"error": "%s: %s" % (type(e).__name__, str(e)[:300])}

    # ---- CPU baseline on this box's host cores (rank 0, N=1 only): the reference binary on a bounded sample -----------
    cpu = None
    if rank == 0 and world == 1 and not a.no_cpu_baseline:
        try:
            pc = 3 if a.precond == "mg" else 1
            v, info = run_reference_binary(size if a.size <= 256 else (256, 256, 256), 2, 1, pc)
            cpu = {"value": v, "unit": UNIT, "cores": info["threads"], "kind": "reference",
                   "sample": "2 timed steps (+1 warm-up) of the same %d^3 plume step, unmodified reference oracle/_ref/manta (OpenMP -O3 float) on %s; step_s=%s"
                             % (min(a.size, 256), cpu_model(), ["%.2f" % t for t in info["step_s"]])}
        except Exception as e:   # the baseline is reported, never required for the GPU numbers
            cpu = {"value": None, "unit": UNIT, "cores": os.cpu_count(), "kind": "reference", "sample": "unavailable: %s" % str(e)[:200]}

    if rank == 0:
        nit = sum(iters) / float(max(len(iters), 1))
        out = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": max(a.warmup, 3),
               "ms_per_step": 1e3 * dev_s / a.steps, "higher_is_better": True, "scaling": "strong" if (a.strong and world > 1) else "weak", "vs_baseline": None,
               "dtype": "f32", "data": "synthetic", "config": config, "clocks": clk, "e2e": e2e, "gpu_launches": int(launches),
               "roofline": roofline, "cpu_baseline": cpu, "pcg_iterations": iters,
               "ms_per_step_per_pcg_iteration": 1e3 * dev_s / a.steps / max(nit, 1e-9),
               "pcg_ms_per_iteration": pcg_ms, "strong_1024": strong_1024, "parity": parity,
               "wall_ms_per_step": 1e3 * wall / a.steps, "kernels": kernels[:16],
               "device_barriers_per_step": barriers_per_step}
        print(json.dumps(out))
    if comm is not None:
        barrier()
        comm.destroy()
    if dist is not None:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
