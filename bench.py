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
`--impl reference` times the reference's own CPU implementation of the same step on the host cores.
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
    "k_add_buoyancy": 32.0,    # flags 4 + density 4 + vel 12 r + 12 w
    "k_wall_buoyancy": 32.0,
    "k_make_rhs": 20.0,        # flags 4 + vel 12 + rhs 4
    "k_make_rhs_plain": 20.0,
    "k_flag_stats": 4.0,
    "k_make_laplace_code": 6.0,  # flags 4 + row code 2
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


def main():
    a = _parse()
    if not a.code_cache:
        os.environ["MF_CG_NOCODECACHE"] = "1"     # every step builds its matrix, as the reference's solvePressureSystem does
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    strong = a.strong or world == 1
    size = (a.size, a.size, a.size if strong else a.size * world)
    cells = size[0] * size[1] * size[2]            # whole job
    pcname = "GridMg-preconditioned PCG (PcMGStatic)" if a.precond == "mg" else "red-black MIC PCG (PcMIC)"
    config = {"workload": "smoke plume %dx%dx%d, MacCormack (order 2) advection of density+velocity, setWallBcs, addBuoyancy, "
                          "solvePressure %s, cgAccuracy 1e-3, zeroPressureFixing" % (size + (pcname,)),
              "grid": list(size), "cells_per_gpu": cells // world, "dt": 1.0,
              "matrix": "row codes reused while the flags are unchanged" if a.code_cache else "rebuilt every step (MF_CG_NOCODECACHE=1)",
              "l2": "working set per step > 1 GiB (>> 126 MB L2), no flush needed" if a.size >= 256 else "L2-resident working set",
              "parallelism": ("z-slab x%d (one global address space over NVLink, device flag barrier per launch group), %s scaling"
                              % (world, "strong" if a.strong else "weak")) if world > 1 else "single GPU"}

    if a.impl == "reference":
        if rank != 0:
            return 0
        pc = 3 if a.precond == "mg" else 1
        ref_size = (min(a.size, 256),) * 3      # bounded sample: one GPU's share of the (weak-scaled) workload
        # keep the whole run within a few minutes: a 256^3 reference step costs several seconds on the host cores
        if a.size > 128 and (a.steps + a.warmup) > 16:
            ref_size = (128, 128, 128)
        t0 = time.time()
        v, info = run_reference_binary(ref_size, a.steps, a.warmup, pc)
        wall = time.time() - t0
        sample = "%d^3 plume step x%d (+%d warm-up), reference binary oracle/_ref/manta (OpenMP, -O3, float), %d threads on %s" % (
            ref_size[0], a.steps, a.warmup, info["threads"], cpu_model())
        out = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup,
               "ms_per_step": 1e3 * sum(info["step_s"]) / len(info["step_s"]), "higher_is_better": True, "scaling": "weak",
               "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": dict(config, sample_grid=list(ref_size)),
               "cpu_baseline": {"value": v, "unit": UNIT, "cores": info["threads"], "kind": "reference", "sample": sample},
               "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
               "gpu_launches": 0, "wall_s": wall}
        print(json.dumps(out))
        return 0

    # ---- our arm ------------------------------------------------------------------------------------------------
    from mantaflow_b200._lib import lib, check
    L = lib()
    dist = None
    comm = None
    if world > 1:
        # torch.distributed is plumbing only (host barrier + max over ranks); the data path is the library's own
        # peer-mapped address space and device barrier
        import torch.distributed as dist
        dist.init_process_group("gloo")
        from mantaflow_b200 import manta as M
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

    P = Plume(size, local, a.precond, comm)
    s = P.s
    cells_gpu = cells // world
    for _ in range(max(a.warmup, 3)):
        P.step()
    s.sync()

    # ---- timed region: K steps, CUDA events on the solver stream, barrier + sync on both sides ---------------------
    clocks = ClockSampler(local)
    barrier(); s.sync()
    clocks.start()
    l0 = s.launchCount()
    P.iterations = []
    check(L.mf_solver_event_record(s._h, 0))
    t0 = time.perf_counter()
    for _ in range(a.steps):
        P.step()
    check(L.mf_solver_event_record(s._h, 1))
    s.sync()
    wall = time.perf_counter() - t0
    barrier()
    ms = C.c_float()
    check(L.mf_solver_event_elapsed_ms(s._h, 0, 1, C.byref(ms)))
    clk = clocks.stop()
    launches = s.launchCount() - l0
    dev_s = max_over_ranks(ms.value * 1e-3)
    value = cells * a.steps / dev_s
    iters = list(P.iterations)

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
        for line in buf.value.decode().splitlines():
            name, n, tot = line.split()
            n, tot = int(n), float(tot)
            e = {"kernel": name, "launches_per_step": n / a.profile_steps, "ms_per_step": tot / a.profile_steps,
                 "share_of_step": tot / prof_ms, "avg_us": 1e3 * tot / n}
            base = name[:-3] if name.endswith("@L0") else name     # MF_PROFILE_LEVELS=1 tags multigrid kernels with their level
            if base in ALGO_BYTES_PER_CELL:
                e["algo_bytes_per_cell"] = ALGO_BYTES_PER_CELL[base]
                e["achieved_gbs"] = ALGO_BYTES_PER_CELL[base] * cells_gpu / (tot / n * 1e-3) / 1e9
                e["frac_of_peak"] = e["achieved_gbs"] / peak
            kernels.append(e)
        if a.kernels_out and rank == 0:
            with open(a.kernels_out, "w") as f:
                json.dump({"ms_per_step_profiled": prof_ms / a.profile_steps, "kernels": kernels}, f, indent=1)
        top = next((k for k in kernels if "achieved_gbs" in k), None)
        if top:
            traffic = None
            try:   # dram bytes per launch of the dominant kernel from the committed ncu capture, if any
                tr = json.load(open(os.path.join(ROOT, "profiles", "dram_traffic.json")))
                traffic = tr.get(str(a.size), {}).get(top["kernel"].split("@")[0])
            except Exception:
                pass
            roofline = {"bound": "hbm", "kernel": top["kernel"], "achieved": top["achieved_gbs"], "peak": peak, "unit": "GB/s",
                        "frac": top["achieved_gbs"] / peak, "traffic": traffic, "peak_source": peak_src,
                        "algo_bytes_per_launch": top["algo_bytes_per_cell"] * cells_gpu, "avg_launch_us": top["avg_us"],
                        "share_of_step": top["share_of_step"], "timed": "CUDA events around every launch, %d extra steps after the timed region" % a.profile_steps}

    # ---- PCG iteration time: whole projection (RHS, matrix, solve, correction) / iterations, 3 extra steps ------------
    pcg_ms = None
    if a.profile_steps > 0:
        P.solve_ms = []
        n0 = len(P.iterations)
        for _ in range(3):
            P.step()
        its = sum(P.iterations[n0:])
        pcg_ms = {"ms_per_iteration": sum(P.solve_ms) / max(its, 1), "solve_ms_per_step": sum(P.solve_ms) / 3.0, "iterations_per_step": its / 3.0,
                  "what": "mf_solve_pressure (MakeRhs + matrix + PCG incl. one V-cycle per iteration + correctVelocity) divided by the PCG iterations"}
        P.solve_ms = None

    # ---- e2e: host buffers, copies inside the timed region -------------------------------------------------------
    e2e = None
    if not a.no_e2e:
        import numpy as np
        k0, k1 = s.slab()
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
        out = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": max(a.warmup, 3),
               "ms_per_step": 1e3 * dev_s / a.steps, "higher_is_better": True, "scaling": "strong" if (a.strong and world > 1) else "weak", "vs_baseline": None,
               "dtype": "f32", "data": "synthetic", "config": config, "clocks": clk, "e2e": e2e, "gpu_launches": int(launches),
               "roofline": roofline, "cpu_baseline": cpu, "pcg_iterations": iters,
               "pcg_ms_per_iteration": pcg_ms, "wall_ms_per_step": 1e3 * wall / a.steps, "kernels": kernels[:14],
               "device_barriers_per_step": (L.mf_solver_barrier_count(s._h) / float(max(a.warmup, 3) + a.steps + a.profile_steps + (3 if a.profile_steps > 0 else 0) + (0 if a.no_e2e else max(2, min(a.steps, 10)) + 2 + max(2, min(a.steps, 3))))) if world > 1 else 0}
        print(json.dumps(out))
    del P
    s = None
    import gc
    gc.collect()
    if comm is not None:
        barrier()
        comm.destroy()
    if dist is not None:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
