"""One plume step inside a cudaProfilerStart/Stop window, for ncu (never a bench number):

  ncu --profile-from-start off --cache-control none --clock-control none --metrics gpu__time_duration.sum,... \
      --csv --log-file gpurun_out/step.csv python tools/profile_step.py --size 256 --warmup 6

The window holds exactly one step (all kernels of advect + forces + pressure solve) after --warmup untimed steps.
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--size", type=int, default=256)
    ap.add_argument("--warmup", type=int, default=6)
    ap.add_argument("--steps", type=int, default=1)
    ap.add_argument("--precond", default="mg")
    a = ap.parse_args()
    import bench
    P = bench.Plume((a.size,) * 3, 0, a.precond)
    for _ in range(a.warmup):
        P.step()
    P.s.sync()
    import torch
    rt = torch.cuda.cudart()
    rt.cudaProfilerStart()
    for _ in range(a.steps):
        P.step()
    P.s.sync()
    rt.cudaProfilerStop()
    print("iterations", P.iterations[-a.steps:])


if __name__ == "__main__":
    main()
