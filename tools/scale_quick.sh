#!/bin/bash
# 8-GPU confirmation of a round's kernels: one parity case on N real ranks (the only committed case with enough z planes for 8), the weak-scaling bench line, strong scaling at 1024^3
# usage: tools/scale_quick.sh <ngpus> <tag>
N=${1:-8}; TAG=${2:-scale}
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
mkdir -p gpurun_out
timeout 300 $TR --nproc-per-node $N --master-port 29542 tools/dist_cases.py --out gpurun_out/${TAG}_dist_cases_${N}gpu.json \
  plume3d_mg > gpurun_out/${TAG}_cases${N}.log 2>&1; echo "cases rc=$?"; tail -3 gpurun_out/${TAG}_cases${N}.log
timeout 600 $TR --nproc-per-node $N --master-port 29543 bench.py --gpus $N --steps 10 --warmup 5 --no-cpu-baseline \
  > gpurun_out/${TAG}_bench_weak_${N}gpu.json 2> gpurun_out/${TAG}_weak${N}.log; echo "weak rc=$?"; head -c 600 gpurun_out/${TAG}_bench_weak_${N}gpu.json; echo
timeout 900 $TR --nproc-per-node $N --master-port 29544 bench.py --gpus $N --steps 5 --warmup 3 --size 1024 --strong --no-e2e --no-cpu-baseline \
  > gpurun_out/${TAG}_bench_strong1024_${N}gpu.json 2> gpurun_out/${TAG}_strong${N}.log; echo "strong rc=$?"; head -c 600 gpurun_out/${TAG}_bench_strong1024_${N}gpu.json; echo
tail -3 gpurun_out/${TAG}_strong${N}.log
