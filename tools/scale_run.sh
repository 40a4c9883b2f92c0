#!/bin/bash
# multi-GPU measurements on one box: parity cases at 4 and 8 ranks, weak-scaling plume at N ranks, strong scaling at 1024^3
# usage: tools/scale_run.sh <ngpus> <tag>      (writes gpurun_out/<tag>_*.json|log)
N=${1:-8}; TAG=${2:-scale}
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
mkdir -p gpurun_out
if [ "$N" -ge 4 ]; then
timeout 300 $TR --nproc-per-node 4 --master-port 29541 tools/dist_cases.py --out gpurun_out/${TAG}_cases4.json \
  adv_real_o2_c2 adv_mac_o2_c2 solve_mgstatic_tight solve_mic_tight plume3d_mg > gpurun_out/${TAG}_cases4.log 2>&1; echo "cases4 rc=$?"
fi
if [ "$N" -ge 8 ]; then
timeout 300 $TR --nproc-per-node 8 --master-port 29542 tools/dist_cases.py --out gpurun_out/${TAG}_cases8.json \
  plume3d_mg > gpurun_out/${TAG}_cases8.log 2>&1; echo "cases8 rc=$?"
fi
timeout 600 $TR --nproc-per-node $N --master-port 29543 bench.py --gpus $N --steps 10 --warmup 5 --kernels-out gpurun_out/${TAG}_weak${N}_kernels.json \
  > gpurun_out/${TAG}_weak${N}.json 2> gpurun_out/${TAG}_weak${N}.log; echo "weak rc=$?"; tail -c 1500 gpurun_out/${TAG}_weak${N}.json
timeout 900 $TR --nproc-per-node $N --master-port 29544 bench.py --gpus $N --steps 5 --warmup 3 --size 1024 --strong --no-e2e \
  > gpurun_out/${TAG}_strong1024_${N}.json 2> gpurun_out/${TAG}_strong1024_${N}.log; echo "strong rc=$?"; tail -c 1500 gpurun_out/${TAG}_strong1024_${N}.json
tail -5 gpurun_out/${TAG}_strong1024_${N}.log
