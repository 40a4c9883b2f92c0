#!/bin/bash
# short weak + strong (1024^3) bench lines on N GPUs, no end-to-end leg (fits a small GPU-minute budget): tools/scale_short.sh <N>
N=${1:-4}; TAG=r1l
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 80 $TR --nproc-per-node $N --master-port 29543 bench.py --gpus $N --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/${TAG}_bench_weak_${N}gpu.json 2> gpurun_out/${TAG}_weak${N}.log; echo "weak rc=$?"; head -c 300 gpurun_out/${TAG}_bench_weak_${N}gpu.json; echo
timeout 80 $TR --nproc-per-node $N --master-port 29544 bench.py --gpus $N --steps 3 --warmup 3 --size 1024 --strong --no-e2e --no-cpu-baseline > gpurun_out/${TAG}_bench_strong1024_${N}gpu.json 2> gpurun_out/${TAG}_strong${N}.log; echo "strong rc=$?"; head -c 300 gpurun_out/${TAG}_bench_strong1024_${N}gpu.json; echo
