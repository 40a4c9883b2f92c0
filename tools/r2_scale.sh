#!/bin/bash
# bench.py under torchrun on N GPUs, as the driver launches it (N = $1); extra bench flags in $2...
N=${1:-2}; shift
mkdir -p gpurun_out
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $N --steps 10 --warmup 3 --kernels-out gpurun_out/kernels_n$N.json "$@" > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo rc=$?; tail -c 1500 gpurun_out/bench_n$N.err | grep -v "OMP_NUM\|\*\*\*\*\|^$"
python tools/show_bench.py gpurun_out/bench_n$N.json | grep -v "^roofline\|^pcg\|^cpu"
