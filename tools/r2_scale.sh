#!/bin/bash
# bench.py under torchrun on N GPUs, as the driver launches it (N = $1), plus the reference arm at the same N
N=${1:-2}
mkdir -p gpurun_out
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo rc=$?; tail -c 1500 gpurun_out/bench_n$N.err
python tools/show_bench.py gpurun_out/bench_n$N.json
