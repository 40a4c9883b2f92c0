#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py -q -m gpu -p no:cacheprovider -k "tma" > gpurun_out/tma_tests.log 2>&1; tail -8 gpurun_out/tma_tests.log
for cfg in "0 32" "1 32" "1 64" "1 16"; do
set -- $cfg
MF_ADVECT_TMA=$1 MF_TMA_ZC=$2 timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e --no-strong1024 > gpurun_out/bench_tma.json 2> gpurun_out/bench_tma.err; tail -c 300 gpurun_out/bench_tma.err
python - <<PY
import json
d = json.load(open("gpurun_out/bench_tma.json"))
ks = {k["kernel"]: k for k in d["kernels"]}
print("TMA=$1 ZC=$2 step %.3f ms" % d["ms_per_step"], {n: round(ks[n]["avg_us"], 1) for n in ks if "mac" in n or n.startswith("k_sl") or n.startswith("k_mc")})
PY
done
