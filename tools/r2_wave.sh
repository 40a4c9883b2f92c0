#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_kernels_gpu.py tests/test_parity_gpu.py tests/test_multi_gpu.py -q -m gpu -p no:cacheprovider -x > gpurun_out/wave_tests.log 2>&1; tail -3 gpurun_out/wave_tests.log
python tools/vcycle_time.py --size 256 --reps 50 --levels 2>&1 | head -3
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e --no-strong1024 > gpurun_out/bench_wave.json 2> gpurun_out/bench_wave.err; tail -c 300 gpurun_out/bench_wave.err
python - <<PY
import json
d = json.load(open("gpurun_out/bench_wave.json"))
ks = {k["kernel"]: k for k in d["kernels"]}
print("step %.3f ms" % d["ms_per_step"], {n: round(ks[n]["avg_us"], 1) for n in ks if "@L0" in n}, d["pcg_iterations"][:5])
PY
