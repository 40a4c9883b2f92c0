#!/bin/bash
# binary16 level 1 with per-colour float-row lists: parity tests, V-cycle level times, bench with and without
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py -q -m gpu -p no:cacheprovider -x -k "vcycle or pressure_solve or gridcg or half" > gpurun_out/half_tests.log 2>&1; tail -4 gpurun_out/half_tests.log
python tools/vcycle_time.py --size 256 --reps 50 --levels 2>&1 | head -3
for v in new old; do
if [ $v = old ]; then export MF_MG_HALF_EXACT_ONLY=1; fi
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e --no-strong1024 > gpurun_out/bench_half_$v.json 2> gpurun_out/bench_half.err; tail -c 300 gpurun_out/bench_half.err
python - <<PY
import json
d = json.load(open("gpurun_out/bench_half_$v.json"))
ks = {k["kernel"]: k for k in d["kernels"]}
print("$v step %.3f ms" % d["ms_per_step"], {n: round(ks[n]["avg_us"], 1) for n in ks if "@L1" in n}, d["pcg_ms_per_iteration"]["ms_per_iteration"], d["pcg_iterations"][:5], d["gpu_launches"])
PY
done
