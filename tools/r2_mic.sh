#!/bin/bash
# MIC(0) wavefront: parity cases that use PcMIC + the unmodified scenes + PcMIC bench lines (coded fast path vs generic)
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -q -m gpu -p no:cacheprovider -k "mic or simpleplume or scene or plume3d or flip01 or gridcg or solve" > gpurun_out/mic_tests.log 2>&1; tail -8 gpurun_out/mic_tests.log
MF_MIC_TZ=16 MF_MIC_GENERIC=1 timeout 600 compute-sanitizer --tool memcheck python -m pytest tests/test_parity_gpu.py -q -m gpu -p no:cacheprovider -k "solve4_mic_tight or plume3d_mic" > gpurun_out/mic_sanitizer.log 2>&1; grep -E "Invalid|ERROR SUMMARY|passed|failed|at 0x|by thread" gpurun_out/mic_sanitizer.log | head -12
for v in code generic; do
if [ $v = generic ]; then export MF_MIC_GENERIC=1; fi
timeout 600 python bench.py --precond mic --steps 5 --warmup 3 --no-cpu-baseline --no-e2e --no-strong1024 > gpurun_out/bench_mic_$v.json 2> gpurun_out/bench_mic.err; tail -c 300 gpurun_out/bench_mic.err
python - <<PY
import json
d = json.load(open("gpurun_out/bench_mic_$v.json"))
print("$v", d["ms_per_step"], d["pcg_iterations"][:6], d["pcg_ms_per_iteration"]["ms_per_iteration"])
for k in d["kernels"][:8]:
    print("%-24s %6.1f %7.3f %7.1f" % (k["kernel"], k["launches_per_step"], k["ms_per_step"], k["avg_us"]))
PY
done
