#!/bin/bash
# MIC(0) wavefront: parity cases that use PcMIC + the unmodified scenes + a PcMIC bench line
timeout 1200 python -m pytest tests -q -m gpu -p no:cacheprovider -k "mic or simpleplume or scene or plume3d or flip01 or gridcg or solve" > gpurun_out/mic_tests.log 2>&1; tail -15 gpurun_out/mic_tests.log
for tz in 8 16; do
MF_MIC_TZ=$tz timeout 600 python bench.py --precond mic --steps 5 --warmup 3 --no-cpu-baseline --no-e2e --no-strong1024 > gpurun_out/bench_mic_tz$tz.json 2> gpurun_out/bench_mic.err; tail -c 300 gpurun_out/bench_mic.err
python - <<PY
import json
d = json.load(open("gpurun_out/bench_mic_tz$tz.json"))
print("TZ", $tz, d["ms_per_step"], d["pcg_iterations"][:6], d["pcg_ms_per_iteration"])
for k in d["kernels"][:8]:
    print("%-24s %6.1f %7.3f %7.1f" % (k["kernel"], k["launches_per_step"], k["ms_per_step"], k["avg_us"]))
PY
done
