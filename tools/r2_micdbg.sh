#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests -q -m gpu -p no:cacheprovider -k "mic or simpleplume or scene_simple or plume3d or gridcg" > gpurun_out/mic_tests.log 2>&1; tail -4 gpurun_out/mic_tests.log
cd gpurun_out
for occ in 3 2; do
MF_MIC_OCC=$occ timeout 600 python ../bench.py --precond mic --steps 5 --warmup 3 --no-cpu-baseline --no-e2e --no-strong1024 > bench_mic_code.json 2> bench_mic.err; tail -c 300 bench_mic.err
python - <<PY
import json
d = json.load(open("bench_mic_code.json"))
print("occ $occ", d["ms_per_step"], d["pcg_iterations"][:6], d["pcg_ms_per_iteration"]["ms_per_iteration"])
for k in d["kernels"][:3]:
    print("%-24s %6.1f %7.3f %7.1f" % (k["kernel"], k["launches_per_step"], k["ms_per_step"], k["avg_us"]))
PY
done
MF_MIC_DBG=40 timeout 600 python ../bench.py --precond mic --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-strong1024 --profile-steps 0 > bench_micdbg.json 2> bench_micdbg.err; tail -c 300 bench_micdbg.err
python ../tools/mic_dbg.py mic_dbg.bin | head -34
