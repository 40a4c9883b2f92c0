#!/bin/bash
# everything the round-end check runs, on one box: the whole GPU suite, smoke(), one bench line with the per-kernel table
timeout 1700 python -m pytest tests -x -q -m gpu > gpurun_out/full_tests.log 2>&1; tail -4 gpurun_out/full_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1; tail -2 gpurun_out/smoke.log
timeout 300 python bench.py --steps 10 --warmup 5 --no-cpu-baseline --no-e2e "$@" > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; tail -c 600 gpurun_out/bench_quick.err
python - <<'PY'
import json
d = json.load(open("gpurun_out/bench_quick.json"))
print(d["ms_per_step"], d["pcg_iterations"], d["gpu_launches"])
for k in d["kernels"]:
    print("%-24s %6.1f %7.3f %7.1f" % (k["kernel"], k["launches_per_step"], k["ms_per_step"], k["avg_us"]))
PY
