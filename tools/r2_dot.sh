#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_kernels_gpu.py tests/test_parity_gpu.py tests/test_multi_gpu.py -q -m gpu -p no:cacheprovider -x > gpurun_out/dot_tests.log 2>&1; tail -4 gpurun_out/dot_tests.log
for v in new old; do
if [ $v = old ]; then export MF_MG_NODOT=1; fi
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e --no-strong1024 > gpurun_out/bench_dot_$v.json 2> gpurun_out/bench_dot.err; tail -c 300 gpurun_out/bench_dot.err
python - <<PY
import json
d = json.load(open("gpurun_out/bench_dot_$v.json"))
ks = {k["kernel"]: k for k in d["kernels"]}
print("$v step %.3f ms" % d["ms_per_step"], {n: round(ks[n]["avg_us"], 1) for n in ks if "up0w" in n or "dot_beta" in n}, d["pcg_ms_per_iteration"]["ms_per_iteration"], d["pcg_iterations"][:5], d["gpu_launches"])
PY
done
python - <<'PY'
import json
r = json.load(open("gpurun_out/parity_report.json"))
bad = [k for k, v in r.items() if any(isinstance(x, dict) and x.get("mismatching") for x in v.values())]
print("cases", len(r), "not bit-identical:", bad)
PY
