#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_parity_gpu.py tests/test_multi_gpu.py -q -m gpu -p no:cacheprovider -k "spmv or laplace or solve or plume3d or rowcode or gridcg" > gpurun_out/spmv_tests.log 2>&1; tail -5 gpurun_out/spmv_tests.log
for cfg in "1 32" "0 32" "0 16" "0 64"; do
set -- $cfg
if [ $1 = 1 ]; then export MF_SPMV_FLAT=1; else unset MF_SPMV_FLAT; fi
MF_SPMV_ZC=$2 timeout 600 python bench.py --steps 10 --warmup 5 --no-cpu-baseline --no-e2e --no-strong1024 > gpurun_out/bench_spmv.json 2> gpurun_out/bench_spmv.err; tail -c 300 gpurun_out/bench_spmv.err
python - <<PY
import json
d = json.load(open("gpurun_out/bench_spmv.json"))
ks = {k["kernel"]: k for k in d["kernels"]}
print("FLAT=$1 ZC=$2 step %.3f ms" % d["ms_per_step"], {n: round(ks[n]["avg_us"], 1) for n in ks if "spmv" in n}, d["pcg_iterations"][:4])
PY
done
python - <<'PY'
import json
r = json.load(open("gpurun_out/parity_report.json"))
bad = [k for k, v in r.items() if any(isinstance(x, dict) and x.get("mismatching") for x in v.values())]
print("cases", len(r), "not bit-identical:", bad)
PY
