#!/bin/bash
# usage: tools/r2_ab.sh ENVVAR [pytest -k expression]   -- parity suites, then the bench line with ENVVAR unset ("new") and set to 1 ("old")
V=$1; K=${2:-}
mkdir -p gpurun_out
if [ -n "$K" ]; then timeout 1500 python -m pytest tests/test_kernels_gpu.py tests/test_parity_gpu.py tests/test_multi_gpu.py -q -m gpu -p no:cacheprovider -x -k "$K" > gpurun_out/ab_tests.log 2>&1
else timeout 1500 python -m pytest tests/test_kernels_gpu.py tests/test_parity_gpu.py tests/test_multi_gpu.py -q -m gpu -p no:cacheprovider -x > gpurun_out/ab_tests.log 2>&1; fi
tail -4 gpurun_out/ab_tests.log
for v in new old; do
if [ $v = old ]; then export $V=1; fi
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e --no-strong1024 > gpurun_out/bench_ab_$v.json 2> gpurun_out/bench_ab.err; tail -c 300 gpurun_out/bench_ab.err
python - <<PY
import json
d = json.load(open("gpurun_out/bench_ab_$v.json"))
ks = {k["kernel"]: k for k in d["kernels"]}
print({n: round(ks[n]["avg_us"], 1) for n in ks if n in ("k_mc_mac_f", "k_mc_centered_f", "k_sl_mac_f", "k_sl_f", "k_mg_down0@L0", "k_mg_up0w@L0", "k_mg_smooth@L1", "k_mg_interp_add_b@L1", "k_mg_interp_add_c@L1", "k_mg_residual@L1")})
print("$v step %.3f ms" % d["ms_per_step"], "pcg/it %.4f" % d["pcg_ms_per_iteration"]["ms_per_iteration"], "solve %.3f" % d["pcg_ms_per_iteration"]["solve_ms_per_step"], d["pcg_iterations"][:5], d["gpu_launches"], d["roofline"]["kernel"], round(d["roofline"]["frac"] or 0, 3), d["roofline"]["bound"])
PY
done
