#!/bin/bash
# parity + kernel suites, one bench line and the times of the small streaming kernels (development helper)
timeout 900 python -m pytest tests/test_parity_gpu.py tests/test_kernels_gpu.py -x -q -m gpu 2>&1 | tail -4
timeout 300 python bench.py --steps 10 --warmup 5 --no-cpu-baseline --no-e2e --kernels-out gpurun_out/kq.json > gpurun_out/bench_quick.json 2>gpurun_out/bench_quick.err
python - <<'PY'
import json
d = json.load(open("gpurun_out/bench_quick.json")); print(d["ms_per_step"], d["pcg_iterations"])
for k in json.load(open("gpurun_out/kq.json"))["kernels"]:
    if k["kernel"] in ("k_set_wall_bcs", "k_set_wall_bcs_edge<4>", "k_set_wall_bcs_edge", "k_add_buoyancy", "k_cg_dot_sigma", "k_correct_velocity", "k_make_rhs_plain", "k_make_laplace_code", "k_cg_update", "k_cg_search", "k_cg_dot_beta", "k_cg_spmv_dot_code"):
        print("%-22s %7.1f us x %.1f" % (k["kernel"], k["avg_us"], k["launches_per_step"]))
PY
