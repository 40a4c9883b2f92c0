#!/bin/bash
# round-2 check on one GPU box: the whole GPU suite (all failures listed), smoke(), the default bench line and a short reference arm
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -q -m gpu -p no:cacheprovider > gpurun_out/full_tests.log 2>&1; tail -30 gpurun_out/full_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1; tail -2 gpurun_out/smoke.log
timeout 900 python bench.py --kernels-out gpurun_out/kernels.json > gpurun_out/bench.json 2> gpurun_out/bench.err; tail -c 1500 gpurun_out/bench.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; tail -c 400 gpurun_out/bench_ref.json
python - <<'PY'
import json
d = json.load(open("gpurun_out/bench.json"))
print(d["ms_per_step"], d["pcg_iterations"][:6], d["gpu_launches"], "e2e", d["e2e"]["ms_per_step"] if d.get("e2e") else None)
print("roofline", json.dumps(d["roofline"])[:900])
print("pcg", d["pcg_ms_per_iteration"]); print("strong", d["strong_1024"]); print("cpu", d["cpu_baseline"])
for k in d["kernels"]:
    print("%-24s %6.1f %7.3f %7.1f %s" % (k["kernel"], k["launches_per_step"], k["ms_per_step"], k["avg_us"], ("%.2f" % k["frac_of_peak"]) if "frac_of_peak" in k else ""))
PY
