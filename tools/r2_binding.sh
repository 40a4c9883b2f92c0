#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_binding_gpu.py tests/test_binding_cpu.py -q -m gpu -p no:cacheprovider > gpurun_out/binding_tests.log 2>&1; tail -30 gpurun_out/binding_tests.log
python - <<'PY'
import json
r = json.load(open("gpurun_out/parity_report.json"))
for k, v in r.items():
    if k.startswith("binding"):
        print(k, v)
PY
