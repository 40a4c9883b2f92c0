#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_multi_gpu.py tests/test_kernels_gpu.py -q -m gpu -p no:cacheprovider -k "emulated or vcycle" > gpurun_out/mgpu_tests.log 2>&1; tail -4 gpurun_out/mgpu_tests.log
bash tools/r2_scale.sh 2 --no-strong1024 --no-e2e | head -12
