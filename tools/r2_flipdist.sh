#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_multi_gpu.py -q -m gpu -p no:cacheprovider -k "flip or two_processes" > gpurun_out/flipdist_tests.log 2>&1; tail -6 gpurun_out/flipdist_tests.log
