#!/bin/bash
bash tools/r2_full.sh nobench
timeout 900 python tools/double_oracle.py > gpurun_out/double_oracle.json 2> gpurun_out/double_oracle.err; tail -3 gpurun_out/double_oracle.err
