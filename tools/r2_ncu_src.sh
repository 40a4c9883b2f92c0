#!/bin/bash
# ncu --set full with source import for the level-0 V-cycle legs and the MAC advection kernels (per-SASS-instruction counts)
mkdir -p gpurun_out
python tools/profile_step.py --size 256 --warmup 6 > gpurun_out/r2c_profile_step_plain.log 2>&1 || { tail -5 gpurun_out/r2c_profile_step_plain.log; exit 1; }
ncu --set full --clock-control none --cache-control none --import-source on --profile-from-start off -k regex:'k_mg_down0|k_mg_up0w|k_mc_mac_f|k_sl_mac_f' -c 4 \
  -o gpurun_out/r2c_src -f python tools/profile_step.py --size 256 --warmup 6 > gpurun_out/r2c_ncu_src.log 2>&1; echo "ncu rc=$?"
ls -la gpurun_out/r2c_src.ncu-rep
