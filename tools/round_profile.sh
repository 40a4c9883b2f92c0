#!/bin/bash
# end-of-round measurements on ONE B200: bench line, reference arm, ncu launch list of one step, ncu --set full of the top kernels.
# usage: tools/round_profile.sh <tag>     (writes gpurun_out/<tag>_*)
TAG=${1:-rX}
mkdir -p gpurun_out
timeout 900 python bench.py --steps 20 --warmup 5 --kernels-out gpurun_out/${TAG}_kernels_256_mg.json > gpurun_out/${TAG}_bench_256_mg.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${TAG}_bench_reference_arm.json 2>> gpurun_out/${TAG}_bench.err; echo "reference arm rc=$?"
timeout 600 python bench.py --precond mic --steps 5 --warmup 3 --no-cpu-baseline --no-e2e --no-strong1024 --kernels-out gpurun_out/${TAG}_kernels_256_mic.json > gpurun_out/${TAG}_bench_256_mic.json 2>> gpurun_out/${TAG}_bench.err; echo "mic rc=$?"
MF_ADVECT_TMA=1 timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e --no-strong1024 --kernels-out gpurun_out/${TAG}_kernels_256_mg_tma.json > gpurun_out/${TAG}_bench_256_mg_tma.json 2>> gpurun_out/${TAG}_bench.err; echo "tma rc=$?"
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,l1tex__t_sector_hit_rate.pct,sm__warps_active.avg.pct_of_peak_sustained_active,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active
python tools/profile_step.py --size 256 --warmup 6 > gpurun_out/${TAG}_profile_step_plain.log 2>&1 && \
MF_PROFILE_LEVELS=1 ncu --profile-from-start off --cache-control none --clock-control none --metrics $M --csv --log-file gpurun_out/${TAG}_ncu_step_256_mg.csv \
  python tools/profile_step.py --size 256 --warmup 6 > gpurun_out/${TAG}_ncu_step.log 2>&1; echo "ncu step rc=$?"
# the driver's recipe: launch list of the bench command itself (gpu__time_duration.sum only)
ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/${TAG}_ncu_launches_bench.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-strong1024 --profile-steps 0 > gpurun_out/${TAG}_ncu_launches.log 2>&1; echo "ncu launches rc=$?"
ncu --set full --clock-control none --cache-control none --import-source on --profile-from-start off -k regex:'k_mg_smooth|k_mc_mac_f|k_sl_mac_f|k_mg_down0|k_mg_up0w|k_cg_spmv_dot_code' -c 24 \
  -o gpurun_out/${TAG}_full -f python tools/profile_step.py --size 256 --warmup 6 > gpurun_out/${TAG}_ncu_full.log 2>&1; echo "ncu full rc=$?"
MF_ADVECT_TMA=1 ncu --set full --clock-control none --cache-control none --import-source on --profile-from-start off -k regex:'k_sl_mac_tma' -c 1 \
  -o gpurun_out/${TAG}_full_tma -f python tools/profile_step.py --size 256 --warmup 6 > gpurun_out/${TAG}_ncu_full_tma.log 2>&1; echo "ncu tma rc=$?"
ncu --set full --clock-control none --cache-control none --import-source on --profile-from-start off -k regex:'k_mic_code' -c 2 \
  -o gpurun_out/${TAG}_full_mic -f python tools/profile_step.py --size 256 --warmup 4 --precond mic > gpurun_out/${TAG}_ncu_full_mic.log 2>&1; echo "ncu mic rc=$?"
tail -c 400 gpurun_out/${TAG}_bench.err
python tools/show_bench.py gpurun_out/${TAG}_bench_256_mg.json
