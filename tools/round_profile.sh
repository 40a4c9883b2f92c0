#!/bin/bash
# end-of-round measurements on ONE B200: bench line, reference arm, ncu launch list of one step, ncu --set full of the top kernels.
# usage: tools/round_profile.sh <tag>     (writes gpurun_out/<tag>_*)
TAG=${1:-rX}
mkdir -p gpurun_out
timeout 900 python bench.py --steps 10 --warmup 5 --kernels-out gpurun_out/${TAG}_kernels_256_mg.json > gpurun_out/${TAG}_bench_256_mg.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${TAG}_bench_reference_arm.json 2>> gpurun_out/${TAG}_bench.err; echo "reference arm rc=$?"
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,l1tex__t_sector_hit_rate.pct,sm__warps_active.avg.pct_of_peak_sustained_active,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active
python tools/profile_step.py --size 256 --warmup 6 > gpurun_out/${TAG}_profile_step_plain.log 2>&1 && \
ncu --profile-from-start off --cache-control none --clock-control none --metrics $M --csv --log-file gpurun_out/${TAG}_ncu_step_256_mg.csv \
  python tools/profile_step.py --size 256 --warmup 6 > gpurun_out/${TAG}_ncu_step.log 2>&1; echo "ncu step rc=$?"
ncu --set full --clock-control none --cache-control none --import-source on --profile-from-start off -k regex:'k_mc_mac_f|k_mg_down0|k_mg_up0w|k_cg_spmv_dot_code|k_sl_f' -c 6 \
  -o gpurun_out/${TAG}_full -f python tools/profile_step.py --size 256 --warmup 6 > gpurun_out/${TAG}_ncu_full.log 2>&1; echo "ncu full rc=$?"
timeout 600 python bench.py --steps 3 --warmup 3 --size 1024 --no-e2e --no-cpu-baseline > gpurun_out/${TAG}_bench_strong1024_1gpu.json 2>> gpurun_out/${TAG}_bench.err; echo "1024 rc=$?"
tail -c 400 gpurun_out/${TAG}_bench.err
