#!/bin/bash
# ncu launch lists at the round's HEAD: one step with DRAM bytes / hit rates / issue utilisation, and the driver's recipe over the bench command
TAG=${1:-r2d}
mkdir -p gpurun_out
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,l1tex__t_sector_hit_rate.pct,sm__warps_active.avg.pct_of_peak_sustained_active,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active
python tools/profile_step.py --size 256 --warmup 6 > gpurun_out/${TAG}_profile_step_plain.log 2>&1 && \
MF_PROFILE_LEVELS=1 timeout 200 ncu --profile-from-start off --cache-control none --clock-control none --metrics $M --csv --log-file gpurun_out/${TAG}_ncu_step_256_mg.csv \
  python tools/profile_step.py --size 256 --warmup 6 > gpurun_out/${TAG}_ncu_step.log 2>&1; echo "ncu step rc=$?"
timeout 120 ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/${TAG}_ncu_launches_bench.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-strong1024 --profile-steps 0 > gpurun_out/${TAG}_ncu_launches.log 2>&1; echo "ncu launches rc=$?"
