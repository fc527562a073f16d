#!/bin/bash
# round 2, call 56: ncu launch list of the round's final build (default workload, short: the GPU budget is nearly spent)
mkdir -p gpurun_out; cd /root/repo
timeout 65 ncu --metrics gpu__time_duration.sum --clock-control none -c 1000 --csv --log-file gpurun_out/r2_c56_ncu_launches_default.csv python bench.py --no-cpu --steps 2 --warmup 3 > gpurun_out/r2_c56_ncu_bench.log 2>&1; echo "ncu rc=$?"; wc -l gpurun_out/r2_c56_ncu_launches_default.csv
