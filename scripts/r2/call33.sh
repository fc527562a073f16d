#!/bin/bash
# round 2, call 33: ncu --set full + source of one level-1 half sweep, one level-1 k_incr4 and one k_up4 (current state)
mkdir -p gpurun_out; cd /root/repo
timeout 300 ncu --set full --import-source on --clock-control none -k regex:k_gs_half_c -s 40 -c 1 -f -o gpurun_out/r2_c33_gs python scripts/r2/one_step.py sphere 2 > gpurun_out/r2_c33_ncu_gs.log 2>&1; echo "ncu gs rc=$?"
timeout 300 ncu --set full --import-source on --clock-control none -k regex:k_incr4 -s 14 -c 1 -f -o gpurun_out/r2_c33_incr python scripts/r2/one_step.py sphere 2 > gpurun_out/r2_c33_ncu_incr.log 2>&1; echo "ncu incr rc=$?"
timeout 300 ncu --set full --import-source on --clock-control none -k regex:k_proj4 -s 2 -c 1 -f -o gpurun_out/r2_c33_proj python scripts/r2/one_step.py sphere 2 > gpurun_out/r2_c33_ncu_proj.log 2>&1; echo "ncu proj rc=$?"
