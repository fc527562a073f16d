#!/bin/bash
# round 2, call 47: ncu --set full of the final level-1 kernels: k_down4<1>, k_divres4<1>, k_up4, k_proj4, k_gs_half_c (4th sweep of a V-cycle)
mkdir -p gpurun_out; cd /root/repo
cap() { timeout 300 ncu --set full --clock-control none -k "regex:$1" -s $2 -c 1 -f -o gpurun_out/r2_c47_$3 python scripts/r2/one_step.py sphere 2 > gpurun_out/r2_c47_ncu_$3.log 2>&1; echo "ncu $3 rc=$?"; }
cap "k_down4<1>" 2 down4_lean
cap "k_divres4<1>" 2 divres4_lean
cap "k_up4" 14 up4
cap "k_proj4" 2 proj4
cap "k_gs_half_c" 59 gs
ls -la gpurun_out/r2_c47_*.ncu-rep
