#!/bin/bash
# round 2, call 14: ncu of the red-black half sweeps (pattern tiles on), 16 launches = levels 2..4 and level 1 of one solver iteration
mkdir -p gpurun_out; cd /root/repo
export WL_GS_CPT=1
python scripts/r2/one_step.py sphere 1 > gpurun_out/r2_c14_plain.log 2>&1 &&
ncu --section SpeedOfLight --section MemoryWorkloadAnalysis --section MemoryWorkloadAnalysis_Tables --section WarpStateStats --section LaunchStats --section Occupancy --section SchedulerStats --clock-control none -k regex:k_gs_half_c -c 16 -f -o gpurun_out/r2_c14_gs python scripts/r2/one_step.py sphere 1 > gpurun_out/r2_c14_ncu.log 2>&1
echo "rc=$?"; tail -3 gpurun_out/r2_c14_ncu.log
