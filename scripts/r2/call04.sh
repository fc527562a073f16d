#!/bin/bash
# round 2, call 4: ncu --set full of the level-1 launches of the paired sweeps (sphere 512x256x256)
mkdir -p gpurun_out; cd /root/repo
python scripts/r2/one_step.py sphere 1 > gpurun_out/r2_c04_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_gs_pair -s 8 -c 2 -f -o gpurun_out/r2_c04_gs_pair python scripts/r2/one_step.py sphere 1 > gpurun_out/r2_c04_ncu.log 2>&1
echo "rc=$?"; tail -5 gpurun_out/r2_c04_ncu.log
