#!/bin/bash
# round 2, call 32 (2 GPUs): the 2-rank worker with the two-launch form forced on every level; ncu --set full of one k_bc_vec launch
mkdir -p gpurun_out; cd /root/repo
timeout 900 python -m pytest tests/test_multigpu.py -m gpu -x -q > gpurun_out/r2_c32_pytest_mgpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2_c32_pytest_mgpu.log | cut -c1-300
timeout 300 ncu --set full --import-source on --clock-control none -k regex:k_bc_vec -s 6 -c 1 -f -o gpurun_out/r2_c32_bcvec python scripts/r2/one_step.py sphere 2 > gpurun_out/r2_c32_ncu.log 2>&1; echo "ncu rc=$?"
