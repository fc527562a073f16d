#!/bin/bash
mkdir -p gpurun_out; cd /root/repo
( timeout 20 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "csg or viscous_and_total" ) > gpurun_out/pytest_gpu_csg.log 2>&1; echo "pytest rc=$?"; tail -25 gpurun_out/pytest_gpu_csg.log
