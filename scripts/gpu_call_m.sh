#!/bin/bash
mkdir -p gpurun_out; cd /root/repo
( timeout 120 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "meanflow or viscous or golden" ) > gpurun_out/pytest_gpu_mean.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu_mean.log
