#!/bin/bash
mkdir -p gpurun_out; cd /root/repo
( timeout 60 python -m pytest tests/test_zz_io_gpu.py -m gpu -x -q ) > gpurun_out/pytest_gpu_io.log 2>&1; echo "pytest rc=$?"; tail -25 gpurun_out/pytest_gpu_io.log
