#!/bin/bash
mkdir -p gpurun_out; cd /root/repo
( timeout 300 python -m pytest tests -m gpu -x -q ) > gpurun_out/pytest_gpu_final3.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/pytest_gpu_final3.log
timeout 100 python __graft_entry__.py smoke 2>&1 | tail -1
