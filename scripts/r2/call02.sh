#!/bin/bash
# round 2, call 2: parity at the timed sizes + closure-body route + time-dependent Function uBC
mkdir -p gpurun_out; cd /root/repo
( timeout 1500 python -m pytest tests/test_fullsize_gpu.py -m gpu -q -s ) > gpurun_out/r2_c02_fullsize.log 2>&1; echo "fullsize rc=$?"; tail -40 gpurun_out/r2_c02_fullsize.log
