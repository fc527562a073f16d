#!/bin/bash
# round 2, call 58: smoke() at the head of the round
mkdir -p gpurun_out; cd /root/repo
timeout 25 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r2_c58_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/r2_c58_smoke.log | cut -c1-300
