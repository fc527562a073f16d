#!/bin/bash
# usage: scripts/r2/gpu.sh <call script> <timeout s> [gpus]  -- runs it on the GPU box, retrying while the pod is busy
call=$1; to=${2:-900}; g=${3:-1}
out=gpurun_out/$(basename $call .sh).out
for try in $(seq 1 40); do
  if [ "$g" = "1" ]; then /usr/local/graft/bin/gpurun --timeout $to -- "bash $call" > $out 2>&1; rc=$?
  else /usr/local/graft/bin/gpurun --gpus $g --timeout $to -- "bash $call" > $out 2>&1; rc=$?; fi
  if grep -q "status=transient\|rc=3\b" $out && ! grep -q "status=ok" $out; then sleep 90; continue; fi
  if [ $rc -eq 3 ]; then sleep 90; continue; fi
  break
done
echo "gpurun rc=$rc tries=$try" >> $out
