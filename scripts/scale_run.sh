#!/bin/bash
# usage: scripts/scale_run.sh "<Ns>" <workload> [extra bench args]   -> gpurun_out/scale_<workload>_N.json
Ns="$1"; W="$2"; shift 2
for N in $Ns; do
  if [ "$N" = "1" ]; then
    timeout 300 python bench.py --gpus 1 --workload $W --no-cpu --breakdown gpurun_out/bd_${W}_$N.tsv "$@" > gpurun_out/scale_${W}_$N.json 2> gpurun_out/scale_${W}_$N.err
  else
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29700+N)) \
      bench.py --gpus $N --workload $W --no-cpu --breakdown gpurun_out/bd_${W}_$N.tsv "$@" > gpurun_out/scale_${W}_$N.json 2> gpurun_out/scale_${W}_$N.err
  fi
  echo "N=$N rc=$?"; python - <<PY
import json
try:
    l=[x for x in open("gpurun_out/scale_${W}_$N.json") if x.startswith("{")][-1]
    d=json.loads(l); print($N, d["value"], d["ms_per_step"], d["config"]["parallelism"], d["e2e"]["value"])
except Exception as e: print("no result", e)
PY
done
