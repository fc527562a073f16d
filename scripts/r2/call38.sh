#!/bin/bash
# round 2, call 38: L2 prefetch of the lines a later block layer will read, in the half sweep: distance sweep (planes ahead)
mkdir -p gpurun_out; cd /root/repo
for pf in 0 4 8 16 0 8; do
  export WL_GS_PREFETCH=$pf
  timeout 300 python bench.py --no-cpu --steps 20 --breakdown gpurun_out/r2_c38_breakdown_pf$pf.tsv > gpurun_out/r2_c38_bench_pf$pf.json 2> gpurun_out/r2_c38_pf$pf.err; echo "pf=$pf rc=$?"
  python - <<PY
import json
d=json.loads(open("gpurun_out/r2_c38_bench_pf$pf.json").read().strip().splitlines()[-1]); print("pf=$pf", d["value"], d["ms_per_step"])
PY
  grep "k_gs_half_c<3>@" gpurun_out/r2_c38_breakdown_pf$pf.tsv | cut -f1,4 | tr '\n' ' '; echo
done
