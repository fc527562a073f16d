#!/bin/bash
# round 2, call 40: prefetch distance sweep (K-PROJ without prefetch)
mkdir -p gpurun_out; cd /root/repo
for pf in 4 6 8 def 12 6 def; do
  if [ $pf = def ]; then unset WL_PREFETCH; else export WL_PREFETCH=$pf; fi
  timeout 300 python bench.py --no-cpu --steps 20 --breakdown gpurun_out/r2_c40_breakdown_pf$pf.tsv > gpurun_out/r2_c40_bench_pf$pf.json 2> gpurun_out/r2_c40_pf$pf.err; echo "pf=$pf rc=$?"
  python - <<PY
import json
d=json.loads(open("gpurun_out/r2_c40_bench_pf$pf.json").read().strip().splitlines()[-1]); print("pf=$pf", d["value"], d["ms_per_step"])
PY
  grep "@32768\|@65536" gpurun_out/r2_c40_breakdown_pf$pf.tsv | cut -f1,4 | tr '\n' ' '; echo
done
unset WL_PREFETCH
