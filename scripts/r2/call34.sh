#!/bin/bash
# round 2, call 34: register caps -- k_proj4 at 64 (4 blocks/SM); b5 = additionally k_incr4 / k_up4 at 48 (5 blocks/SM, 72 / 32 B spilled); A/B on one box
mkdir -p gpurun_out; cd /root/repo
for v in prev cur b5 prev cur b5; do
  if [ $v = cur ]; then unset WL_B200_LIB; else export WL_B200_LIB=/root/repo/gpurun_tmp/lib_$v.so; fi
  timeout 300 python bench.py --no-cpu --steps 20 --breakdown gpurun_out/r2_c34_breakdown_$v.tsv > gpurun_out/r2_c34_bench_$v.json 2> gpurun_out/r2_c34_$v.err; echo "bench $v rc=$?"
  python - <<PY
import json
d=json.loads(open("gpurun_out/r2_c34_bench_$v.json").read().strip().splitlines()[-1]); print("$v", d["value"], d["ms_per_step"])
PY
  grep "@32768\|@65536" gpurun_out/r2_c34_breakdown_$v.tsv | cut -f1,4 | tr '\n' ' '; echo
done
unset WL_B200_LIB
timeout 300 ncu --set full --import-source on --clock-control none -k regex:k_gs_half_c -s 59 -c 1 -f -o gpurun_out/r2_c34_gs python scripts/r2/one_step.py sphere 2 > gpurun_out/r2_c34_ncu_gs.log 2>&1; echo "ncu gs rc=$?"
