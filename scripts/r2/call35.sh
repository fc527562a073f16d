#!/bin/bash
# round 2, call 35: cur = k_proj4 at 4 blocks + k_up4 at 5; all5 = k_proj4, k_up4, k_down4<1>, k_divres4<1> at 5 blocks/SM; A/B on one box
mkdir -p gpurun_out; cd /root/repo
for v in prev cur all5 prev cur all5; do
  if [ $v = cur ]; then unset WL_B200_LIB; else export WL_B200_LIB=/root/repo/gpurun_tmp/lib_$v.so; fi
  timeout 300 python bench.py --no-cpu --steps 20 --breakdown gpurun_out/r2_c35_breakdown_$v.tsv > gpurun_out/r2_c35_bench_$v.json 2> gpurun_out/r2_c35_$v.err; echo "bench $v rc=$?"
  python - <<PY
import json
d=json.loads(open("gpurun_out/r2_c35_bench_$v.json").read().strip().splitlines()[-1]); print("$v", d["value"], d["ms_per_step"])
PY
  grep "@32768\|@65536" gpurun_out/r2_c35_breakdown_$v.tsv | cut -f1,4 | tr '\n' ' '; echo
done
unset WL_B200_LIB
