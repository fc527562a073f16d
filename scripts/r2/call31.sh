#!/bin/bash
# round 2, call 31 (overlap also while only k_mom3 is profiled): list launch of the split kernels on a higher-priority side stream (concurrent with the main launch); A/B against call 29's build
mkdir -p gpurun_out; cd /root/repo
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2_c31_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2_c31_pytest.log | cut -c1-300
for v in prev cur prev cur; do
  if [ $v = cur ]; then unset WL_B200_LIB; else export WL_B200_LIB=/root/repo/gpurun_tmp/lib_$v.so; fi
  timeout 300 python bench.py --no-cpu --steps 20 --breakdown gpurun_out/r2_c31_breakdown_sphere_$v.tsv > gpurun_out/r2_c31_bench_sphere_$v.json 2> gpurun_out/r2_c31_bench_$v.err; echo "bench $v rc=$?"
  python - <<PY
import json
d=json.loads(open("gpurun_out/r2_c31_bench_sphere_$v.json").read().strip().splitlines()[-1]); print("$v", d["value"], d["ms_per_step"], d["roofline"]["frac"], d["config"]["pattern_tiles_level1"]["general"], d["config"]["pattern_tiles_level1"]["constants"], d["config"]["pattern_tiles_level1"]["constants_with_boundary"])
PY
  grep "@32768\|@65536\|@2368" gpurun_out/r2_c31_breakdown_sphere_$v.tsv | cut -f1,4 | tr '\n' ' '; echo
done
unset WL_B200_LIB
for w in tgv cylinder; do timeout 300 python bench.py --no-cpu --workload $w --steps 20 > gpurun_out/r2_c31_bench_$w.json 2>/dev/null; python -c "
import json; d=json.loads(open('gpurun_out/r2_c31_bench_$w.json').read().strip().splitlines()[-1]); print('$w', d['value'], d['ms_per_step'])"; done
