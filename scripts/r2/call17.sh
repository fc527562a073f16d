#!/bin/bash
# round 2, call 17: same-box A/B of library builds: base (before the hoisted loads), cur (hoisted + runtime accel-field branch in K-MOM),
# tpl (accel field as a template parameter), unr (tpl + four-plane unrolled march)
mkdir -p gpurun_out; cd /root/repo
for v in base cur tpl unr base unr; do
  if [ $v = cur ]; then unset WL_B200_LIB; else export WL_B200_LIB=/root/repo/gpurun_tmp/lib_$v.so; fi
  timeout 300 python bench.py --no-cpu --steps 20 --breakdown gpurun_out/r2_c17_breakdown_sphere_$v.tsv > gpurun_out/r2_c17_bench_sphere_$v.json 2> gpurun_out/r2_c17_bench_$v.err; echo "bench $v rc=$?"
  python - <<PY
import json
d=json.loads(open("gpurun_out/r2_c17_bench_sphere_$v.json").read().strip().splitlines()[-1]); print("$v", d["value"], d["ms_per_step"], d["roofline"]["frac"], d["roofline"]["avg_ms"])
PY
  grep "@32768\|@65536\|@2368" gpurun_out/r2_c17_breakdown_sphere_$v.tsv | cut -f1,4 | tr '\n' ' '; echo
done
export WL_B200_LIB=/root/repo/gpurun_tmp/lib_unr.so
timeout 300 python bench.py --no-cpu --workload tgv --steps 50 > gpurun_out/r2_c17_bench_tgv_unr.json 2>/dev/null; python -c "
import json; d=json.loads(open('gpurun_out/r2_c17_bench_tgv_unr.json').read().strip().splitlines()[-1]); print('tgv unr', d['value'], d['ms_per_step'])"
timeout 900 python -m pytest tests/test_parity_gpu.py tests/test_fullsize_gpu.py tests/test_fnbc_gpu.py -x -q > gpurun_out/r2_c17_pytest_unr.log 2>&1; echo "pytest unr rc=$?"; tail -3 gpurun_out/r2_c17_pytest_unr.log
