#!/bin/bash
# round 2, call 36: cur = k_proj4 / k_up4 / k_down4<1> at 5 blocks per SM; d6 = k_down4<1> at 6 (40 registers, no spill); full suite on cur
mkdir -p gpurun_out; cd /root/repo
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2_c36_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_c36_pytest.log | cut -c1-300
for v in prev cur d6 prev cur d6; do
  if [ $v = cur ]; then unset WL_B200_LIB; else export WL_B200_LIB=/root/repo/gpurun_tmp/lib_$v.so; fi
  timeout 300 python bench.py --no-cpu --steps 20 --breakdown gpurun_out/r2_c36_breakdown_$v.tsv > gpurun_out/r2_c36_bench_$v.json 2> gpurun_out/r2_c36_$v.err; echo "bench $v rc=$?"
  python - <<PY
import json
d=json.loads(open("gpurun_out/r2_c36_bench_$v.json").read().strip().splitlines()[-1]); print("$v", d["value"], d["ms_per_step"])
PY
  grep "@32768\|@65536" gpurun_out/r2_c36_breakdown_$v.tsv | cut -f1,4 | tr '\n' ' '; echo
done
unset WL_B200_LIB
for w in tgv cylinder; do timeout 300 python bench.py --no-cpu --workload $w --steps 20 > gpurun_out/r2_c36_bench_$w.json 2>/dev/null; python -c "
import json; d=json.loads(open('gpurun_out/r2_c36_bench_$w.json').read().strip().splitlines()[-1]); print('$w', d['value'], d['ms_per_step'])"; done
