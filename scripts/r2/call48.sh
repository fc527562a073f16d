#!/bin/bash
# round 2, call 48: two rows per thread in the half sweep (k_gs_half_c2): full suite, A/B by WL_NO_GS_C2
mkdir -p gpurun_out; cd /root/repo
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2_c48_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2_c48_pytest.log | cut -c1-400
for v in off on off on; do
  if [ $v = on ]; then unset WL_NO_GS_C2; else export WL_NO_GS_C2=1; fi
  timeout 300 python bench.py --no-cpu --steps 20 --breakdown gpurun_out/r2_c48_breakdown_$v.tsv > gpurun_out/r2_c48_bench_$v.json 2> gpurun_out/r2_c48_$v.err; echo "c2 $v rc=$?"
  python - <<PY
import json
d=json.loads(open("gpurun_out/r2_c48_bench_$v.json").read().strip().splitlines()[-1]); print("c2 $v", d["value"], d["ms_per_step"])
PY
  grep "k_gs_half_c" gpurun_out/r2_c48_breakdown_$v.tsv | cut -f1,4 | tr '\n' ' '; echo
done
unset WL_NO_GS_C2
for w in tgv cylinder; do timeout 300 python bench.py --no-cpu --workload $w --steps 20 > gpurun_out/r2_c48_bench_$w.json 2>/dev/null; python -c "
import json; d=json.loads(open('gpurun_out/r2_c48_bench_$w.json').read().strip().splitlines()[-1]); print('$w', d['value'], d['ms_per_step'])"; done
