#!/bin/bash
# round 2, call 43: K-GS8 (eight cells per thread on the class-1/3 tiles + one-cell kernel over the class-0/2 list): full suite, A/B by WL_NO_GS8
mkdir -p gpurun_out; cd /root/repo
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2_c43_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2_c43_pytest.log | cut -c1-400
for v in off on off on; do
  if [ $v = on ]; then unset WL_NO_GS8; else export WL_NO_GS8=1; fi
  timeout 300 python bench.py --no-cpu --steps 20 --breakdown gpurun_out/r2_c43_breakdown_$v.tsv > gpurun_out/r2_c43_bench_$v.json 2> gpurun_out/r2_c43_$v.err; echo "gs8 $v rc=$?"
  python - <<PY
import json
d=json.loads(open("gpurun_out/r2_c43_bench_$v.json").read().strip().splitlines()[-1]); print("gs8 $v", d["value"], d["ms_per_step"])
PY
  grep "k_gs8@\|k_gs_half_c<3>@65536\|k_gs_half_c<3, true>" gpurun_out/r2_c43_breakdown_$v.tsv | cut -f1,2,4 | tr '\n' ' '; echo
done
unset WL_NO_GS8
for w in tgv cylinder; do timeout 300 python bench.py --no-cpu --workload $w --steps 20 > gpurun_out/r2_c43_bench_$w.json 2>/dev/null; python -c "
import json; d=json.loads(open('gpurun_out/r2_c43_bench_$w.json').read().strip().splitlines()[-1]); print('$w', d['value'], d['ms_per_step'])"; done
