#!/bin/bash
# round 2, call 49: half sweep with two rows per thread (default) against four cells per thread (WL_GS_C4=1): parity of c4 at the timed sizes, A/B
mkdir -p gpurun_out; cd /root/repo
WL_GS_C4=1 timeout 900 python -m pytest tests/test_fullsize_gpu.py -m gpu -x -q -k "config_B or config_C" > gpurun_out/r2_c49_pytest_c4.log 2>&1; echo "pytest c4 rc=$?"; tail -2 gpurun_out/r2_c49_pytest_c4.log | cut -c1-300
for v in c2 c4 c2 c4; do
  if [ $v = c4 ]; then export WL_GS_C4=1; else unset WL_GS_C4; fi
  timeout 300 python bench.py --no-cpu --steps 20 --breakdown gpurun_out/r2_c49_breakdown_$v.tsv > gpurun_out/r2_c49_bench_$v.json 2> gpurun_out/r2_c49_$v.err; echo "$v rc=$?"
  python - <<PY
import json
d=json.loads(open("gpurun_out/r2_c49_bench_$v.json").read().strip().splitlines()[-1]); print("$v", d["value"], d["ms_per_step"])
PY
  grep "k_gs_half_c" gpurun_out/r2_c49_breakdown_$v.tsv | cut -f1,4 | tr '\n' ' '; echo
done
unset WL_GS_C4
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2_c49_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/r2_c49_pytest.log | cut -c1-300
for w in tgv cylinder; do timeout 300 python bench.py --no-cpu --workload $w --steps 20 > gpurun_out/r2_c49_bench_$w.json 2>/dev/null; python -c "
import json; d=json.loads(open('gpurun_out/r2_c49_bench_$w.json').read().strip().splitlines()[-1]); print('$w', d['value'], d['ms_per_step'])"; done
