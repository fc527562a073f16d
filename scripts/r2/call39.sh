#!/bin/bash
# round 2, call 39: L2 prefetch in the half sweep and the five 4-cells-per-thread kernels: off / default distance / 16 planes; full suite
mkdir -p gpurun_out; cd /root/repo
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2_c39_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_c39_pytest.log | cut -c1-300
for pf in 0 def 16 0 def; do
  if [ $pf = def ]; then unset WL_PREFETCH; else export WL_PREFETCH=$pf; fi
  timeout 300 python bench.py --no-cpu --steps 20 --breakdown gpurun_out/r2_c39_breakdown_pf$pf.tsv > gpurun_out/r2_c39_bench_pf$pf.json 2> gpurun_out/r2_c39_pf$pf.err; echo "pf=$pf rc=$?"
  python - <<PY
import json
d=json.loads(open("gpurun_out/r2_c39_bench_pf$pf.json").read().strip().splitlines()[-1]); print("pf=$pf", d["value"], d["ms_per_step"])
PY
  grep "@32768\|@65536" gpurun_out/r2_c39_breakdown_pf$pf.tsv | cut -f1,4 | tr '\n' ' '; echo
done
unset WL_PREFETCH
for w in tgv cylinder; do timeout 300 python bench.py --no-cpu --workload $w --steps 20 > gpurun_out/r2_c39_bench_$w.json 2>/dev/null; python -c "
import json; d=json.loads(open('gpurun_out/r2_c39_bench_$w.json').read().strip().splitlines()[-1]); print('$w', d['value'], d['ms_per_step'])"; done
