#!/bin/bash
# round 2, call 13: 8-cells-per-thread sweeps (k_gs8, class-aware): parity suite, bench A/B
mkdir -p gpurun_out; cd /root/repo
( timeout 900 python -m pytest tests -m gpu -q -x ) > gpurun_out/r2_c13_pytest_gpu.log 2>&1; echo "pytest rc=$?"; grep -v "^E  " gpurun_out/r2_c13_pytest_gpu.log | tail -5 | cut -c1-250
for v in gs8 nogs8; do
  if [ $v = nogs8 ]; then export WL_NO_GS8=1; else unset WL_NO_GS8; fi
  timeout 300 python bench.py --no-cpu --steps 10 --warmup 3 --breakdown gpurun_out/r2_c13_breakdown_sphere512_$v.tsv 2>gpurun_out/r2_c13_bench.err | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('sphere $v', d['value'], d['ms_per_step'])"
  grep "k_gs" gpurun_out/r2_c13_breakdown_sphere512_$v.tsv
  timeout 300 python bench.py --workload tgv --no-cpu --steps 20 --warmup 5 2>>gpurun_out/r2_c13_bench.err | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('tgv $v', d['value'], d['ms_per_step'])"
done
tail -3 gpurun_out/r2_c13_bench.err
