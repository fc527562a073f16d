#!/bin/bash
# round 2, call 8: first GPU run of the TMA K-MOM (commit a59530b) and of the pruned library: whole GPU suite, then
# sphere / TGV bench with TMA on and off
mkdir -p gpurun_out; cd /root/repo
( timeout 900 python -m pytest tests -m gpu -q -x ) > gpurun_out/r2_c08_pytest_gpu.log 2>&1; echo "pytest rc=$?"; grep -v "^E  " gpurun_out/r2_c08_pytest_gpu.log | tail -12 | cut -c1-300
for v in tma notma; do
  if [ $v = notma ]; then export WL_NO_TMA=1; else unset WL_NO_TMA; fi
  timeout 300 python bench.py --no-cpu --steps 10 --warmup 3 --breakdown gpurun_out/r2_c08_breakdown_sphere512_$v.tsv 2>gpurun_out/r2_c08_bench_$v.err | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('sphere $v', d['value'], d['ms_per_step'], 'kmom avg ms', d['roofline']['avg_ms'], d['roofline']['frac'])"
  timeout 300 python bench.py --workload tgv --no-cpu --steps 20 --warmup 5 2>>gpurun_out/r2_c08_bench_$v.err | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('tgv $v', d['value'], d['ms_per_step'], d['roofline']['kernel'], d['roofline']['avg_ms'])"
done
tail -3 gpurun_out/r2_c08_bench_tma.err
head -6 gpurun_out/r2_c08_breakdown_sphere512_tma.tsv
