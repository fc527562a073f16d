#!/bin/bash
# round 2, call 12: uniform-coefficient tile flags: whole GPU suite (bit-exactness vs oracle), sphere / TGV / cylinder bench on and off
mkdir -p gpurun_out; cd /root/repo
( timeout 900 python -m pytest tests -m gpu -q -x ) > gpurun_out/r2_c12_pytest_gpu.log 2>&1; echo "pytest rc=$?"; grep -v "^E  " gpurun_out/r2_c12_pytest_gpu.log | tail -8 | cut -c1-250
for v in flags noflags; do
  if [ $v = noflags ]; then export WL_NO_TILE_FLAGS=1; else unset WL_NO_TILE_FLAGS; fi
  timeout 300 python bench.py --no-cpu --steps 10 --warmup 3 --breakdown gpurun_out/r2_c12_breakdown_sphere512_$v.tsv 2>gpurun_out/r2_c12_bench_$v.err | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('sphere $v', d['value'], d['ms_per_step'])"
  timeout 300 python bench.py --workload tgv --no-cpu --steps 20 --warmup 5 2>>gpurun_out/r2_c12_bench_$v.err | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('tgv $v', d['value'], d['ms_per_step'])"
done
unset WL_NO_TILE_FLAGS
timeout 300 python bench.py --workload cylinder --no-cpu --steps 5 --warmup 3 --breakdown gpurun_out/r2_c12_breakdown_cylinder.tsv 2>>gpurun_out/r2_c12_bench_flags.err | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('cylinder flags', d['value'], d['ms_per_step'])"
tail -3 gpurun_out/r2_c12_bench_flags.err
cat gpurun_out/r2_c12_breakdown_sphere512_flags.tsv
