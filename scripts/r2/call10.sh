#!/bin/bash
# round 2, call 10: wl_cfl_noalloc test alone in a fresh process, then the whole GPU suite (it runs last), then TGV bench
mkdir -p gpurun_out; cd /root/repo
( timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q ) > gpurun_out/r2_c10_noalloc_alone.log 2>&1; echo "noalloc alone rc=$?"; grep -v "^E  " gpurun_out/r2_c10_noalloc_alone.log | tail -12 | cut -c1-250
( timeout 900 python -m pytest tests -m gpu -q -x ) > gpurun_out/r2_c10_pytest_gpu.log 2>&1; echo "pytest rc=$?"; grep -v "^E  " gpurun_out/r2_c10_pytest_gpu.log | tail -6 | cut -c1-250
timeout 300 python bench.py --workload tgv --no-cpu --steps 20 --warmup 5 --breakdown gpurun_out/r2_c10_breakdown_tgv128.tsv 2>gpurun_out/r2_c10_bench.err | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('tgv', d['value'], d['ms_per_step'], d['gpu_launches'], d['roofline']['kernel'], d['roofline']['timing'])"
tail -3 gpurun_out/r2_c10_bench.err
