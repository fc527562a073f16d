#!/bin/bash
# round 2, call 5: L2-blocked smoother: parity (default and forced thin slabs on small grids) + bench sweep over slab sizes
mkdir -p gpurun_out; cd /root/repo
( timeout 1500 python -m pytest tests -m gpu -q -x ) > gpurun_out/r2_c05_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2_c05_pytest_gpu.log
( WL_L2_BLOCK_MIN_CELLS=1 WL_L2_SLAB_PLANES=4 timeout 1500 python -m pytest tests/test_parity_gpu.py tests/test_fullsize_gpu.py tests/test_golden.py -m gpu -q -x ) > gpurun_out/r2_c05_pytest_gpu_thin_slabs.log 2>&1; echo "pytest thin rc=$?"; tail -4 gpurun_out/r2_c05_pytest_gpu_thin_slabs.log
for mb in 24 48 72 96; do
WL_L2_SLAB_MB=$mb timeout 600 python bench.py --no-cpu --steps 10 --warmup 3 --breakdown gpurun_out/r2_c05_breakdown_sphere512_slab$mb.tsv 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('sphere slab_mb=$mb', d['value'], d['ms_per_step'], d['gpu_launches'])"
done
WL_NO_L2_BLOCK=1 timeout 600 python bench.py --no-cpu --steps 10 --warmup 3 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('sphere unblocked', d['value'], d['ms_per_step'], d['gpu_launches'])"
head -12 gpurun_out/r2_c05_breakdown_sphere512_slab48.tsv
