#!/bin/bash
# round 2, call 45: after removing the dormant in-kernel-finish switch from the 4-cells-per-thread kernels: full suite + default bench
mkdir -p gpurun_out; cd /root/repo
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2_c45_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_c45_pytest.log | cut -c1-300
timeout 300 python bench.py --no-cpu --steps 20 > gpurun_out/r2_c45_bench.json 2> gpurun_out/r2_c45.err; echo "bench rc=$?"
python -c "
import json; d=json.loads(open('gpurun_out/r2_c45_bench.json').read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], d['gpu_launches'])"
